/*
 * trt_oracle.c -- CPU ORACLE (test infrastructure; pinned bit for bit to oracle/_ref, see trt_oracle.h).
 * Build: oracle/Makefile  (gcc -O2 -ffp-contract=off -fno-fast-math -mfma -fopenmp).
 * -mfma only makes the *explicit* fmaf() calls (pinned sin/cos, conv 1 helpers) one instruction;
 * -ffp-contract=off guarantees the compiler never fuses anything on its own.
 */
#include "trt_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

typedef struct { float x, y, z; } v3;
typedef struct {
    v3 p, normal;
    float t;
    int front_face;
    v3 color, emission;
    float reflectivity;
    uint32_t index;
} hitrec;                                   /* shader.wgsl:39-47 */

typedef struct {
    const oprim *prims;
    const ouniforms *u;
    float tan_half_fov;
    int math_mode;
} ctx_t;

/* f32 values of the WGSL literals: 3.14159265359, 2.0*3.14159265359 and 4.0*3.14159265359 are
 * abstract-float constant expressions, rounded to f32 once (shader.wgsl:318,358,369,412). */
#define PI_F 3.14159265359f
#define TWO_PI_F ((float)(2.0 * 3.14159265359))
#define FOUR_PI_F ((float)(4.0 * 3.14159265359))

static inline v3 ld3(const ovec3 *p) { v3 r = { p->x, p->y, p->z }; return r; }

/* shader.wgsl:106-110 */
static inline uint32_t pcg_hash(uint32_t input) {
    uint32_t state = input * 747796405u + 2891336453u;
    uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
    return (word >> 22u) ^ word;
}
/* shader.wgsl:112-115: the literal 4294967295.0 rounds to 2^32 in f32; f32(u32) rounds to nearest */
static inline float rnd(uint32_t *state) {
    *state = pcg_hash(*state);
    return (float)(*state) / 4294967296.0f;
}
/* WGSL u32(f32): clamp to [0, 2^32-1] then truncate (shader.wgsl:562) */
static inline uint32_t f32_to_u32_sat(float v) {
    if (!(v > 0.0f)) return 0u;
    if (v >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)v;
}

#include "trt_sincos_pinned.h"   /* the pinned sin/cos of conv 0 / 1 */
static inline void sincos_mode(int math_mode, float phi, float *s, float *c) {
    if (math_mode == 1) { *s = sinf(phi); *c = cosf(phi); }
    else sincos_pinned(phi, s, c);
}

#define CAT2(a, b) a##b
#define CAT(a, b) CAT2(a, b)

#define CONV 0
#define COUNT 0
#define FN(n) CAT(n, _c0)
#include "trt_oracle_core.inc"
#undef FN
#undef COUNT
#define COUNT 1
#define FN(n) CAT(n, _c0k)
#include "trt_oracle_core.inc"
#undef FN
#undef COUNT
#undef CONV

#define CONV 1
#define COUNT 0
#define FN(n) CAT(n, _c1)
#include "trt_oracle_core.inc"
#undef FN
#undef COUNT
#define COUNT 1
#define FN(n) CAT(n, _c1k)
#include "trt_oracle_core.inc"
#undef FN
#undef COUNT
#undef CONV

/* ---- exported helpers ------------------------------------------------------------------------ */
uint32_t oracle_pcg_hash(uint32_t v) { return pcg_hash(v); }

void oracle_rng_stream(uint32_t x, uint32_t y, uint32_t w, uint32_t seed, uint32_t frame, uint32_t n,
                       uint32_t *states, float *values) {
    uint32_t rng = (y * w + x) * 1973u + seed * 9277u + frame * 12345u;   /* shader.wgsl:507 */
    rng = pcg_hash(rng + 0u * 5096u);                                      /* shader.wgsl:517, i = 0 */
    for (uint32_t k = 0; k < n; k++) {
        float v = rnd(&rng);
        if (states) states[k] = rng;
        if (values) values[k] = v;
    }
}

void oracle_sincos_pinned(float phi, float *s, float *c) { sincos_pinned(phi, s, c); }

/* camera.rs:17-25 (f32 cos/sin = libm cosf/sinf on Linux); up = right x forward (vec3.rs:23-29) */
void oracle_camera_vectors(float yaw, float pitch, float *f, float *r, float *u) {
    f[0] = cosf(yaw) * cosf(pitch);
    f[1] = sinf(pitch);
    f[2] = sinf(yaw) * cosf(pitch);
    r[0] = -sinf(yaw);
    r[1] = 0.0f;
    r[2] = cosf(yaw);
    u[0] = r[1] * f[2] - r[2] * f[1];
    u[1] = r[2] * f[0] - r[0] * f[2];
    u[2] = r[0] * f[1] - r[1] * f[0];
}

/* shader.wgsl:522: tan(f32(fov_rad) / 2.0) */
float oracle_tan_half(float fov_rad) { return tanf(fov_rad / 2.0f); }

static v3 mk3(const float *p) { v3 r = { p[0], p[1], p[2] }; return r; }

float oracle_hit_sphere(int conv, const float *o, const float *d, const float *c, float r, float tmin, float tmax) {
    int fr;
    return conv ? hit_sphere_c1(mk3(o), mk3(d), mk3(c), r, tmin, tmax, &fr)
                : hit_sphere_c0(mk3(o), mk3(d), mk3(c), r, tmin, tmax, &fr);
}
float oracle_hit_plane(int conv, const float *o, const float *d, const float *p, const float *n, float tmin, float tmax) {
    int st;
    return conv ? hit_plane_c1(mk3(o), mk3(d), mk3(p), mk3(n), tmin, tmax, &st)
                : hit_plane_c0(mk3(o), mk3(d), mk3(p), mk3(n), tmin, tmax, &st);
}
float oracle_hit_triangle(int conv, const float *o, const float *d, const float *a, const float *b, const float *c,
                          float tmin, float tmax) {
    int st;
    return conv ? hit_triangle_c1(mk3(o), mk3(d), mk3(a), mk3(b), mk3(c), tmin, tmax, &st)
                : hit_triangle_c0(mk3(o), mk3(d), mk3(a), mk3(b), mk3(c), tmin, tmax, &st);
}

int oracle_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

static void counters_add(ocounters *dst, const ocounters *src) {
    uint64_t *d = (uint64_t *)dst;
    const uint64_t *s = (const uint64_t *)src;
    for (size_t i = 0; i < sizeof(ocounters) / sizeof(uint64_t); i++) d[i] += s[i];
}

typedef uint32_t (*pixel_fn)(const ctx_t *, ocounters *, uint32_t, uint32_t, float *, float *, ovariance *);

int oracle_render(const oprim *prims, const ouniforms *u, int conv, int math_mode, int threads,
                  uint32_t y0, uint32_t y1, const uint32_t *idx, uint64_t n_idx,
                  float *accumulation, float *pixels, ovariance *variance, ocounters *counters,
                  uint64_t *samples_traced) {
    uint64_t traced_total = 0;
    if (!prims || !u || !accumulation || u->width == 0 || u->height == 0) return -1;
    if (conv != 0 && conv != 1) return -1;
    ctx_t cx = { prims, u, tanf(u->fov_rad / 2.0f), math_mode };
    pixel_fn fn = counters ? (conv ? pixel_c1k : pixel_c0k) : (conv ? pixel_c1 : pixel_c0);
#ifdef _OPENMP
    int nt = threads > 0 ? threads : omp_get_max_threads();
#else
    int nt = 1;
    (void)threads;
#endif
    const uint32_t w = u->width;
    if (idx) {
        ocounters total;
        memset(&total, 0, sizeof total);
#pragma omp parallel num_threads(nt)
        {
            ocounters local;
            uint64_t traced = 0;
            memset(&local, 0, sizeof local);
#pragma omp for schedule(dynamic, 16)
            for (int64_t k = 0; k < (int64_t)n_idx; k++) {
                uint32_t gi = idx[k];
                traced += fn(&cx, &local, gi % w, gi / w, accumulation + 4 * k, pixels ? pixels + 4 * k : NULL,
                             variance ? variance + k : NULL);
            }
#pragma omp critical
            { counters_add(&total, &local); traced_total += traced; }
        }
        if (counters) counters_add(counters, &total);
        if (samples_traced) *samples_traced = traced_total;
        return 0;
    }
    if (y1 > u->height) y1 = u->height;
    if (y0 > y1) return -1;
    ocounters total;
    memset(&total, 0, sizeof total);
    /* rows in parallel, dynamic schedule: the analogue of the reference's rayon into_par_iter over
     * rows (lib.rs:503); pixels are independent (shader.wgsl:499-621). */
#pragma omp parallel num_threads(nt)
    {
        ocounters local;
        uint64_t traced = 0;
        memset(&local, 0, sizeof local);
#pragma omp for schedule(dynamic, 1)
        for (int64_t y = y0; y < (int64_t)y1; y++) {
            for (uint32_t x = 0; x < w; x++) {
                size_t k = (size_t)y * w + x;
                traced += fn(&cx, &local, x, (uint32_t)y, accumulation + 4 * k, pixels ? pixels + 4 * k : NULL,
                             variance ? variance + k : NULL);
            }
        }
#pragma omp critical
        { counters_add(&total, &local); traced_total += traced; }
    }
    if (counters) counters_add(counters, &total);
    if (samples_traced) *samples_traced = traced_total;
    return 0;
}
