/*
 * trt_oracle.h -- CPU ORACLE for the Terminal-Raytracer per-pixel path-tracing hot path.
 *
 * THIS IS TEST INFRASTRUCTURE, NOT PRODUCT CODE.  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it.  The product (libtrt_b200.so)
 * never links, loads or calls anything in oracle/.
 *
 * PARITY PIN: the reference (Gooey321/Terminal-Raytracer) has no tests and no golden vectors, and it cannot be
 * run through wgpu here (no Rust toolchain, no graphics adapter).  Its path tracer, however, is one WGSL file of
 * plain scalar code, and oracle/_ref/libref.so is that file compiled here: oracle/wgsl2c.py transliterates
 * /root/reference/src/shader.wgsl to C at build time (one statement per statement; the generated text never
 * enters the repo) and oracle/Makefile compiles it with -ffp-contract=off.  tests/test_oracle_ref.py shows this
 * hand-written restatement (conv 0) returns the SAME BITS as that build -- pixels, accumulation, variance -- on
 * the three shipped scenes, spp 1..128, depth 0..32, frames 0..3 and a reduced C4 scene, and the conv 0 golden
 * fixtures (tests/golden/frames.*) are outputs of that build.  What remains a hand restatement pinned only by the
 * source text: the Rust quantiser (src/lib.rs:503-529, trt_oracle_quant.c) and the host data path (oracle.py);
 * the implementation-defined arithmetic of WGSL (sqrt / sin / cos / tan / division accuracy, contraction) is
 * fixed by the conventions below for both builds alike.
 *
 * Arithmetic conventions (the reference leaves them to naga + the GPU driver, so they are chosen
 * here and stated; see DESIGN.md "Arithmetic conventions"):
 *   conv 0 ("as written"): IEEE-754 binary32, round-to-nearest-even, one rounding per WGSL
 *           operator, no contraction.
 *   conv 1 ("fused helpers"): same, except that the vector helpers dot / cross / a+b*s /
 *           a*s+b*t+c*u and three scalar a*b+-c patterns in the sphere test are evaluated with
 *           fused multiply-adds in a fixed, documented order (what a driver compiler does to the
 *           same WGSL when it contracts).
 *   sin/cos: `trt_sincos_pinned` -- a documented Cody-Waite + minimax polynomial (<= 2 ulp from
 *           glibc sinf/cosf on [0, 2*pi], checked in tests) so that CPU and GPU agree bit for bit;
 *           math_mode 1 switches the oracle to glibc sinf/cosf for a cross-check.
 *   tan(fov/2), camera cos/sin, quantiser powf/sqrt: glibc (what Rust's f32 methods call on Linux).
 */
#ifndef TRT_ORACLE_H
#define TRT_ORACLE_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* vec3.rs:5-12 -- 16-byte padded float3 */
typedef struct { float x, y, z, pad; } ovec3;

/* primitive.rs:4-34 / shader.wgsl:8-32 -- 180-byte tagged primitive */
typedef struct {
    uint32_t type, pad1, pad2, pad3;        /* 0 sphere, 1 plane, 2 triangle */
    ovec3 color, emission;
    float reflectivity, pad4, pad5, pad6;
    ovec3 sphere_center;
    float sphere_radius, sphere_pad[3];
    ovec3 plane_point, plane_normal;
    ovec3 v0, v1, v2;
    float tri_pad;
} oprim;

/* lib.rs:23-50 (host, 160 B); the shader reads the first 112 B (shader.wgsl:49-69) */
typedef struct {
    uint32_t width, height, samples_per_pixel, max_depth;
    uint32_t seed, frame_number, primitive_count, pad0;
    float aspect_ratio, char_aspect_ratio, fov_rad, pad1;
    ovec3 camera_pos, camera_forward, camera_right, camera_up;
    ovec3 grid_min, inv_cell_size;          /* dead fields (uniform grid never read) */
    uint32_t grid_dims[3], pad3;
} ouniforms;

/* shader.wgsl:71-74 */
typedef struct { float variance; uint32_t sample_count; } ovariance;

/* Work counters; the FLOP cost table is DESIGN.md "Algorithmic FLOPs" (SURVEY 8(d)). */
typedef struct {
    uint64_t pixels;
    uint64_t samples_base, samples_extra;
    uint64_t closest_scans;      /* hit_scene calls from ray_color            */
    uint64_t hits, misses;
    uint64_t shadow_scans;       /* hit_scene calls from sample_direct_lighting */
    uint64_t shadow_unoccluded, shadow_lit;   /* lit = both cosines > 0 and unoccluded */
    uint64_t shadow_needed;      /* both cosines > 0 (occluded or not)        */
    uint64_t mirror, diffuse, roulette_tests, roulette_kills;
    uint64_t prim_tests;         /* primitive tests executed as written (closest + shadow) */
    uint64_t prim_tests_min;     /* tests in closest scans + needed shadow scans up to 1st occluder */
    uint64_t flops_written, flops_min;
    uint64_t rng_draws;
} ocounters;

uint32_t oracle_pcg_hash(uint32_t v);
/* states and floats produced by main()'s seeding for sample 0 (shader.wgsl:507,517) + n draws */
void oracle_rng_stream(uint32_t x, uint32_t y, uint32_t w, uint32_t seed, uint32_t frame,
                       uint32_t n, uint32_t *states, float *values);
void oracle_sincos_pinned(float phi, float *s, float *c);
void oracle_camera_vectors(float yaw, float pitch, float *fwd3, float *right3, float *up3);
float oracle_tan_half(float fov_rad);

/* intersection KAT entry points; return t or -1 (shader.wgsl:155-223) */
float oracle_hit_sphere(int conv, const float *o3, const float *d3, const float *c3, float r,
                        float tmin, float tmax);
float oracle_hit_plane(int conv, const float *o3, const float *d3, const float *p3, const float *n3,
                       float tmin, float tmax);
float oracle_hit_triangle(int conv, const float *o3, const float *d3, const float *v0,
                          const float *v1, const float *v2, float tmin, float tmax);

/*
 * One frame of shader.wgsl `main` over a set of pixels.
 *  idx == NULL: pixels are all (x, y) with y in [y0, y1); accumulation/pixels/variance are full
 *               w*h arrays indexed y*w+x.
 *  idx != NULL: n_idx global pixel indices; the three arrays are compact (entry k <-> idx[k]).
 * accumulation is read when frame_number != 0 and always written; pixels/variance/counters may be NULL.
 * conv: 0 or 1; math_mode: 0 pinned sin/cos, 1 glibc; threads: OpenMP threads (<=0: all).
 * Returns 0, or -1 on bad arguments.
 */
int oracle_render(const oprim *prims, const ouniforms *u, int conv, int math_mode, int threads,
                  uint32_t y0, uint32_t y1, const uint32_t *idx, uint64_t n_idx,
                  float *accumulation, float *pixels, ovariance *variance, ocounters *counters,
                  uint64_t *samples_traced /* optional: ray_color calls made */);

/* lib.rs:503-532: rows block of one frame.  mode 0 ASCII, 1 truecolor.  Returns bytes needed
 * (writes at most cap).  row_lengths (h entries) optional. */
size_t oracle_quantise(const float *pixels4, uint32_t w, uint32_t h, int mode, int threads,
                       uint8_t *out, size_t cap, uint32_t *row_lengths);
/* number of pixels whose pre-truncation channel/glyph value lies within eps of an integer */
uint64_t oracle_quantise_boundary_count(const float *pixels4, uint32_t w, uint32_t h, int mode,
                                        float eps);
int oracle_max_threads(void);

#ifdef __cplusplus
}
#endif
#endif
