/*
 * ref_driver.c -- host side of oracle/_ref/libref.so: "the reference's shader compiled here"
 * (TEST INFRASTRUCTURE; hand-written; the only file that includes the generated transliteration).
 *
 * _ref/shader_wgsl.inc is produced at build time by oracle/wgsl2c.py from /root/reference/src/shader.wgsl
 * (never committed, never copied).  This driver plays the part of the wgpu host code around it:
 *   lib.rs:442        queue.write_buffer(uniform)         -> memcpy of the 112-byte shader prefix of Uniforms
 *   lib.rs:301-305    primitive storage buffer            -> g_primitives points at the caller's 180-byte records
 *   lib.rs:450-453    dispatch_workgroups(ceil(w/8), ceil(h/8), 1) of @workgroup_size(8,8,1)
 *                                                         -> every invocation id of that grid calls w_main();
 *                                                            the ids beyond the frame return at shader.wgsl:503
 *   bindings 2..4     pixels / accumulation / variance    -> the caller's arrays
 * Invocations are independent (each touches only its own index), so they are spread over OpenMP threads
 * row by row, the analogue of the GPU's parallel dispatch.  var<private> rand_state is _Thread_local.
 *
 * The signature mirrors oracle_render() (trt_oracle.h) so that tests can run both on the same arguments.
 */
#include <stddef.h>
#include <stdint.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "wgsl_rt.h"
#include "_ref/shader_wgsl.inc"

#if !defined(WGSL_ENTRY_main) || !defined(WGSL_BINDING_0_uniforms) || !defined(WGSL_BINDING_1_primitives) || \
    !defined(WGSL_BINDING_2_pixels) || !defined(WGSL_BINDING_3_accumulation) || !defined(WGSL_BINDING_4_variance)
#error "shader.wgsl no longer has the entry point / bindings this driver was written for (shader.wgsl:76-80,499)"
#endif
_Static_assert(sizeof(W_Vec3) == 16, "Vec3 is 16 bytes (vec3.rs:5-12)");
_Static_assert(sizeof(W_Primitive) == 180, "Primitive is 180 bytes (primitive.rs:4-34)");
_Static_assert(sizeof(W_Uniforms) == 112, "the shader's Uniforms is the 112-byte prefix of lib.rs:23-50");
_Static_assert(sizeof(W_VarianceBuffer) == 8, "VarianceBuffer is 8 bytes");

/* Point a storage binding at `base` so that element [index] of the shader is element [k] of the caller's
 * compact array (idx mode).  Integer arithmetic on the address, not out-of-bounds pointer arithmetic. */
#define REBASE(type, base, k, index) ((type *)((uintptr_t)(base) + ((uintptr_t)(k) - (uintptr_t)(index)) * sizeof(type)))

int ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

uint32_t ref_pcg_hash(uint32_t v) { return w_pcg_hash(v); }

/* rand_state after main()'s seeding for sample 0 is not separable from main; this KAT entry repeats the two
 * statements shader.wgsl:507 and :517 around the transliterated pcg_hash / random_f32. */
void ref_rng_stream(uint32_t x, uint32_t y, uint32_t w, uint32_t seed, uint32_t frame, uint32_t n,
                    uint32_t *states, float *values) {
    g_rand_state = (y * w + x) * 1973u + seed * 9277u + frame * 12345u;
    g_rand_state = w_pcg_hash(g_rand_state + 0u * 5096u);
    for (uint32_t k = 0; k < n; k++) {
        float v = w_random_f32();
        if (states) states[k] = g_rand_state;
        if (values) values[k] = v;
    }
}

static W_Ray mk_ray(const float *o, const float *d) {
    W_Ray r;
    memset(&r, 0, sizeof r);
    r.origin.x = o[0]; r.origin.y = o[1]; r.origin.z = o[2];
    r.direction.x = d[0]; r.direction.y = d[1]; r.direction.z = d[2];
    return r;
}
/* intersection KAT entries: the transliterated hit_sphere / hit_plane / hit_triangle on one 180-byte primitive */
float ref_hit(const void *prim180, const float *o3, const float *d3, float tmin, float tmax) {
    W_Primitive p;
    memcpy(&p, prim180, sizeof p);
    W_Ray r = mk_ray(o3, d3);
    if (p.primitive_type == 0u) return w_hit_sphere(r, p, tmin, tmax);
    if (p.primitive_type == 1u) return w_hit_plane(r, p, tmin, tmax);
    return w_hit_triangle(r, p, tmin, tmax);
}

/*
 * One dispatch of the shader.  idx == NULL: rows [y0, y1) of the frame, arrays are full w*h; idx != NULL:
 * n_idx global pixel indices, arrays are compact.  pixels / variance may be NULL (the shader still writes
 * them: into scratch).  accumulation is read when frame_number != 0.  Returns 0, -1 on bad arguments.
 */
int ref_render(const void *prims180, const void *uniforms, int math_mode, int threads,
               uint32_t y0, uint32_t y1, const uint32_t *idx, uint64_t n_idx,
               float *accumulation, float *pixels, void *variance) {
    if (!prims180 || !uniforms || !accumulation) return -1;
    memcpy(&g_uniforms, uniforms, sizeof g_uniforms);
    if (g_uniforms.width == 0 || g_uniforms.height == 0) return -1;
    wgsl_math_mode = math_mode;
    const uint32_t w = g_uniforms.width, h = g_uniforms.height;
#ifdef _OPENMP
    int nt = threads > 0 ? threads : omp_get_max_threads();
#else
    int nt = 1;
    (void)threads;
#endif
    if (idx) {
#pragma omp parallel num_threads(nt)
        {
            W_Vec3 scratch_px;
            W_VarianceBuffer scratch_var;
            g_primitives = (W_Primitive *)prims180;
#pragma omp for schedule(dynamic, 16)
            for (int64_t k = 0; k < (int64_t)n_idx; k++) {
                uint32_t gi = idx[k];
                g_accumulation = REBASE(W_Vec3, accumulation, k, gi);
                g_pixels = pixels ? REBASE(W_Vec3, pixels, k, gi) : REBASE(W_Vec3, &scratch_px, 0, gi);
                g_variance = variance ? REBASE(W_VarianceBuffer, variance, k, gi)
                                      : REBASE(W_VarianceBuffer, &scratch_var, 0, gi);
                wgsl_vec3u id = { gi % w, gi / w, 0u };
                w_main(id);
            }
        }
        return 0;
    }
    if (y1 > h) y1 = h;
    if (y0 > y1) return -1;
    /* the grid of lib.rs:450-453 restricted to rows [y0, y1): x runs over ceil(w/8)*8 invocation ids; when the
     * range ends at the frame's last row the ids below it (up to ceil(h/8)*8) are dispatched too. */
    const uint32_t gx = (w + 7u) / 8u * 8u;
    const uint32_t gy_end = (y1 == h) ? (h + 7u) / 8u * 8u : y1;
#pragma omp parallel num_threads(nt)
    {
        W_Vec3 scratch_px;
        W_VarianceBuffer scratch_var;
        g_primitives = (W_Primitive *)prims180;
        g_accumulation = (W_Vec3 *)accumulation;
#pragma omp for schedule(dynamic, 1)
        for (int64_t y = y0; y < (int64_t)gy_end; y++) {
            for (uint32_t x = 0; x < gx; x++) {
                uint32_t gi = (uint32_t)y * w + x;
                g_pixels = pixels ? (W_Vec3 *)pixels : REBASE(W_Vec3, &scratch_px, 0, gi);
                g_variance = variance ? (W_VarianceBuffer *)variance : REBASE(W_VarianceBuffer, &scratch_var, 0, gi);
                wgsl_vec3u id = { x, (uint32_t)y, 0u };
                w_main(id);
            }
        }
    }
    return 0;
}
