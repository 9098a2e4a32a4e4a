/*
 * wgsl_rt.h -- C run-time for the transliterated reference shader (TEST INFRASTRUCTURE; hand-written).
 *
 * oracle/wgsl2c.py turns /root/reference/src/shader.wgsl into C; every WGSL builtin it meets becomes a
 * wgsl_<name>_<type>() call defined here.  These are exactly the points where WGSL leaves the result to
 * the implementation (naga 25.0.1 + the GPU driver's compiler, Cargo.lock:567-568), so this file is where
 * the oracle's arithmetic convention ("conv 0") is stated for the reference build:
 *   + - * /          IEEE-754 binary32, round to nearest even, one rounding per operator; the translation
 *                    unit is compiled with -ffp-contract=off (WGSL allows 2.5 ulp for / and contraction)
 *   sqrt             IEEE sqrtf                         (WGSL: inherited from 1/inverseSqrt, ~2 ulp)
 *   sin, cos         wgsl_math_mode 0: the pinned implementation (trt_sincos_pinned.h), 1: glibc sinf/cosf
 *                                                        (WGSL: 2^-11 absolute on [-pi, pi])
 *   tan              glibc tanf
 *   min, max, abs    fminf / fmaxf / fabsf              (WGSL: either operand if one is NaN)
 *   f32(u32)         round to nearest                   (WGSL: correctly rounded)
 *   u32(f32), i32(f32)  truncate, clamped to the target range, NaN -> 0  (WGSL: clamped)
 *   integer / and %  x / 0 = x, x % 0 = 0               (WGSL)
 *   integer + - *    wrap modulo 2^32
 */
#ifndef WGSL_RT_H
#define WGSL_RT_H
#include <math.h>
#include <stdbool.h>
#include <stdint.h>
#include "trt_sincos_pinned.h"

typedef struct { uint32_t x, y, z; } wgsl_vec3u;

static int wgsl_math_mode = 0;      /* set by the driver before the parallel region; read-only inside */

static inline float wgsl_sqrt_f32(float x) { return sqrtf(x); }
static inline float wgsl_tan_f32(float x) { return tanf(x); }
static inline float wgsl_sin_f32(float x) {
    if (wgsl_math_mode == 1) return sinf(x);
    float s, c;
    sincos_pinned(x, &s, &c);
    return s;
}
static inline float wgsl_cos_f32(float x) {
    if (wgsl_math_mode == 1) return cosf(x);
    float s, c;
    sincos_pinned(x, &s, &c);
    return c;
}
static inline float wgsl_floor_f32(float x) { return floorf(x); }
static inline float wgsl_ceil_f32(float x) { return ceilf(x); }
static inline float wgsl_abs_f32(float x) { return fabsf(x); }
static inline int32_t wgsl_abs_i32(int32_t x) { return x < 0 ? (int32_t)(0u - (uint32_t)x) : x; }
static inline uint32_t wgsl_abs_u32(uint32_t x) { return x; }
static inline float wgsl_min_f32(float a, float b) { return fminf(a, b); }
static inline float wgsl_max_f32(float a, float b) { return fmaxf(a, b); }
static inline uint32_t wgsl_min_u32(uint32_t a, uint32_t b) { return a < b ? a : b; }
static inline uint32_t wgsl_max_u32(uint32_t a, uint32_t b) { return a > b ? a : b; }
static inline int32_t wgsl_min_i32(int32_t a, int32_t b) { return a < b ? a : b; }
static inline int32_t wgsl_max_i32(int32_t a, int32_t b) { return a > b ? a : b; }
static inline float wgsl_mod_f32(float a, float b) { return a - b * truncf(a / b); }

static inline uint32_t wgsl_div_u32(uint32_t a, uint32_t b) { return b ? a / b : a; }
static inline uint32_t wgsl_mod_u32(uint32_t a, uint32_t b) { return b ? a % b : 0u; }
static inline int32_t wgsl_div_i32(int32_t a, int32_t b) {
    return (b == 0 || (a == INT32_MIN && b == -1)) ? a : a / b;
}
static inline int32_t wgsl_mod_i32(int32_t a, int32_t b) {
    return (b == 0 || (a == INT32_MIN && b == -1)) ? 0 : a % b;
}

static inline float wgsl_f32_from_u32(uint32_t v) { return (float)v; }
static inline float wgsl_f32_from_i32(int32_t v) { return (float)v; }
static inline float wgsl_f32_from_bool(bool v) { return v ? 1.0f : 0.0f; }
static inline uint32_t wgsl_u32_from_f32(float v) {
    if (!(v > 0.0f)) return 0u;                         /* negatives, -0, NaN */
    if (v >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)v;
}
static inline int32_t wgsl_i32_from_f32(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;
}
static inline uint32_t wgsl_u32_from_i32(int32_t v) { return (uint32_t)v; }
static inline int32_t wgsl_i32_from_u32(uint32_t v) { return (int32_t)v; }
static inline uint32_t wgsl_u32_from_bool(bool v) { return v ? 1u : 0u; }
static inline int32_t wgsl_i32_from_bool(bool v) { return v ? 1 : 0; }

#endif
