// trt_device.cuh -- device-side data layout and math of the sm_100a path tracer.
//
// Written from scratch for B200; behaviour follows /root/reference/src/shader.wgsl (cited per
// function).  This translation unit is compiled with --fmad=false: nvcc never contracts a*b+c on
// its own, so TRT_CONV_AS_WRITTEN is one IEEE rounding per WGSL operator, and TRT_CONV_FUSED fuses
// exactly where __fmaf_rn is written below (DESIGN.md "Arithmetic conventions").
#pragma once
#ifdef __CUDACC_RTC__
// run-time compilation (trt_jit.cpp): NVRTC has the CUDA built-ins but no host headers
typedef unsigned char uint8_t;
typedef int int32_t;
typedef unsigned int uint32_t;
typedef unsigned long long uint64_t;
struct trt_variance { float variance; uint32_t sample_count; };     // include/trt_b200.h
#define TRT_CONSTANT(decl) extern "C" { __constant__ decl; }        // findable by name in the loaded module
#else
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/trt_b200.h"
#define TRT_CONSTANT(decl) static __constant__ decl
#endif

namespace trt {

// ---- scene in HBM: structure of arrays, one float4 per field per primitive ----------------------
// (the reference scans a 180-byte AoS struct per test, shader.wgsl:255; here a test touches 32-48 B)
//   sphere   : ga = (cx, cy, cz, r)     gb = (r*r, -, -, -)
//   plane    : ga = (px, py, pz, -)     gb = (nx, ny, nz, -) raw     gn = normalize(n)
//   triangle : ga = v0                  gb = e1 = v1-v0   gc = e2 = v2-v0   gn = normalize(e1 x e2)
// material   : mc = (color.rgb, reflectivity)   me = (emission.rgb, type as float bits)
// Every derived value (r*r, e1, e2, normals, light area) is computed with the same f32 operation
// sequence the shader uses, so it is bit-identical to recomputing it per test.
struct SceneView {
    const float4 *ga, *gb, *gc, *gn, *mc, *me;
    // triangle and sphere tiles of the cooperative scan, two primitives per entry pair: (x_a, x_b, y_a, y_b), (z_a, z_b,
    // -R_a, -R_b) with (x, y, z) = v0 / the centre and R = ga.w / r^2 (the cull of scan_run on packed FP32 pairs); a
    // tile's entries start at tiles[k].w
    const float4 *gd;
    const uint8_t *type;        // per primitive: 0 sphere, 1 plane, 2 triangle, 255 unknown
    const int4 *runs;           // (type, begin, end, -) maximal runs of equal type, in scan order
    const int4 *lights;         // (prim index, type, area as float bits, -) emissive spheres/triangles in scan order
    int n_prims, n_runs, n_lights;
    // tiles of at most TRT_TILE consecutive primitives of one run, in scan order: (type, first index, count, first gd entry);
    // the CTA-cooperative scan of large scenes stages one tile at a time in shared memory (trt_trace.cuh)
    const int4 *tiles;
    int n_tiles;
    // opt-in acceleration structure (trt_config.accel = TRT_ACCEL_GRID; OUTSIDE the parity contract's linear scan, though
    // it returns the same (t, index)): the reference's abandoned uniform grid (lib.rs:181-244) over spheres and
    // triangles -- CSR cell lists in scan order -- plus the list of planes (unbounded: always tested).  NULL = none.
    const uint32_t *grid_off, *grid_idx;
    const int *grid_planes;
    int n_grid_planes;
    int gx, gy, gz;
    float3 grid_min, grid_cell, grid_inv_cell;
};
#define TRT_TILE 96             // x 4 arrays (ga, gb, gc, gd) x 2 buffers x 16 B = 12 KB of shared memory per CTA

#define TRT_CONST_PRIMS 256     // scenes up to this size are also mirrored in __constant__ memory

#define TRT_STATS_BYTES (8 * (8 + 2 * 512))     // 8 words of frame statistics + two 512-bin histograms (measurement builds)
struct FrameParams {
    uint32_t width, height, spp, max_depth;
    uint32_t seed, frame_number, base_samples, n_local_pixels;
    uint32_t all_ones;                                  // 0xffffffff, as a value the compiler cannot fold (TRT_V_KEY: a mask kept in a register)
    float char_aspect, half_h, half_w;                  // half_h = tan(fov/2) (host, glibc), half_w = aspect*half_h
    float wm1, hm1;                                     // f32(width-1), f32(height-1)
    float spp_f, base_f;
    // conv 2 camera ray with the frame-uniform factors folded on the host (TRT_V_CAMFOLD): view-plane coordinates
    // vx = (x + r) * cam_kx - half_w, vy = (height-1-y + r) * cam_ky - cam_cy  (shader.wgsl:519-531 multiplied out)
    float cam_kx, cam_ky, cam_cy;
    float3 cam_pos, cam_fwd, cam_right, cam_up;
    const uint32_t *row_map;     // local row -> global row (row sharding); NULL = identity
    const uint32_t *order;       // hand-out order of the local pixels (longest first, from the previous frame's variance); NULL = natural
    float4 *pixels, *accumulation;
    trt_variance *variance;
    unsigned long long *counters;   // trt_counters as 19 x u64 (instrumented build only)
    unsigned int *work_counter;     // persistent kernel: next pixel to hand out
    unsigned long long *stats;      // [0] samples traced, [1] closest-hit scans (bounces) of this frame; [2..]: diagnostics
    // frame batch (trt_render_frames): n_batch frames that differ only in (seed, frame_number) are traced by one
    // launch; work item = (frame, local pixel); a finished pixel stores its mean into batch_cur[frame][local pixel]
    // and blend_frames_kernel applies the temporal blend (shader.wgsl:610-619) in frame order afterwards
    const uint2 *batch;             // (seed, frame_number) per frame; NULL: single frame
    float4 *batch_cur;
    uint32_t n_batch;
    uint32_t ctas_hint;             // host only: resident CTAs per SM the launcher should use (0: all that fit)
    // bounce count of every pixel of this frame (global pixel index), written by the production kernel: the measured
    // length of the pixel's serial chain, which orders the next frame's hand-out (order_*_kernel); NULL: not recorded
    uint32_t *cost;
};

// shader.wgsl:610-619: the temporal blend of one pixel, with explicit roundings so that every call site
// (finish_pixel, blend_frames_kernel) produces the same bits in a given convention
template <int CONV> __device__ __forceinline__ float4 blend_pixel(float4 a, float3 cur, uint32_t frame_number) {
    if (frame_number == 0u) return make_float4(cur.x, cur.y, cur.z, 0.0f);
    const float n = __uint2float_rn(frame_number + 1u);
    float alpha;
    if (CONV == 2) asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(alpha) : "f"(n));
    else alpha = __fdiv_rn(1.0f, n);
    const float oma = __fsub_rn(1.0f, alpha);
    if (CONV == 2)
        return make_float4(__fmaf_rn(cur.x, alpha, __fmul_rn(a.x, oma)), __fmaf_rn(cur.y, alpha, __fmul_rn(a.y, oma)),
                           __fmaf_rn(cur.z, alpha, __fmul_rn(a.z, oma)), 0.0f);
    return make_float4(__fadd_rn(__fmul_rn(a.x, oma), __fmul_rn(cur.x, alpha)), __fadd_rn(__fmul_rn(a.y, oma), __fmul_rn(cur.y, alpha)),
                       __fadd_rn(__fmul_rn(a.z, oma), __fmul_rn(cur.z, alpha)), 0.0f);
}

__device__ __forceinline__ float3 f3(float x, float y, float z) { return make_float3(x, y, z); }
__device__ __forceinline__ float3 xyz(float4 v) { return make_float3(v.x, v.y, v.z); }

// ---- correctly rounded division of three numerators by one divisor -------------------------------
// vec3_div (shader.wgsl:99-101) is `a.x / s, a.y / s, a.z / s`: three IEEE divisions.  Compiled as
// written, each one is MUFU.RCP + Newton steps + FCHK, and FCHK sends every zero numerator (axis-
// aligned normals: most of the Cornell box) to a ~60-instruction slow-path subroutine (ncu:
// profiles/r1_trace_wave_pre_div.txt, 32 % of all instructions).  div3_rn computes the same three
// correctly rounded quotients with ONE reciprocal:
//     y  = rcp(s) refined once           (error < 1 ulp)
//     q0 = RN(a*y);  r = a - s*q0 (exact, FMA);  q = RN(q0 + r*y)      -- Markstein's correction step
// which is RN(a/s) whenever a, s and a/s are normal numbers away from the exponent limits, and keeps
// the signed zero of a zero numerator (q0).  Anything else (tiny, huge, infinite, NaN) goes to the
// plain IEEE division in div3_slow.  tests/test_gpu_division.py checks it against `/` bit for bit.
static __device__ __noinline__ float3 div3_slow(float ax, float ay, float az, float s) { return make_float3(ax / s, ay / s, az / s); }

__device__ __forceinline__ float3 div3_rn(float3 a, float s) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(s));
    y = __fmaf_rn(y, __fmaf_rn(-s, y, 1.0f), y);
    const float qx = a.x * y, qy = a.y * y, qz = a.z * y;
    float ox = __fmaf_rn(__fmaf_rn(-s, qx, a.x), y, qx);
    float oy = __fmaf_rn(__fmaf_rn(-s, qy, a.y), y, qy);
    float oz = __fmaf_rn(__fmaf_rn(-s, qz, a.z), y, qz);
    ox = (a.x == 0.0f) ? qx : ox;
    oy = (a.y == 0.0f) ? qy : oy;
    oz = (a.z == 0.0f) ? qz : oz;
    // guard: |s| in [2^-40, 2^40); every numerator is zero or has |a| in [2^-60, 2^60)
    // (bits << 1 drops the sign; "- 1" sends zero to 0xffffffff so that it passes the lower bound)
    const uint32_t us = __float_as_uint(s) << 1;
    const uint32_t ux = (__float_as_uint(a.x) << 1) - 1u, uy = (__float_as_uint(a.y) << 1) - 1u, uz = (__float_as_uint(a.z) << 1) - 1u;
    const uint32_t lo = min(min(ux, uy), uz);
    const float big = fmaxf(fmaxf(fabsf(a.x), fabsf(a.y)), fabsf(a.z));
    const bool ok = (us - 0x57000000u < 0x50000000u) && lo >= 0x43000000u - 1u && big < 1.152921504606846976e18f;
    if (!ok) return div3_slow(a.x, a.y, a.z, s);
    return make_float3(ox, oy, oz);
}

// ---- conv 2 "device native": the arithmetic a GPU shader compiler gives the same WGSL --------------
// WGSL only bounds the error of `/` (2.5 ulp), sqrt / inverseSqrt (2 ulp), sin / cos (2^-11 absolute)
// and allows fused multiply-add everywhere; conv 2 uses the SM's native units within those bounds:
// MUFU.RCP / RSQ / SQRT / SIN / COS (1-2 ulp) and FFMA contraction (this convention is only
// instantiated in trt_native.cu, which is compiled with --fmad=true -use_fast_math).
// Packed FP32 pairs (sm_100: add / mul / fma .f32x2 -> FADD2 / FMUL2 / FFMA2, two IEEE operations per issue slot; the FMA
// pipe needs two cycles for them, so the FLOP rate is that of the scalar forms -- measured 73.4 against 70.9 TFLOP/s,
// tools/micro/ffma2_rate.cu -- but a kernel bound by ISSUE slots gets the second operation for free).
typedef unsigned long long f2_t;
__device__ __forceinline__ f2_t f2_pack(float lo, float hi) { f2_t r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void f2_unpack(f2_t v, float &lo, float &hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ f2_t f2_add(f2_t a, f2_t b) { f2_t r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_sub(f2_t a, f2_t b) { f2_t r; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_mul(f2_t a, f2_t b) { f2_t r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2_fma(f2_t a, f2_t b, f2_t c) { f2_t r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ float rcp_native(float x) { float y; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float rsqrt_native(float x) { float y; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ float sqrt_native(float x) { float y; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }

// TRT_V_PACKED: conv 2's element-wise float3 helpers put the x and y components through ONE packed instruction (FADD2 /
// FMUL2 / FFMA2, a scalar operand is broadcast for free) and z through a scalar one: two issue slots instead of three.
// Every lane's operation is the IEEE operation of the scalar form (the products and sums conv 2 contracts to FFMA are
// written as fma here), so only values the compiler used to contract differently can move by an ulp.
#ifndef TRT_V_PACKED
#define TRT_V_PACKED 1
#endif
// The strict conventions pack ADDITIONS AND SUBTRACTIONS only (and, in conv 1, the helpers that are one fused multiply-add
// by definition).  ptxas 12.9 contracts `mul.rn.f32x2` followed by `add.rn.f32x2` into FFMA2 whatever --fmad says, and
// even sees through fma(a, b, -0.0) used as an exact product (tools/micro: z.cu, z2.cu) -- a packed product feeding a
// sum would silently lose a rounding.  A packed lane of an addition is bit for bit the scalar IEEE addition.
// Measured (profiles/r2_ab_packed_fp32.txt): bit-exact (the strict suite passes), pool machine 81.6 -> 79.9 ms, one pixel per
// lane 85.1 -> 85.6 ms on the A/B frame: not worth depending on what ptxas leaves alone -- off.
#ifndef TRT_V_PACKED_STRICT
#define TRT_V_PACKED_STRICT 0
#endif
__device__ __forceinline__ f2_t f2z_add(f2_t a, f2_t b) { f2_t r; asm("add.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2z_sub(f2_t a, f2_t b) { f2_t r; asm("sub.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2z_mul(f2_t a, f2_t b) { f2_t r; asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ f2_t f2z_fma(f2_t a, f2_t b, f2_t c) { f2_t r; asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
// by convention: conv 2 flushes denormals like the rest of its translation unit (-use_fast_math); the strict conventions
// use the plain IEEE forms, so a packed lane is bit for bit the scalar operation it replaces
template <int CONV> __device__ __forceinline__ f2_t p_add(f2_t a, f2_t b) { return CONV == 2 ? f2z_add(a, b) : f2_add(a, b); }
template <int CONV> __device__ __forceinline__ f2_t p_sub(f2_t a, f2_t b) { return CONV == 2 ? f2z_sub(a, b) : f2_sub(a, b); }
template <int CONV> __device__ __forceinline__ f2_t p_mul(f2_t a, f2_t b) { return CONV == 2 ? f2z_mul(a, b) : f2_mul(a, b); }
template <int CONV> __device__ __forceinline__ f2_t p_fma(f2_t a, f2_t b, f2_t c) { return CONV == 2 ? f2z_fma(a, b, c) : f2_fma(a, b, c); }
__device__ __forceinline__ f2_t xy2(float3 a) { return f2_pack(a.x, a.y); }
__device__ __forceinline__ f2_t ss2(float s) { return f2_pack(s, s); }
__device__ __forceinline__ float3 f3z(f2_t xy, float z) { float x, y; f2_unpack(xy, x, y); return f3(x, y, z); }

// ---- vector helpers (shader.wgsl:83-101,126-140,225-232) ----------------------------------------
template <int CONV> struct M {
    static constexpr bool kPacked = CONV == 2 && TRT_V_PACKED != 0;                 // products, sums and fused forms
    static constexpr bool kPackedAdd = kPacked || (CONV != 2 && TRT_V_PACKED_STRICT != 0);   // sums only
    static constexpr bool kPackedFma = kPacked || (CONV == 1 && TRT_V_PACKED_STRICT != 0);   // helpers that ARE one fma
    static __device__ __forceinline__ float3 add(float3 a, float3 b) {
        if (kPackedAdd) return f3z(p_add<CONV>(xy2(a), xy2(b)), a.z + b.z);
        return f3(a.x + b.x, a.y + b.y, a.z + b.z);
    }
    static __device__ __forceinline__ float3 sub(float3 a, float3 b) {
        if (kPackedAdd) return f3z(p_sub<CONV>(xy2(a), xy2(b)), a.z - b.z);
        return f3(a.x - b.x, a.y - b.y, a.z - b.z);
    }
    static __device__ __forceinline__ float3 scale(float3 a, float s) {
        if (kPacked) return f3z(p_mul<CONV>(xy2(a), ss2(s)), a.z * s);
        return f3(a.x * s, a.y * s, a.z * s);
    }
    static __device__ __forceinline__ float3 mulv(float3 a, float3 b) {
        if (kPacked) return f3z(p_mul<CONV>(xy2(a), xy2(b)), a.z * b.z);
        return f3(a.x * b.x, a.y * b.y, a.z * b.z);
    }
    static __device__ __forceinline__ float3 divs(float3 a, float s) {                              // == a.x/s, a.y/s, a.z/s
        if (CONV == 2) { const float y = rcp_native(s); return scale(a, y); }
        return div3_rn(a, s);
    }
    static __device__ __forceinline__ float div(float a, float b) { return CONV == 2 ? a * rcp_native(b) : a / b; }
    static __device__ __forceinline__ float sqrt(float a) { return CONV == 2 ? sqrt_native(a) : sqrtf(a); }
    static __device__ __forceinline__ float dot(float3 a, float3 b) {
        if (CONV) return __fmaf_rn(a.z, b.z, __fmaf_rn(a.y, b.y, a.x * b.x));
        return a.x * b.x + a.y * b.y + a.z * b.z;
    }
    static __device__ __forceinline__ float3 cross(float3 a, float3 b) {
        if (CONV) return f3(__fmaf_rn(a.y, b.z, -(a.z * b.y)), __fmaf_rn(a.z, b.x, -(a.x * b.z)), __fmaf_rn(a.x, b.y, -(a.y * b.x)));
        return f3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
    }
    static __device__ __forceinline__ float3 axpy(float3 a, float3 b, float s) {          // a + b*s
        if (kPackedFma) return f3z(p_fma<CONV>(xy2(b), ss2(s), xy2(a)), __fmaf_rn(b.z, s, a.z));
        if (CONV) return f3(__fmaf_rn(b.x, s, a.x), __fmaf_rn(b.y, s, a.y), __fmaf_rn(b.z, s, a.z));
        return f3(a.x + b.x * s, a.y + b.y * s, a.z + b.z * s);
    }
    static __device__ __forceinline__ float3 vmad(float3 a, float3 b, float3 c) {         // a + b*c
        if (kPackedFma) return f3z(p_fma<CONV>(xy2(b), xy2(c), xy2(a)), __fmaf_rn(b.z, c.z, a.z));
        if (CONV) return f3(__fmaf_rn(b.x, c.x, a.x), __fmaf_rn(b.y, c.y, a.y), __fmaf_rn(b.z, c.z, a.z));
        return f3(a.x + b.x * c.x, a.y + b.y * c.y, a.z + b.z * c.z);
    }
    static __device__ __forceinline__ float3 lin3(float3 a, float s, float3 b, float t, float3 c, float u) {
        if (kPacked)
            return f3z(p_fma<CONV>(xy2(c), ss2(u), p_fma<CONV>(xy2(b), ss2(t), p_mul<CONV>(xy2(a), ss2(s)))), __fmaf_rn(c.z, u, __fmaf_rn(b.z, t, a.z * s)));
        if (CONV)
            return f3(__fmaf_rn(c.x, u, __fmaf_rn(b.x, t, a.x * s)), __fmaf_rn(c.y, u, __fmaf_rn(b.y, t, a.y * s)),
                      __fmaf_rn(c.z, u, __fmaf_rn(b.z, t, a.z * s)));
        return f3(a.x * s + b.x * t + c.x * u, a.y * s + b.y * t + c.y * u, a.z * s + b.z * t + c.z * u);
    }
    static __device__ __forceinline__ float length(float3 v) { return sqrt(dot(v, v)); }
    static __device__ __forceinline__ float3 normalize(float3 v) {
        if (CONV == 2) { const float y = rsqrt_native(dot(v, v)); return scale(v, y); }
        return divs(v, length(v));
    }
    static __device__ __forceinline__ float3 reflect(float3 v, float3 n) {                 // shader.wgsl:138-140
        float k = 2.0f * dot(v, n);
        if (kPackedFma) return f3z(p_fma<CONV>(xy2(n), ss2(-k), xy2(v)), __fmaf_rn(n.z, -k, v.z));
        if (CONV) return f3(__fmaf_rn(n.x, -k, v.x), __fmaf_rn(n.y, -k, v.y), __fmaf_rn(n.z, -k, v.z));
        return sub(v, scale(n, k));
    }
    static __device__ __forceinline__ float msub(float a, float b, float c) {              // a*b - c
        if (CONV) return __fmaf_rn(a, b, -c);
        return a * b - c;
    }
    static __device__ __forceinline__ float nmadd(float c, float a, float b) {             // c - a*b
        if (CONV) return __fmaf_rn(-a, b, c);
        return c - a * b;
    }
};

#ifndef TRT_V_PCG_MULHI
#define TRT_V_PCG_MULHI 0   // measured: slower (IMAD.HI is no cheaper than the shift), kept off
#endif
// ---- RNG (shader.wgsl:104-115) --------------------------------------------------------------------
__device__ __forceinline__ uint32_t pcg_hash(uint32_t input) {
    uint32_t state = input * 747796405u + 2891336453u;
    uint32_t word = ((state >> ((state >> 28u) + 4u)) ^ state) * 277803737u;
#if TRT_V_PCG_MULHI
    // word >> 22 as the high half of word * 2^10: the same bits from the multiplier pipe instead of the ALU pipe
    uint32_t hi;
    asm("mul.hi.u32 %0, %1, 1024;" : "=r"(hi) : "r"(word));
    return hi ^ word;
#else
    return (word >> 22u) ^ word;
#endif
}
// f32(state) / 4294967295.0: the literal is 2^32 in f32, so this is an exact scaling by 2^-32
__device__ __forceinline__ float rnd(uint32_t &state) {
    state = pcg_hash(state);
    return __uint2float_rn(state) * 2.3283064365386963e-10f;
}

// ---- pinned sin/cos: the arithmetic convention for WGSL sin()/cos() --------------------
// Cody-Waite reduction by pi/2 in three fused steps, degree-7/8 minimax polynomials, all fused.
__device__ __forceinline__ void sincos_pinned(float phi, float &sn, float &cs) {
    float q = rintf(phi * 0.6366197466850281f);
    float r = __fmaf_rn(q, -1.5703125f, phi);
    r = __fmaf_rn(q, -0.0004837512969970703f, r);
    r = __fmaf_rn(q, -7.549790126404332e-08f, r);
    float s = r * r;
    float ps = __fmaf_rn(-0.00019517233886290342f, s, 0.008332177996635437f);
    ps = __fmaf_rn(ps, s, -0.16666655242443085f);
    float sr = __fmaf_rn(r * s, ps, r);
    float pc = __fmaf_rn(2.4438379114144482e-05f, s, -0.0013887366512790322f);
    pc = __fmaf_rn(pc, s, 0.04166664555668831f);
    pc = __fmaf_rn(pc, s, -0.5f);
    float cr = __fmaf_rn(pc, s, 1.0f);
    int k = __float2int_rn(q) & 3;
    float so = (k & 1) ? cr : sr;
    float co = (k & 1) ? sr : cr;
    if (k == 1 || k == 2) co = -co;
    if (k >= 2) so = -so;
    sn = so;
    cs = co;
}

// conv 2: MUFU.SIN / MUFU.COS on the angle folded into (-pi, pi] (their best range); r2 in [0, 1]
template <int CONV> __device__ __forceinline__ void sincos_turn(float r2, float &sn, float &cs) {
    if (CONV == 2) {
        const float phi = 6.28318530718f * (r2 > 0.5f ? r2 - 1.0f : r2);
        sn = __sinf(phi);
        cs = __cosf(phi);
    } else {
        sincos_pinned(6.28318530718f * r2, sn, cs);
    }
}

#define TRT_PI_F 3.14159265359f
#define TRT_TWO_PI_F 6.28318530718f
#define TRT_FOUR_PI_F 12.56637061436f

// ---- counters -----------------------------------------------------------------------------------------
enum {
    K_PIXELS, K_SAMPLES_BASE, K_SAMPLES_EXTRA, K_CLOSEST_SCANS, K_HITS, K_MISSES, K_SHADOW_SCANS,
    K_SHADOW_UNOCCLUDED, K_SHADOW_LIT, K_SHADOW_NEEDED, K_MIRROR, K_DIFFUSE, K_ROULETTE_TESTS,
    K_ROULETTE_KILLS, K_PRIM_TESTS, K_PRIM_TESTS_MIN, K_FLOPS_WRITTEN, K_FLOPS_MIN, K_RNG_DRAWS, K_COUNT
};
template <bool COUNT> struct Counters;
template <> struct Counters<false> {
    __device__ __forceinline__ void add(int, uint32_t) {}
    __device__ __forceinline__ void flops(uint32_t) {}          // written and min
    __device__ __forceinline__ void flush(unsigned long long *) {}
};
template <> struct Counters<true> {
    unsigned long long v[K_COUNT];
    __device__ Counters() { for (int i = 0; i < K_COUNT; i++) v[i] = 0; }
    __device__ __forceinline__ void add(int k, uint32_t n) { v[k] += n; }
    __device__ __forceinline__ void flops(uint32_t n) { v[K_FLOPS_WRITTEN] += n; v[K_FLOPS_MIN] += n; }
    __device__ void flush(unsigned long long *g) {
        for (int i = 0; i < K_COUNT; i++) {
            unsigned long long s = v[i];
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((threadIdx.x & 31) == 0 && s) atomicAdd(&g[i], s);
        }
    }
};

}  // namespace trt
