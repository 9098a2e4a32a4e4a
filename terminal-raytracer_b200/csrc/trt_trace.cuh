// trt_trace.cuh -- ray generation, linear-scan intersection, NEE, bounce, roulette: the per-ray
// device functions shared by the trace kernels.  Follows /root/reference/src/shader.wgsl (cited).
#pragma once
#include "trt_device.cuh"
#ifndef __CUDACC_RTC__
#include <cuda_pipeline.h>
#endif

#ifndef TRT_TAIL_DIAG
#define TRT_TAIL_DIAG 0  // measurement build (tools/tail_hist.py): 1 = histograms of when warps run dry / leave, 2 = last pixels
#endif
namespace trt {

struct LightView {               // emissive spheres / triangles in scan order (shader.wgsl:342-392)
    const float4 *la, *lb, *lc;  // sphere: la = (c, r); triangle: la = v0, lb = v1, lc = v2; lb.w = dot(n, v0), lc.w = magnitude scale
    const float4 *ln;            // (light normal xyz [triangle], area)
    const float4 *le;            // (emission rgb, type as float bits)
};

// ---- where the scan reads geometry from -----------------------------------------------------------
// Every lane of a warp tests the same primitive at the same time, so the geometry address is
// warp-uniform.  Small scenes (Cornell box: 11 primitives) live in __constant__ memory: operands come
// from the constant bank / uniform registers, no per-lane load, no load latency in the scan.
// Large scenes (20 k primitives) are read from the SoA arrays in HBM through the read-only path
// (one broadcast transaction per warp, L1/L2 resident: 0.85 MB).
#define TRT_CONST_LIGHTS 16
#define TRT_CONST_RUNS 16
struct ConstScene {
    float4 ga[TRT_CONST_PRIMS], gb[TRT_CONST_PRIMS], gc[TRT_CONST_PRIMS];
    float4 la[TRT_CONST_LIGHTS], lb[TRT_CONST_LIGHTS], lc[TRT_CONST_LIGHTS], ln[TRT_CONST_LIGHTS], le[TRT_CONST_LIGHTS];
    int4 runs[TRT_CONST_RUNS];
};
TRT_CONSTANT(ConstScene c_scene);      // one mirror per translation unit (strict / native builds)

// Canonical small scene (conv 2 fast scan): the reference always pushes spheres, then planes, then
// triangles (lib.rs:122-154), so a scene of at most TRT_SMALL_S/P/T of each is mirrored once more with
// one fixed constant-bank address per primitive: the unrolled scan reads its operands straight from
// the constant bank (no address arithmetic, no loop control, no run dispatch).  Index of sphere j is
// j, of plane j n_sph + j, of triangle j n_sph + n_pl + j.
#define TRT_SMALL_S 24         // capacity of the mirror: what a run-time specialised kernel (trt_jit.cuh) can take
#define TRT_SMALL_P 8
#define TRT_SMALL_T 8
#define TRT_SMALL_LPAIRS 8      // pairs of lights mirrored for the packed reach test (= TRT_NV_MAX_LIGHTS / 2)
#define TRT_STATIC_S 4         // slots the statically compiled run-time-count scan unrolls (ok: the scene fits these)
#define TRT_STATIC_P 8
#define TRT_STATIC_T 4
struct SmallScene {
    int ok, n_sph, n_pl, n_tri;       // ok: canonical order and within TRT_STATIC_*; jit_ok (below): within TRT_SMALL_*
    int n_lights, light_spheres, n_light_spheres, jit_ok;
    unsigned tri_zero, pad0_, pad1_, pad2_;               // ShapeFixed::TZ of this scene   // light list: count, bit k set = light k is a sphere
    float4 sph[TRT_SMALL_S];          // (c, r)
    float4 sph_r2[TRT_SMALL_S];       // (r*r, -, -, -)
    // spheres 2k and 2k + 1 side by side for the packed pre-reject of scan_small_native (TRT_V_PACKED):
    // [3k] = (x_a, x_b, y_a, y_b), [3k + 1] = (z_a, z_b, r_a, r_b), [3k + 2] = (r_a^2, r_b^2, -, -)
    float4 sph_pair[3 * (TRT_SMALL_S / 2)];
    float4 pl_p[TRT_SMALL_P];         // (point, kind as int bits): kind 0 / 1 / 2 = normal is +-unit x / y / z, 3 = general
    float4 pl_n[TRT_SMALL_P];         // raw normal (general planes)
    // planes 2k and 2k + 1 when both are axis-aligned on the SAME axis (the walls of a box): their coordinates on that
    // axis side by side, for one packed multiply-add (scan_small_native with a compile-time shape, TRT_V_PACKED)
    float2 pl_pair[TRT_SMALL_P / 2];
    // triangle lights 2k and 2k + 1 side by side (the reach test of the bounce, trt_native.cuh): [3k] = (nx_a, nx_b, ny_a,
    // ny_b), [3k + 1] = (nz_a, nz_b, off_a, off_b), [3k + 2] = (|off_a|, |off_b|, -, -); off = dot(n, v0)
    float4 lt_pair[3 * TRT_SMALL_LPAIRS];
    float4 tri_v0[TRT_SMALL_T], tri_e1[TRT_SMALL_T], tri_e2[TRT_SMALL_T];
};
TRT_CONSTANT(SmallScene c_small);

struct SrcGlobal {
    static constexpr bool kGrid = false;
    static constexpr bool kSmall = false;
    static constexpr bool kTiled = false;     // every warp reads the geometry on its own (L1 / L2)
    static constexpr int kTileSlots = 1;
    static constexpr int kCtasPerSm = 0;      // 0: the kernel's default
    static constexpr int kUnroll = 4;       // HBM/L2 latency: keep several primitives' loads in flight
    const SceneView &sc;
    const LightView &lv;
    __device__ SrcGlobal(const SceneView &s, const LightView &l) : sc(s), lv(l) {}
    __device__ __forceinline__ float4 ga(int i) const { return __ldg(&sc.ga[i]); }
    __device__ __forceinline__ float4 gb(int i) const { return __ldg(&sc.gb[i]); }
    __device__ __forceinline__ float4 gc(int i) const { return __ldg(&sc.gc[i]); }
    __device__ __forceinline__ float ga_axis(int i, int axis) const { return __ldg(reinterpret_cast<const float *>(&sc.ga[i]) + axis); }
    __device__ __forceinline__ int4 run(int r) const { return __ldg(&sc.runs[r]); }
    __device__ __forceinline__ float4 la(int k) const { return __ldg(&lv.la[k]); }
    __device__ __forceinline__ float4 lb(int k) const { return __ldg(&lv.lb[k]); }
    __device__ __forceinline__ float4 lc(int k) const { return __ldg(&lv.lc[k]); }
    __device__ __forceinline__ float4 ln(int k) const { return __ldg(&lv.ln[k]); }
    __device__ __forceinline__ float4 le(int k) const { return __ldg(&lv.le[k]); }
};
struct SrcConst {
    static constexpr bool kGrid = false;
    static constexpr bool kSmall = true;     // the canonical small-scene mirror may be used (c_small.ok)
    static constexpr bool kTiled = false;
    static constexpr int kTileSlots = 1;
    static constexpr int kCtasPerSm = 0;
    static constexpr int kUnroll = 1;       // constant bank: no load latency; keep the loop body small (I-cache)
    const SceneView &sc;
    __device__ SrcConst(const SceneView &s, const LightView &) : sc(s) {}
    __device__ __forceinline__ float4 ga(int i) const { return c_scene.ga[i]; }
    __device__ __forceinline__ float4 gb(int i) const { return c_scene.gb[i]; }
    __device__ __forceinline__ float4 gc(int i) const { return c_scene.gc[i]; }
    __device__ __forceinline__ float ga_axis(int i, int axis) const { return reinterpret_cast<const float *>(&c_scene.ga[i])[axis]; }
    __device__ __forceinline__ int4 run(int r) const { return c_scene.runs[r]; }
    __device__ __forceinline__ float4 la(int k) const { return c_scene.la[k]; }
    __device__ __forceinline__ float4 lb(int k) const { return c_scene.lb[k]; }
    __device__ __forceinline__ float4 lc(int k) const { return c_scene.lc[k]; }
    __device__ __forceinline__ float4 ln(int k) const { return c_scene.ln[k]; }
    __device__ __forceinline__ float4 le(int k) const { return c_scene.le[k]; }
};

// Compile-time run list of the strict production scan (scan_lean): (device run type, begin, end) per run, as
// trt_upload_scene builds them (trt_api.cu: type runs, axis-aligned planes 3..5, slab pairs 6..8).  With a fixed list
// the run loop, the type dispatch and the inner loops unroll and every operand has a fixed constant-bank address.
struct RunsAny {
    static constexpr bool kFixed = false;
    static constexpr int kCount = 0;
    static __host__ __device__ __forceinline__ constexpr int4 run(int) { return int4{0, 0, 0, 0}; }
};
// the default Cornell box: 3 spheres, slab pairs y (floor / ceiling), x (right / left), z (front / back), 2 triangles
struct RunsCornell {
    static constexpr bool kFixed = true;
    static constexpr int kCount = 5;
    static __host__ __device__ __forceinline__ constexpr int4 run(int r) {
        return r == 0 ? int4{0, 0, 3, 0} : r == 1 ? int4{7, 3, 5, 0} : r == 2 ? int4{6, 5, 7, 0} : r == 3 ? int4{8, 7, 9, 0} : int4{2, 9, 11, 0};
    }
};
// Compile-time description of a canonical small scene for scan_small_native: primitive counts, the kind of every
// plane (2 bits each, first plane in the low bits; kinds as in SmallScene::pl_p.w) and, when NL >= 0, the number of
// lights and which of them are spheres (bit k of LSPH; the others are triangles).
struct ShapeAny {
    typedef RunsAny Runs;
    static constexpr bool kFixed = false, kFixedLights = false;
    static constexpr int NS = 0, NP = 0, NT = 0, NL = -1;
    static constexpr unsigned KINDS = 0u, LSPH = 0u;
    static __host__ __device__ __forceinline__ constexpr int kind(int) { return 3; }
    static __host__ __device__ __forceinline__ constexpr bool light_is_sphere(int, int type) { return type == 0; }
    static __host__ __device__ __forceinline__ constexpr unsigned tri_zero(int) { return 0u; }
};
template <int NS_, int NP_, int NT_, unsigned KINDS_, int NL_ = -1, unsigned LSPH_ = 0u, unsigned TZ_ = 0u, class RUNS_ = RunsAny> struct ShapeFixed {
    typedef RUNS_ Runs;
    static constexpr bool kFixed = true, kFixedLights = NL_ >= 0;
    static constexpr int NS = NS_, NP = NP_, NT = NT_, NL = NL_;
    static constexpr unsigned KINDS = KINDS_, LSPH = LSPH_;
    static __host__ __device__ __forceinline__ constexpr int kind(int j) { return (int)((KINDS_ >> (2 * j)) & 3u); }
    static __host__ __device__ __forceinline__ constexpr bool light_is_sphere(int k, int type) {
        return NL_ < 0 ? type == 0 : LSPH_ == 0u ? false : LSPH_ == (1u << NL_) - 1u ? true : ((LSPH_ >> k) & 1u) != 0u;
    }
    // 6 bits per triangle (the first five): e1.x e1.y e1.z e2.x e2.y e2.z are exactly zero (axis-aligned edges)
    static constexpr unsigned TZ = TZ_;
    static __host__ __device__ __forceinline__ constexpr unsigned tri_zero(int j) { return j < 5 ? (TZ_ >> (6 * j)) & 63u : 0u; }
};
// the reference's embedded default scene (lib.rs:107, src/scenes/Cornell_Box.json): 3 spheres, floor / ceiling
// (y), right / left (x), front / back (z) walls, 2 triangles, both of them lights
typedef ShapeFixed<3, 6, 2, 0xA05u, 2, 0u, 0x4DEu> ShapeCornell;      // light quad: edges along x and z
typedef ShapeFixed<3, 6, 2, 0xA05u, 2, 0u, 0x4DEu, RunsCornell> ShapeCornellStrict;   // + the strict scan's run list

// Large scenes in the production kernel: the CTA scans together.  Each warp scanning 20 k primitives on its
// own streams ~1 MB through a 256 KB L1 per scan (C4 measured: 0.65 % of FP32 peak, L2 -> L1 bound); here the
// four warps of a CTA walk the scene in step, tile by tile: TRT_TILE primitives (3 x float4 each) are staged
// in shared memory with cp.async, double-buffered, and every lane tests its own ray against the tile through
// conflict-free broadcast reads.  Light data and hit records still come through the read-only path.
// Large scenes through the opt-in uniform grid (scan_grid): geometry through the read-only path, per-lane traversal.
struct SrcGrid : SrcGlobal {
    static constexpr bool kGrid = true;
    __device__ SrcGrid(const SceneView &s, const LightView &l) : SrcGlobal(s, l) {}
};
#ifndef TRT_TILED_CTAS
#define TRT_TILED_CTAS 6
#endif
struct SrcTiled : SrcGlobal {
    static constexpr bool kTiled = true;
    static constexpr int kTileSlots = 2 * 4 * TRT_TILE;      // two buffers of (ga, gb, gc, gd)[TRT_TILE]
    static constexpr int kCtasPerSm = TRT_TILED_CTAS;        // 6: 80 registers; the scan loop must not spill
    __device__ SrcTiled(const SceneView &s, const LightView &l) : SrcGlobal(s, l) {}
};
// geometry accessor over one staged tile: index i is the primitive's scan-order index
struct TileAcc {
    const float4 *a, *b, *c;
    int base;
    const float4 *d;     // triangle tiles: the pair-packed cull entries (SceneView::gd), tile-relative
    __device__ __forceinline__ float4 ga(int i) const { return a[i - base]; }
    __device__ __forceinline__ float4 gb(int i) const { return b[i - base]; }
    __device__ __forceinline__ float4 gc(int i) const { return c[i - base]; }
    __device__ __forceinline__ float ga_axis(int i, int axis) const { return reinterpret_cast<const float *>(&a[i - base])[axis]; }
};

// Products with a constant that is exactly zero, dropped at compile time (run-time specialised scan: the flags are
// template constants once the loop is unrolled).  Exact for finite operands: x * 0 = 0, y - 0 = y, and a chain of
// fused multiply-adds that starts at 0 equals the chain without that term.  a * b - c * d, constants b and d:
__device__ __forceinline__ float dpz(float a, float b, bool zb, float c, float d, bool zd) {
    if (zb && zd) return 0.0f;
    if (zd) return a * b;
    if (zb) return -(c * d);
    return __fmaf_rn(a, b, -(c * d));
}
// dot(a, b) = fma(a.z, b.z, fma(a.y, b.y, a.x * b.x)) over the terms that are not known to be zero
__device__ __forceinline__ float dotz(float3 a, float3 b, bool z0, bool z1, bool z2) {
    float r = 0.0f;
    bool have = false;
    if (!z0) { r = a.x * b.x; have = true; }
    if (!z1) { r = have ? __fmaf_rn(a.y, b.y, r) : a.y * b.y; have = true; }
    if (!z2) { r = have ? __fmaf_rn(a.z, b.z, r) : a.z * b.z; }
    return r;
}

// guarded kernel variants (A/B through TRT_B200_JIT_DEFS or the Makefile's VARIANTS)
#ifndef TRT_V_CAMFOLD
#define TRT_V_CAMFOLD 1
#endif
// TRT_V_SHIFT_T: plane / triangle accept tests of the unrolled small-scene scan on t' = t - t_min compared as unsigned
// integers (one compare instead of two and a predicate merge).  t' carries one more rounding than t (within WGSL's
// latitude for the division it replaces).  B200 A/B frame: 39.94 -> 38.83 ms; the deviation from the strict frame is
// unchanged (C3 parity subset: 1726 of 20 000 pixels over 1e-4 with and without it; profiles/r2_ab_variants.txt).
#ifndef TRT_V_SHIFT_T
#define TRT_V_SHIFT_T 1
#endif
// TRT_V_KEY: the bound of the unrolled small-scene scan as ONE ordered key -- the bits of t' = t - t_min with the low
// index bits replaced by the primitive index.  For t' >= +0 the float order is the unsigned order, a negative or NaN t'
// sorts above every valid bound, so `t_min <= t < closest, remember who` is FFMA + LOP3 + (half a) VIMNMX3.U32 per plane
// (as written: FFMA + 2 FSETP + PLOP3 + FSEL + SEL).  Equal keys keep the lower index = the first primitive in scan
// order (shader.wgsl:279).  Cost: t is known to 2^-19 relative (4 index bits) instead of 2^-24 -- conv 2 only,
// inside its tolerance (tests/test_gpu_native.py).  B200, A/B frame: 39.94 -> 38.65 ms, i.e. 0.5 % better than
// TRT_V_SHIFT_T -- but the coarser t DOUBLES the deviation from the strict frame on deep paths (C3 parity subset: 3893
// of 20 000 pixels over 1e-4 against 1726, RMSE 0.34 of the noise floor against 0.23; profiles/r2_ab_variants.txt):
// measured, not adopted.
#ifndef TRT_V_KEY
#define TRT_V_KEY 0
#endif
// TRT_V_NO_RENORM: cosine_hemisphere_unit returns u*x + v*y + w*z without the final normalize (:334): (u, v, w) is
// orthonormal and x^2 + y^2 + z^2 = 1 up to rounding, so the length is 1 +- a few ulp already.  -0.6 % (38.65 -> 38.41 ms).
// TRT_V_TMA_TILES: the geometry tiles of the CTA-cooperative scan (SrcTiled) are staged by the TMA unit -- one thread
// issues three cp.async.bulk copies per tile (the SoA arrays ga / gb / gc are contiguous per tile) that complete on an
// mbarrier -- instead of three cp.async per thread.  0 = the per-thread cp.async staging of round 1 (A/B).
#ifndef TRT_V_TMA_TILES
#define TRT_V_TMA_TILES 1
#endif
#ifndef TRT_V_NO_RENORM
#define TRT_V_NO_RENORM 1
#endif
// TRT_V_TRI_CULL: triangle runs of 8 or more in the run-loop scans (scan_run: SrcTiled / SrcGlobal / unspecialised
// scenes, conv 2) test the ray against a sphere around v0 first.  0 = every triangle takes the full test (round 1, A/B).
#ifndef TRT_V_TRI_CULL
#define TRT_V_TRI_CULL 1
#endif

template <int CONV, bool COUNT, class SRC>
struct Tracer {
    typedef M<CONV> V;
    const SceneView &sc;
    const FrameParams &fp;
    Counters<COUNT> &cn;
    SRC src;
    uint32_t n_samples = 0, n_bounces = 0;      // always-on frame statistics (trt_frame_stats)

    __device__ Tracer(const SceneView &s, const LightView &l, const FrameParams &f, Counters<COUNT> &c)
        : sc(s), fp(f), cn(c), src(s, l) {}

    // warp-reduce the per-thread statistics and add them to the frame totals (all 32 lanes call this)
    __device__ __forceinline__ void flush_stats() {
        unsigned long long s = n_samples, b = n_bounces;
        for (int o = 16; o > 0; o >>= 1) {
            s += __shfl_xor_sync(0xffffffffu, s, o);
            b += __shfl_xor_sync(0xffffffffu, b, o);
        }
        if ((threadIdx.x & 31) == 0 && fp.stats) {
            atomicAdd(&fp.stats[0], s);
            atomicAdd(&fp.stats[1], b);
        }
    }

    // ---- single-primitive tests; `closest` is the current upper bound (shader.wgsl:258-277) -------
    // sphere: pre-reject shader.wgsl:260-269, then hit_sphere shader.wgsl:155-175 (oc == to_center,
    // h == t_ca bit for bit, so they are computed once)
    __device__ __forceinline__ float test_sphere(float3 ro, float3 rd, float4 a, float r2, float t_min, float closest,
                                                 uint32_t &cost) const {
        float3 tc = V::sub(xyz(a), ro);
        float t_ca = V::dot(tc, rd);
        float tcc = V::dot(tc, tc);
        float dist2 = V::nmadd(tcc, t_ca, t_ca);
        cost = 18;
        if (dist2 > r2 || t_ca + a.w < t_min || t_ca - a.w > closest) return -1.0f;
        float c = tcc - r2;
        float disc = V::msub(t_ca, t_ca, c);
        cost = 37;
        if (disc < 0.0f) return -1.0f;
        float sq = sqrtf(disc);
        float root = t_ca - sq;
        if (root <= t_min || closest <= root) {
            cost = 38;
            root = t_ca + sq;
            if (root <= t_min || closest <= root) return -1.0f;
        }
        return root;
    }
    // test_sphere from its pre-reject on, given t_ca = tc . d, |tc|^2 and dist2 (the packed pair path of scan_small_native)
    // (far = t_ca + r, near = t_ca - r)
    __device__ __forceinline__ float sphere_tail(float t_ca, float tcc, float dist2, float far, float near, float r2, float t_min, float closest) const {
        if (dist2 > r2 || far < t_min || near > closest) return -1.0f;
        const float c = tcc - r2;
        const float disc = V::msub(t_ca, t_ca, c);
        if (disc < 0.0f) return -1.0f;
        const float sq = sqrtf(disc);
        float root = t_ca - sq;
        if (root <= t_min || closest <= root) {
            root = t_ca + sq;
            if (root <= t_min || closest <= root) return -1.0f;
        }
        return root;
    }
    // shader.wgsl:178-189
    __device__ __forceinline__ float test_plane(float3 ro, float3 rd, float4 p, float4 n, float t_min, float closest,
                                                uint32_t &cost) const {
        float denom = V::dot(xyz(n), rd);
        cost = 5;
        if (fabsf(denom) < 0.0001f) return -1.0f;
        cost = 14;
        float t = V::dot(V::sub(xyz(p), ro), xyz(n)) / denom;
        if (t < t_min || t > closest) return -1.0f;
        return t;
    }
    // shader.wgsl:192-223 with edge1/edge2 precomputed (same f32 subtraction)
    __device__ __forceinline__ float test_triangle(float3 ro, float3 rd, float4 v0, float4 e1, float4 e2, float t_min,
                                                   float closest, uint32_t &cost) const {
        float3 h = V::cross(rd, xyz(e2));
        float a = V::dot(xyz(e1), h);
        cost = 20;
        if (a > -0.00001f && a < 0.00001f) return -1.0f;
        float f = 1.0f / a;
        float3 s = V::sub(ro, xyz(v0));
        float u = f * V::dot(s, h);
        cost = 30;
        if (u < 0.0f || u > 1.0f) return -1.0f;
        float3 q = V::cross(s, xyz(e1));
        float v = f * V::dot(rd, q);
        cost = 46;
        if (v < 0.0f || u + v > 1.0f) return -1.0f;
        cost = 52;
        float t = f * V::dot(xyz(e2), q);
        if (t > t_min && t < closest) return t;
        return -1.0f;
    }

    __device__ __forceinline__ float test_any(int type, int i, float3 ro, float3 rd, float t_min, float closest,
                                              uint32_t &cost) const {
        cost = 0;
        if (type == 0) return test_sphere(ro, rd, src.ga(i), src.gb(i).x, t_min, closest, cost);
        if (type == 1 || (type >= 3 && type <= 8)) return test_plane(ro, rd, src.ga(i), src.gb(i), t_min, closest, cost);   // 3..8: axis-aligned planes / slab pairs
        if (type == 2) return test_triangle(ro, rd, src.ga(i), src.gb(i), src.gc(i), t_min, closest, cost);
        return -1.0f;
    }

    // ---- hit_scene as written (shader.wgsl:248-294): full scan, narrowing bound, first index wins
    // ties.  The record is completed once from the final (t, index) -- same values as updating it on
    // every improvement.  shadow_needed: -1 closest-hit scan; 0/1 shadow scan (counting only).
    __device__ __forceinline__ float scan_closest(float3 ro, float3 rd, float t_min, float t_max, int &best,
                                                  int shadow_needed) {
        float closest = t_max;
        best = -1;
        bool found = false;
        for (int r = 0; r < sc.n_runs; r++) {
            const int4 run = src.run(r);
#pragma unroll 1                        // one copy of each test: the loop body must stay inside the instruction cache
            for (int i = run.y; i < run.z; i++) {
                uint32_t cost;
                const float t = test_any(run.x, i, ro, rd, t_min, closest, cost);
                if (COUNT) {
                    cn.add(K_PRIM_TESTS, 1);
                    cn.add(K_FLOPS_WRITTEN, cost);
                    if (shadow_needed < 0 || (shadow_needed == 1 && !found)) {
                        cn.add(K_PRIM_TESTS_MIN, 1);
                        cn.add(K_FLOPS_MIN, cost);
                    }
                }
                if (t > 0.0f && t < closest) {
                    closest = t;
                    best = i;
                    if (COUNT) {
                        // improving-hit cost needs the facing of this candidate
                        float3 p = V::axpy(ro, rd, t);
                        float3 n = normal_at(i, run.x, p);
                        bool front = V::dot(rd, n) < 0.0f;
                        uint32_t hc = 6u + (run.x == 0 ? 15u : run.x == 2 ? 24u : 9u) + 5u + (front ? 0u : 3u);
                        cn.add(K_FLOPS_WRITTEN, hc);
                        if (shadow_needed < 0) cn.add(K_FLOPS_MIN, hc);
                        found = true;
                    }
                }
            }
        }
        return best >= 0 ? closest : -1.0f;
    }

    // ---- production scan: same result as scan_closest(ro, rd, 0.001, t_max), fewer instructions ----------
    // Bit-exact shortcuts (DESIGN.md "Scan"), each guarded so that every input takes either the shortcut
    // or the as-written code above:
    //  * fused accept: a test's own range check and hit_scene's `t > 0 && t < closest` (:279) collapse
    //    into one condition per primitive type (t_min = 0.001 > 0 at every call site, :400,:445);
    //  * axis-aligned planes (normal = +-unit axis, found at upload): dot(n, d) == +-d_axis and
    //    dot(p - o, n) == +-(p_axis - o_axis) exactly for finite rays, so t = (p_a - o_a) / d_a;
    //  * two adjacent parallel axis-aligned planes ("slab pair", the walls of a box): one division for both;
    //  * triangles: u = (1/a) * dot(s, h) is clearly < 0 or > 1 from the signs / magnitudes of dot(s, h)
    //    and a (margins cover the two roundings); only the survivors pay 1/a.
    __device__ __noinline__ float scan_as_written(float3 ro, float3 rd, float t_max, int &best) {
        return scan_closest(ro, rd, 0.001f, t_max, best, -1);
    }

    // the tests of one run of scan_lean; `fallback`: an input the shortcuts do not cover (the caller rescans as written)
    template <int UN>
    __device__ __forceinline__ void lean_run(const int4 run, float3 ro, float3 rd, float t_min, float &closest, int &best, bool &fallback) {
            if (run.x == 0) {                                   // spheres: as written, accept == (t > 0)
#pragma unroll UN
                for (int i = run.y; i < run.z; i++) {
                    uint32_t cost;
                    const float t = test_sphere(ro, rd, src.ga(i), src.gb(i).x, t_min, closest, cost);
                    if (t > 0.0f) { closest = t; best = i; }
                }
            } else if (run.x >= 6 && run.x <= 8) {              // slab pairs: two parallel axis-aligned planes
                // Of two planes p_lo < p_hi on one axis a ray can only hit the first one past its origin:
                // d > 0: the smallest p above o; d < 0: the largest p below o.  The other plane has a larger
                // t (or t <= 0), so as written it either loses to this one or is rejected: one division
                // per pair.  Exceptions take the as-written scan: 0 < t < t_min on the chosen plane (the
                // other one may then be the hit) and |num| so large that both quotients could round alike.
                const float o_a = run.x == 6 ? ro.x : run.x == 7 ? ro.y : ro.z;
                const float d_a = run.x == 6 ? rd.x : run.x == 7 ? rd.y : rd.z;
                const bool live = fabsf(d_a) >= 0.0001f;        //  abs(denom) < 0.0001 -> miss (:180)
#pragma unroll UN
                for (int i = run.y; i < run.z; i += 2) {
                    const float4 pr = src.gc(i);                // (p_lo, p_hi, 1 if the first plane is p_hi, 2^20 * gap)
                    const bool take_hi = d_a > 0.0f ? !(o_a < pr.x) : (o_a > pr.y);
                    const float num = (take_hi ? pr.y : pr.x) - o_a;
                    const float t = num / d_a;
                    if (live) {
                        if ((t > 0.0f && t < t_min) || !(fabsf(num) < pr.w)) { fallback = true; return; }
                        if (t >= t_min && t < closest) { closest = t; best = i + ((take_hi ? 1 : 0) ^ __float2int_rz(pr.z)); }
                    }
                }
            } else if (run.x >= 3 && run.x <= 5) {              // single axis-aligned planes
                const float o_a = run.x == 3 ? ro.x : run.x == 4 ? ro.y : ro.z;
                const float d_a = run.x == 3 ? rd.x : run.x == 4 ? rd.y : rd.z;
                const bool live = fabsf(d_a) >= 0.0001f;
#pragma unroll UN
                for (int i = run.y; i < run.z; i++) {
                    const float t = (src.ga_axis(i, run.x - 3) - o_a) / d_a;
                    if (live && t >= t_min && t < closest) { closest = t; best = i; }
                }
            } else if (run.x == 1) {                            // general planes: as written, fused accept
#pragma unroll UN
                for (int i = run.y; i < run.z; i++) {
                    const float4 n = src.gb(i);
                    const float denom = V::dot(xyz(n), rd);
                    const float t = V::dot(V::sub(xyz(src.ga(i)), ro), xyz(n)) / denom;
                    if (fabsf(denom) >= 0.0001f && t >= t_min && t < closest) { closest = t; best = i; }
                }
            } else if (run.x == 2) {                            // triangles
#pragma unroll UN
                for (int i = run.y; i < run.z; i++) {
                    const float3 e1 = xyz(src.gb(i)), e2 = xyz(src.gc(i));
                    const float3 h = V::cross(rd, e2);
                    const float a = V::dot(e1, h);
                    if (a > -0.00001f && a < 0.00001f) continue;
                    const float3 s = V::sub(ro, xyz(src.ga(i)));
                    const float du = V::dot(s, h);
                    const float aa = fabsf(a), adu = fabsf(du);
                    // u = RN(RN(1/a) * du): negative (and not flushed to zero) or above 1 beyond the rounding margin
                    const bool opposite = (__float_as_int(du) ^ __float_as_int(a)) < 0;
                    if ((opposite && adu > 1e-30f && aa < 1e6f) || adu > aa * 1.000001f) continue;
                    const float f = 1.0f / a;
                    const float u = f * du;
                    if (u < 0.0f || u > 1.0f) continue;
                    const float3 q = V::cross(s, e1);
                    const float v = f * V::dot(rd, q);
                    if (v < 0.0f || u + v > 1.0f) continue;
                    const float t = f * V::dot(e2, q);
                    if (t > t_min && t < closest) { closest = t; best = i; }
                }
            }
    }

    template <class RUNS = RunsAny>
    __device__ __forceinline__ float scan_lean(float3 ro, float3 rd, float t_max, int &best) {
        const float t_min = 0.001f;
        // non-finite rays (never produced by a sane scene) take the as-written code: 0 * inf is NaN there
        // (origins beyond 1e30 too: there p - o could overflow; x * 1e8 overflows to inf first)
        float fin = __fmaf_rn(ro.x * 1e8f, 0.0f, 0.0f);
        fin = __fmaf_rn(ro.y * 1e8f, 0.0f, fin); fin = __fmaf_rn(ro.z * 1e8f, 0.0f, fin);
        fin = __fmaf_rn(rd.x, 0.0f, fin); fin = __fmaf_rn(rd.y, 0.0f, fin); fin = __fmaf_rn(rd.z, 0.0f, fin);
        if (!(fin == 0.0f)) return scan_as_written(ro, rd, t_max, best);
        float closest = t_max;
        best = -1;
        bool fallback = false;
        if (RUNS::kFixed) {
#pragma unroll
            for (int r = 0; r < RUNS::kCount; r++) lean_run<16>(RUNS::run(r), ro, rd, t_min, closest, best, fallback);
        } else {
            for (int r = 0; r < sc.n_runs && !fallback; r++) lean_run<SRC::kUnroll>(src.run(r), ro, rd, t_min, closest, best, fallback);
        }
        if (fallback) return scan_as_written(ro, rd, t_max, best);
        return best >= 0 ? closest : -1.0f;
    }

    // ---- conv 2 scan: hit_scene (shader.wgsl:248-294) in the device-native convention ---------------------
    // Same tests, same order, same strict `t < closest` accept (first index wins ties); divisions are
    // MUFU.RCP products and the compiler contracts.  One reciprocal per ray axis serves every axis-
    // aligned plane; of a slab pair only the plane ahead of the origin is evaluated (the other one has
    // t <= 0 or a larger t), except when that plane is closer than t_min: then the second plane is
    // tested too, as hit_scene would.
    // ---- the cull of scan_run's long triangle runs (TRT_V_TRI_CULL); returns the first index it has not handled ----
    static __device__ __forceinline__ bool tri_far_away(const float4 v0r, float3 ro, float3 rd) {
        const float3 tc = V::sub(xyz(v0r), ro);
        const float t_ca = V::dot(tc, rd);
        const float tcc = V::dot(tc, tc);
        return V::nmadd(tcc, t_ca, t_ca) > __fmaf_rn(2e-5f, tcc, v0r.w);
    }
    template <class ACC, class FULL>
    __device__ __forceinline__ int tri_cull(const ACC &acc, const int4 run, float3 ro, float3 rd, FULL &full) {
        int i = run.y;
        // four at a time: the four loads in flight together, four independent chains, then the branches
        for (; i + 4 <= run.z; i += 4) {
            const float4 w0 = acc.ga(i), w1 = acc.ga(i + 1), w2 = acc.ga(i + 2), w3 = acc.ga(i + 3);
            const bool f0 = tri_far_away(w0, ro, rd), f1 = tri_far_away(w1, ro, rd), f2 = tri_far_away(w2, ro, rd), f3_ = tri_far_away(w3, ro, rd);
            if (!f0) full(i);
            if (!f1) full(i + 1);
            if (!f2) full(i + 2);
            if (!f3_) full(i + 3);
        }
        for (; i < run.z; i++)
            if (!tri_far_away(acc.ga(i), ro, rd)) full(i);
        return i;
    }
    // A staged tile: TWO triangles per step on packed FP32 pairs (f2_*: FADD2 / FMUL2 / FFMA2).  The same inequality,
    // |tc|^2 - (tc . d)^2 > 2e-5 |tc|^2 + R, arranged as |tc|^2 * (1 - 2e-5) + (-R) > (tc . d)^2 so that it needs no negation:
    // 11 packed instructions and two compares for two triangles (scalar: 12 + 1 each).
    template <class FULL>
    __device__ __forceinline__ int tri_cull(const TileAcc &acc, const int4 run, float3 ro, float3 rd, FULL &full) {
        const f2_t nrx = f2_pack(-ro.x, -ro.x), nry = f2_pack(-ro.y, -ro.y), nrz = f2_pack(-ro.z, -ro.z);
        const f2_t dx = f2_pack(rd.x, rd.x), dy = f2_pack(rd.y, rd.y), dz = f2_pack(rd.z, rd.z);
        const f2_t k2 = f2_pack(1.0f - 2e-5f, 1.0f - 2e-5f);
        const ulonglong2 *d = reinterpret_cast<const ulonglong2 *>(acc.d);
        const int n_pairs = (run.z - run.y) >> 1;
        // (lhs, rhs) of the inequality for the two primitives of pair p: far away <=> lhs > rhs
        auto pair_test = [&](int p, float &l0, float &l1, float &r0, float &r1) {
            const ulonglong2 q0 = d[2 * p], q1 = d[2 * p + 1];
            const f2_t tcx = f2_add(q0.x, nrx), tcy = f2_add(q0.y, nry), tcz = f2_add(q1.x, nrz);
            const f2_t tca = f2_fma(tcz, dz, f2_fma(tcy, dy, f2_mul(tcx, dx)));
            const f2_t tcc = f2_fma(tcz, tcz, f2_fma(tcy, tcy, f2_mul(tcx, tcx)));
            f2_unpack(f2_fma(tcc, k2, q1.y), l0, l1);
            f2_unpack(f2_mul(tca, tca), r0, r1);
        };
        int p = 0;
        // two pairs per round, ONE branch for the four primitives in the common case (all four far away for all lanes)
        for (; p + 2 <= n_pairs; p += 2) {
            float a0, a1, b0, b1, c0, c1, e0, e1;
            pair_test(p, a0, a1, b0, b1);
            pair_test(p + 1, c0, c1, e0, e1);
            if (!(a0 > b0 && a1 > b1 && c0 > e0 && c1 > e1)) {
                if (!(a0 > b0)) full(run.y + 2 * p);
                if (!(a1 > b1)) full(run.y + 2 * p + 1);
                if (!(c0 > e0)) full(run.y + 2 * p + 2);
                if (!(c1 > e1)) full(run.y + 2 * p + 3);
            }
        }
        for (; p < n_pairs; p++) {
            float l0, l1, r0, r1;
            pair_test(p, l0, l1, r0, r1);
            if (!(l0 > r0 && l1 > r1)) {
                if (!(l0 > r0)) full(run.y + 2 * p);
                if (!(l1 > r1)) full(run.y + 2 * p + 1);
            }
        }
        int i = run.y + 2 * n_pairs;
        if (run.x == 2)
            for (; i < run.z; i++)
                if (!tri_far_away(acc.ga(i), ro, rd)) full(i);
        return i;
    }
    // spheres: only a staged tile has cull entries for them (an odd last sphere is left to the caller's loop)
    template <class ACC, class FULL>
    __device__ __forceinline__ int sph_cull(const ACC &, const int4 run, float3, float3, FULL &) { return run.y; }
    template <class FULL>
    __device__ __forceinline__ int sph_cull(const TileAcc &acc, const int4 run, float3 ro, float3 rd, FULL &full) { return tri_cull(acc, run, ro, rd, full); }

    // the tests of one run [begin, end) of primitives of one type, geometry through `acc` (global / constant /
    // staged tile); `closest` and `best` carry the scan state (shader.wgsl:255-292)
    template <int UNROLL, class ACC>
    __device__ __forceinline__ void scan_run(const ACC &acc, const int4 run, float3 ro, float3 rd, float3 inv, float t_min,
                                             float &closest, int &best) {
            if (run.x == 0) {
                auto full = [&](int i) {
                    uint32_t cost;
                    const float t = test_sphere(ro, rd, acc.ga(i), acc.gb(i).x, t_min, closest, cost);
                    if (t > 0.0f) { closest = t; best = i; }
                };
                int i = run.y;
#if TRT_V_TRI_CULL
                // a staged tile: the packed-pair cull (centre and r^2 in gd) in front of the sphere test as well
                if (run.z - run.y >= 8) i = sph_cull(acc, run, ro, rd, full);
#endif
#pragma unroll UNROLL
                for (; i < run.z; i++) full(i);
            } else if (run.x >= 3 && run.x <= 8) {
                const int axis = run.x >= 6 ? run.x - 6 : run.x - 3;
                const float o_a = axis == 0 ? ro.x : axis == 1 ? ro.y : ro.z;
                const float d_a = axis == 0 ? rd.x : axis == 1 ? rd.y : rd.z;
                const float i_a = axis == 0 ? inv.x : axis == 1 ? inv.y : inv.z;
                const bool live = fabsf(d_a) >= 0.0001f;
                if (run.x >= 6) {
#pragma unroll UNROLL
                    for (int i = run.y; i < run.z; i += 2) {
                        const float4 pr = acc.gc(i);            // (p_lo, p_hi, 1 if the first plane is p_hi, -)
                        const bool take_hi = d_a > 0.0f ? !(o_a < pr.x) : (o_a > pr.y);
                        const float t = ((take_hi ? pr.y : pr.x) - o_a) * i_a;
                        const int first_hi = __float2int_rz(pr.z);
                        if (live && t >= t_min && t < closest) { closest = t; best = i + ((take_hi ? 1 : 0) ^ first_hi); }
                        if (live && t > 0.0f && t < t_min) {    // rare: the other plane of the pair may be the hit
                            const float t2 = ((take_hi ? pr.x : pr.y) - o_a) * i_a;
                            if (t2 >= t_min && t2 < closest) { closest = t2; best = i + ((take_hi ? 0 : 1) ^ first_hi); }
                        }
                    }
                } else {
#pragma unroll UNROLL
                    for (int i = run.y; i < run.z; i++) {
                        const float t = (acc.ga_axis(i, axis) - o_a) * i_a;
                        if (live && t >= t_min && t < closest) { closest = t; best = i; }
                    }
                }
            } else if (run.x == 1) {
#pragma unroll UNROLL
                for (int i = run.y; i < run.z; i++) {
                    const float4 n = acc.gb(i);
                    const float denom = V::dot(xyz(n), rd);
                    const float t = V::dot(V::sub(xyz(acc.ga(i)), ro), xyz(n)) * rcp_native(denom);
                    if (fabsf(denom) >= 0.0001f && t >= t_min && t < closest) { closest = t; best = i; }
                }
            } else if (run.x == 2) {
                // hit_triangle (shader.wgsl:192-223) for one triangle, as written in this convention
                auto full = [&](int i) {
                    const float3 e1 = xyz(acc.gb(i)), e2 = xyz(acc.gc(i));
                    const float3 h = V::cross(rd, e2);
                    const float a = V::dot(e1, h);
                    if (a > -0.00001f && a < 0.00001f) return;
                    const float f = rcp_native(a);
                    const float3 s = V::sub(ro, xyz(acc.ga(i)));
                    const float u = f * V::dot(s, h);
                    if (u < 0.0f || u > 1.0f) return;
                    const float3 q = V::cross(s, e1);
                    const float v = f * V::dot(rd, q);
                    if (v < 0.0f || u + v > 1.0f) return;
                    const float t = f * V::dot(e2, q);
                    if (t > t_min && t < closest) { closest = t; best = i; }
                };
                int i = run.y;
#if TRT_V_TRI_CULL
                // Long runs (the tiled / global scans of a large scene): a conservative cull in front of the test.  A ray
                // that hits the triangle passes within r = max(|e1|, |e2|) of v0, so when its LINE stays further away
                // than that the as-written test would reject it too (u or v outside [0, 1]).  One 16-byte load (v0 and
                // r^2 in ga.w, written by repack_kernel with 10 % to spare) and 12 arithmetic instructions instead of
                // three loads and ~25; d2 = |tc|^2 - (tc . d)^2 is as exact as |tc|^2 * 1e-6 (|d| = 1 to a few ulp,
                // cancellation), hence the second term.  Survivors run the test as written: accepted hits are unchanged.
                if (run.z - run.y >= 8) i = tri_cull(acc, run, ro, rd, full);
#endif
#pragma unroll UNROLL
                for (; i < run.z; i++) full(i);
            }
    }

    __device__ __forceinline__ float scan_native(float3 ro, float3 rd, float t_max, int &best) {
        const float t_min = 0.001f;
        float closest = t_max;
        best = -1;
        const float3 inv = f3(rcp_native(rd.x), rcp_native(rd.y), rcp_native(rd.z));
        for (int r = 0; r < sc.n_runs; r++) scan_run<SRC::kUnroll>(src, src.run(r), ro, rd, inv, t_min, closest, best);
        return best >= 0 ? closest : -1.0f;
    }

    // ---- opt-in: hit_scene through the uniform grid (lib.rs:181-244 built one and never used it) ----------------
    // 3-D DDA over the cells the ray crosses; a cell's primitives are tested with the same functions as the linear
    // scan, a hit counts once its t lies inside the cell being visited (a primitive spans several cells), and the walk
    // stops when the best hit lies before the next cell.  Candidates are compared as (t, index) pairs, lowest index on
    // equal t, so the result is the linear scan's `first primitive wins ties` (shader.wgsl:279) -- but this is an
    // acceleration structure the reference does not execute: outside the parity contract, opt-in, reported separately.
    __device__ __forceinline__ float scan_grid(float3 ro, float3 rd, float t_max, int &best) {
        const float t_min = 0.001f;
        float closest = t_max;
        best = -1;
        auto consider = [&](int i, int type) {
            uint32_t cost;
            const float t = test_any(type, i, ro, rd, t_min, type == 0 ? closest : closest * 1.0000005f + 1e-30f, cost);
            if (t > 0.0f && (t < closest || (t == closest && i < best))) { closest = t; best = i; }
        };
        for (int k = 0; k < sc.n_grid_planes; k++) consider(__ldg(&sc.grid_planes[k]), 1);
        if (!(t_max > t_min)) return -1.0f;
        // clip the ray to the grid's box
        const float3 inv = f3(rcp_native(rd.x), rcp_native(rd.y), rcp_native(rd.z));
        const float3 hi = f3(sc.grid_min.x + sc.grid_cell.x * sc.gx, sc.grid_min.y + sc.grid_cell.y * sc.gy, sc.grid_min.z + sc.grid_cell.z * sc.gz);
        float t0 = t_min, t1 = closest;
        {
            const float ax = (sc.grid_min.x - ro.x) * inv.x, bx = (hi.x - ro.x) * inv.x;
            const float ay = (sc.grid_min.y - ro.y) * inv.y, by = (hi.y - ro.y) * inv.y;
            const float az = (sc.grid_min.z - ro.z) * inv.z, bz = (hi.z - ro.z) * inv.z;
            t0 = fmaxf(t0, fmaxf(fminf(ax, bx), fmaxf(fminf(ay, by), fminf(az, bz))));
            t1 = fminf(t1, fminf(fmaxf(ax, bx), fminf(fmaxf(ay, by), fmaxf(az, bz))));
        }
        if (!(t0 <= t1)) return best >= 0 ? closest : -1.0f;
        const float3 p = f3(ro.x + rd.x * t0, ro.y + rd.y * t0, ro.z + rd.z * t0);
        int cx = min(max((int)floorf((p.x - sc.grid_min.x) * sc.grid_inv_cell.x), 0), sc.gx - 1);
        int cy = min(max((int)floorf((p.y - sc.grid_min.y) * sc.grid_inv_cell.y), 0), sc.gy - 1);
        int cz = min(max((int)floorf((p.z - sc.grid_min.z) * sc.grid_inv_cell.z), 0), sc.gz - 1);
        const int sx = rd.x > 0.0f ? 1 : -1, sy = rd.y > 0.0f ? 1 : -1, sz = rd.z > 0.0f ? 1 : -1;
        const float big = 3.0e38f;
        // t at which the ray leaves the current cell along each axis, and the t per cell step
        float nx = fabsf(rd.x) > 1e-20f ? (sc.grid_min.x + sc.grid_cell.x * (cx + (sx > 0 ? 1 : 0)) - ro.x) * inv.x : big;
        float ny = fabsf(rd.y) > 1e-20f ? (sc.grid_min.y + sc.grid_cell.y * (cy + (sy > 0 ? 1 : 0)) - ro.y) * inv.y : big;
        float nz = fabsf(rd.z) > 1e-20f ? (sc.grid_min.z + sc.grid_cell.z * (cz + (sz > 0 ? 1 : 0)) - ro.z) * inv.z : big;
        const float dx = fabsf(rd.x) > 1e-20f ? fabsf(sc.grid_cell.x * inv.x) : big;
        const float dy = fabsf(rd.y) > 1e-20f ? fabsf(sc.grid_cell.y * inv.y) : big;
        const float dz = fabsf(rd.z) > 1e-20f ? fabsf(sc.grid_cell.z * inv.z) : big;
        for (;;) {
            const float t_exit = fminf(nx, fminf(ny, nz));
            const int cell = cx + sc.gx * (cy + sc.gy * cz);
            const uint32_t b = __ldg(&sc.grid_off[cell]), e = __ldg(&sc.grid_off[cell + 1]);
            for (uint32_t k = b; k < e; k++) {
                const int i = (int)__ldg(&sc.grid_idx[k]);
                consider(i, (int)__ldg(&sc.type[i]));
            }
            // a hit inside the cells visited so far ends the walk (margin: the cell boundary's own rounding)
            if (best >= 0 && closest <= t_exit * 1.000001f + 1e-6f) break;
            if (t_exit >= t1) break;
            if (nx <= ny && nx <= nz) { cx += sx; nx += dx; if (cx < 0 || cx >= sc.gx) break; }
            else if (ny <= nz) { cy += sy; ny += dy; if (cy < 0 || cy >= sc.gy) break; }
            else { cz += sz; nz += dz; if (cz < 0 || cz >= sc.gz) break; }
        }
        return best >= 0 ? closest : -1.0f;
    }

#ifndef __CUDACC_RTC__
    // CTA-cooperative form of scan_native (SrcTiled): every thread of the CTA must call it, the same number of
    // times.  `tile` = 2 x 4 x TRT_TILE float4 of shared memory.  any_hit: the caller only reads t < 0
    // (shader.wgsl:400-402), so a warp whose lanes have all found an occluder (or have nothing to look for)
    // stops testing; it keeps taking part in the staging and the barriers.
    // `mbar`: two mbarriers in shared memory (initialised once by the kernel, count 1), `phase`: the parity each of them
    // completes next (bit b for buffer b), carried by the caller across scans (TRT_V_TMA_TILES)
    __device__ __forceinline__ float scan_native_tiled(float3 ro, float3 rd, float t_max, int &best, float4 *tile, bool any_hit,
                                                       unsigned long long *mbar, unsigned &phase) {
        const float t_min = 0.001f;
        float closest = t_max;
        best = -1;
        const float3 inv = f3(rcp_native(rd.x), rcp_native(rd.y), rcp_native(rd.z));
        const int tid = (int)threadIdx.x;
        const int n_tiles = sc.n_tiles;
#if TRT_V_TMA_TILES
        auto stage = [&](int k) {
            if (tid != 0) return;
            const int4 td = __ldg(&sc.tiles[k]);
            float4 *buf = tile + (k & 1) * 4 * TRT_TILE;
            const unsigned bytes = (unsigned)td.z * 16u;
            const unsigned dbytes = td.x == 2 || td.x == 0 ? (unsigned)((td.z + 1) & ~1) * 16u : 0u;      // cull entries of a triangle / sphere tile
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&mbar[k & 1]);
            // the buffer was read with ordinary loads up to the barrier the caller has just passed: order them before
            // the asynchronous proxy's writes
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(3u * bytes + dbytes) : "memory");
            const float4 *src[3] = {sc.ga + td.y, sc.gb + td.y, sc.gc + td.y};
#pragma unroll
            for (int a = 0; a < 3; a++)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 (unsigned)__cvta_generic_to_shared(buf + a * TRT_TILE)),
                             "l"(src[a]), "r"(bytes), "r"(bar)
                             : "memory");
            if (dbytes)
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                                 (unsigned)__cvta_generic_to_shared(buf + 3 * TRT_TILE)),
                             "l"(sc.gd + td.w), "r"(dbytes), "r"(bar)
                             : "memory");
        };
        auto wait_tile = [&](int k) {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&mbar[k & 1]);
            const unsigned parity = (phase >> (k & 1)) & 1u;
            unsigned done = 0;
            while (!done)
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(done) : "r"(bar), "r"(parity) : "memory");
            phase ^= 1u << (k & 1);
        };
#else
        (void)mbar; (void)phase;
        auto stage = [&](int k) {
            const int4 td = __ldg(&sc.tiles[k]);
            float4 *buf = tile + (k & 1) * 4 * TRT_TILE;
            if (td.x == 2 || td.x == 0)
                for (int j = tid; j < ((td.z + 1) & ~1); j += (int)blockDim.x) __pipeline_memcpy_async(&buf[3 * TRT_TILE + j], &sc.gd[td.w + j], sizeof(float4));
            for (int j = tid; j < td.z; j += (int)blockDim.x) {
                __pipeline_memcpy_async(&buf[j], &sc.ga[td.y + j], sizeof(float4));
                __pipeline_memcpy_async(&buf[TRT_TILE + j], &sc.gb[td.y + j], sizeof(float4));
                __pipeline_memcpy_async(&buf[2 * TRT_TILE + j], &sc.gc[td.y + j], sizeof(float4));
            }
            __pipeline_commit();
        };
        auto wait_tile = [&](int) { __pipeline_wait_prior(0); };
#endif
        if (n_tiles > 0) stage(0);
        bool warp_done = false;
        for (int k = 0; k < n_tiles; k++) {
            wait_tile(k);
            __syncthreads();                                  // tile k has landed; buffer (k + 1) & 1 is free again
            if (k + 1 < n_tiles) stage(k + 1);
            if (!warp_done) {
                const int4 td = __ldg(&sc.tiles[k]);
                const float4 *buf = tile + (k & 1) * 4 * TRT_TILE;
                TileAcc acc{buf, buf + TRT_TILE, buf + 2 * TRT_TILE, td.y, buf + 3 * TRT_TILE};
                scan_run<4>(acc, make_int4(td.x, td.y, td.y + td.z, 0), ro, rd, inv, t_min, closest, best);
                if (any_hit) warp_done = __all_sync(0xffffffffu, best >= 0 || !(t_max > t_min));
            }
        }
        __syncthreads();                                      // the next scan restages buffer 0
        return best >= 0 ? closest : -1.0f;
    }
#endif

    // conv 2, canonical small scene: the same tests in the same order, fully unrolled over fixed constant-bank
    // slots (a slot beyond the scene's count is skipped by a warp-uniform branch).  With a compile-time SHAPE
    // (ShapeFixed: primitive counts and plane kinds known to the compiler) the slot checks and the per-plane
    // kind dispatch vanish: a plane test is one subtract, one multiply and the accept.
    template <class SHAPE>
    __device__ __forceinline__ float scan_small_native(float3 ro, float3 rd, float t_max, int &best) {
        const float t_min = 0.001f;
        float closest = t_max;
        best = -1;
        const int n_sph = SHAPE::kFixed ? SHAPE::NS : c_small.n_sph;
        const int n_pl = SHAPE::kFixed ? SHAPE::NP : c_small.n_pl;
        const int n_tri = SHAPE::kFixed ? SHAPE::NT : c_small.n_tri;
#if TRT_V_PACKED
        // hit_sphere's distance pre-reject for TWO spheres at a time on packed pairs (sphere a in the low lane): tc = c - o,
        // t_ca = tc . d, |tc|^2 and dist2 = |tc|^2 - t_ca^2 are 10 packed instructions for the pair instead of 10 each; the
        // rest of the test (sphere_tail) is per sphere and sees the narrowed bound of its predecessor, as written.
        const f2_t nox = ss2(-ro.x), noy = ss2(-ro.y), noz = ss2(-ro.z), pdx = ss2(rd.x), pdy = ss2(rd.y), pdz = ss2(rd.z);
#pragma unroll
        for (int k = 0; 2 * k < (SHAPE::kFixed ? SHAPE::NS : TRT_STATIC_S); k++) {
            const int j = 2 * k;
            if (j + 1 < (SHAPE::kFixed ? SHAPE::NS : TRT_STATIC_S) && j + 1 < n_sph) {
                const float4 q0 = c_small.sph_pair[3 * k], q1 = c_small.sph_pair[3 * k + 1], q2 = c_small.sph_pair[3 * k + 2];
                const f2_t tcx = f2z_add(f2_pack(q0.x, q0.y), nox), tcy = f2z_add(f2_pack(q0.z, q0.w), noy), tcz = f2z_add(f2_pack(q1.x, q1.y), noz);
                const f2_t tca = f2z_fma(tcz, pdz, f2z_fma(tcy, pdy, f2z_mul(tcx, pdx)));
                const f2_t tcc = f2z_fma(tcz, tcz, f2z_fma(tcy, tcy, f2z_mul(tcx, tcx)));
                const f2_t d2 = f2z_sub(tcc, f2z_mul(tca, tca));          // (contracted to one FFMA2 with a negated product)
                const f2_t rr = f2_pack(q1.z, q1.w);
                float ta, tb, ca, cb_, da, db, fa, fb, na, nb;
                f2_unpack(tca, ta, tb); f2_unpack(tcc, ca, cb_); f2_unpack(d2, da, db);
                f2_unpack(f2z_add(tca, rr), fa, fb); f2_unpack(f2z_sub(tca, rr), na, nb);      // t_ca + r, t_ca - r
                float t = sphere_tail(ta, ca, da, fa, na, q2.x, t_min, closest);
                if (t > 0.0f) { closest = t; best = j; }
                t = sphere_tail(tb, cb_, db, fb, nb, q2.y, t_min, closest);
                if (t > 0.0f) { closest = t; best = j + 1; }
            } else if (j < n_sph) {
                uint32_t cost;
                const float t = test_sphere(ro, rd, c_small.sph[j], c_small.sph_r2[j].x, t_min, closest, cost);
                if (t > 0.0f) { closest = t; best = j; }
            }
        }
#else
#pragma unroll
        for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NS : TRT_STATIC_S); j++) {
            if (j < n_sph) {
                uint32_t cost;
                const float t = test_sphere(ro, rd, c_small.sph[j], c_small.sph_r2[j].x, t_min, closest, cost);
                if (t > 0.0f) { closest = t; best = j; }
            }
        }
#endif
#if TRT_V_KEY
        // index field: 4 / 5 / 6 bits by the scene's primitive count (the same rule for a compile-time and a run-time
        // count, so that a specialised kernel and the run-time-count kernel quantise t alike); all ones = nothing found.
        // The mask is derived from a kernel parameter (FrameParams::all_ones), a value the compiler cannot fold, so
        // that it sits in a register and (bits & mask) | index is ONE LOP3 (register, register, immediate).
        const int n_tot = n_sph + n_pl + n_tri;
        const int kb = n_tot < 15 ? 4 : n_tot < 31 ? 5 : 6;
        const unsigned nkm = fp.all_ones << kb;
        const unsigned KM = (1u << kb) - 1u;                    // a literal when the shape is fixed
        // A bound at or below t_min accepts nothing: key 0 is below every candidate.
        const bool open = closest > t_min;
        unsigned key = open ? ((__float_as_uint(closest - t_min) & nkm) | (best >= 0 ? (unsigned)best : KM)) : 0u;
#define TRT_KEY_OF(tp, idx) ((__float_as_uint(tp) & nkm) | (unsigned)(idx))
#elif TRT_V_SHIFT_T
        // From here on the bound is kept as the bits of closest' = closest - t_min: for t' >= +0 the float order is the
        // unsigned integer order, and a negative or NaN t' has its sign / exponent bits above every valid bound, so
        // `t_min <= t < closest` is ONE unsigned compare.  (conv 2: t' is rounded once more than t; a triangle at
        // exactly t_min is accepted, :214 has `>`.)  A bound at or below t_min accepts nothing.
        unsigned cb = closest > t_min ? __float_as_uint(closest - t_min) : 0u;
#endif
        if (n_pl > 0) {
            // a dead axis (|d| < 1e-4, :180) gets a NaN reciprocal: its t is NaN and fails both accept compares
            const float nan = __int_as_float(0x7fc00000);
            const float3 inv = f3(fabsf(rd.x) >= 0.0001f ? rcp_native(rd.x) : nan, fabsf(rd.y) >= 0.0001f ? rcp_native(rd.y) : nan,
                                  fabsf(rd.z) >= 0.0001f ? rcp_native(rd.z) : nan);
            float3 roi = V::mulv(ro, inv);
            float t_pair_hi = 0.0f;
#if TRT_V_SHIFT_T || TRT_V_KEY
            roi = V::add(roi, f3(t_min, t_min, t_min));                    // axis-aligned planes yield t' directly
#endif
#pragma unroll
            for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NP : TRT_STATIC_P); j++) {
                if (j < n_pl) {
                    const float4 p = c_small.pl_p[j];
                    const int kind = SHAPE::kFixed ? SHAPE::kind(j) : __float_as_int(p.w);
                    float t;
#if TRT_V_PACKED && (TRT_V_SHIFT_T || TRT_V_KEY)
                    // two walls on one axis (compile-time shape): both t' in one FFMA2, the second one is kept for the next round
                    if (SHAPE::kFixed && kind < 3 && (j & 1) && SHAPE::kind(j - 1) == kind) {
                        t = t_pair_hi;
                    } else if (SHAPE::kFixed && kind < 3 && !(j & 1) && j + 1 < SHAPE::NP && SHAPE::kind(j + 1) == kind) {
                        const float2 pp = c_small.pl_pair[j >> 1];
                        const float i_a = kind == 0 ? inv.x : kind == 1 ? inv.y : inv.z, r_a = kind == 0 ? roi.x : kind == 1 ? roi.y : roi.z;
                        f2_unpack(f2z_fma(f2_pack(pp.x, pp.y), ss2(i_a), ss2(-r_a)), t, t_pair_hi);
                    } else
#endif
                    if (kind == 0) t = __fmaf_rn(p.x, inv.x, -roi.x);          // (p - o) / d as p * (1/d) - o * (1/d)
                    else if (kind == 1) t = __fmaf_rn(p.y, inv.y, -roi.y);
                    else if (kind == 2) t = __fmaf_rn(p.z, inv.z, -roi.z);
                    else {
                        const float3 n = xyz(c_small.pl_n[j]);
                        const float denom = V::dot(n, rd);
                        t = V::dot(V::sub(xyz(p), ro), n) * (fabsf(denom) >= 0.0001f ? rcp_native(denom) : nan);
#if TRT_V_SHIFT_T || TRT_V_KEY
                        t -= t_min;
#endif
                    }
#if TRT_V_KEY
                    key = min(key, TRT_KEY_OF(t, n_sph + j));
#elif TRT_V_SHIFT_T
                    if (__float_as_uint(t) < cb) { cb = __float_as_uint(t); best = n_sph + j; }
#else
                    if (t >= t_min && t < closest) { closest = t; best = n_sph + j; }
#endif
                }
            }
        }
#pragma unroll
        for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NT : TRT_STATIC_T); j++) {
            if (j < n_tri) {
                // hit_triangle (shader.wgsl:192-223); z = which edge components are exactly zero (ShapeFixed::TZ)
                const unsigned z = SHAPE::tri_zero(j);
                const bool e1x0 = z & 1u, e1y0 = z & 2u, e1z0 = z & 4u, e2x0 = z & 8u, e2y0 = z & 16u, e2z0 = z & 32u;
                const float3 e1 = xyz(c_small.tri_e1[j]), e2 = xyz(c_small.tri_e2[j]);
                const float3 h = f3(dpz(rd.y, e2.z, e2z0, rd.z, e2.y, e2y0), dpz(rd.z, e2.x, e2x0, rd.x, e2.z, e2z0),
                                    dpz(rd.x, e2.y, e2y0, rd.y, e2.x, e2x0));                       // cross(rd, e2)
                const bool hx0 = e2z0 && e2y0, hy0 = e2x0 && e2z0, hz0 = e2y0 && e2x0;
                const float a = dotz(e1, h, e1x0 || hx0, e1y0 || hy0, e1z0 || hz0);
                if (!(a > -0.00001f && a < 0.00001f)) {
                    const float f = rcp_native(a);
                    const float3 s = V::sub(ro, xyz(c_small.tri_v0[j]));
                    const float u = f * dotz(s, h, hx0, hy0, hz0);
                    if (!(u < 0.0f || u > 1.0f)) {
                        const float3 q = f3(dpz(s.y, e1.z, e1z0, s.z, e1.y, e1y0), dpz(s.z, e1.x, e1x0, s.x, e1.z, e1z0),
                                            dpz(s.x, e1.y, e1y0, s.y, e1.x, e1x0));                   // cross(s, e1)
                        const bool qx0 = e1z0 && e1y0, qy0 = e1x0 && e1z0, qz0 = e1y0 && e1x0;
                        const float v = f * dotz(rd, q, qx0, qy0, qz0);
                        if (!(v < 0.0f || u + v > 1.0f)) {
                            const float t = f * dotz(e2, q, e2x0 || qx0, e2y0 || qy0, e2z0 || qz0);
#if TRT_V_KEY
                            key = min(key, TRT_KEY_OF(t - t_min, n_sph + n_pl + j));
#elif TRT_V_SHIFT_T
                            const unsigned tb = __float_as_uint(t - t_min);
                            if (tb < cb) { cb = tb; best = n_sph + n_pl + j; }
#else
                            if (t > t_min && t < closest) { closest = t; best = n_sph + n_pl + j; }
#endif
                        }
                    }
                }
            }
        }
#if TRT_V_KEY
#undef TRT_KEY_OF
        best = (int)(key & KM);
        if (!open || best == (int)KM) best = -1;
        closest = __uint_as_float(key & nkm) + t_min;
#elif TRT_V_SHIFT_T
        closest = __uint_as_float(cb) + t_min;
#endif
        return best >= 0 ? closest : -1.0f;
    }

#if TRT_V_PACKED && TRT_V_SHIFT_T && !TRT_V_KEY
    // TWO shadow rays from one origin (the samples of two lights for one queued hit) through the canonical small scene
    // in ONE pass, any-hit: the caller reads only `hit.t < 0` (shader.wgsl:400-402), and some primitive is accepted by
    // hit_scene's narrowing scan iff some primitive has t in [t_min, t_max) -- the first one it accepts is tested
    // against t_max itself -- so no bound is narrowed and no index kept.  Every t is computed as scan_small_native
    // computes it (same operations, same order); the rays share the origin, so of a sphere test only tc . d and
    // dist2 are per ray -- on packed pairs, ray A in the low lane -- and a wall's t' for both rays is one FFMA2.
    // t_max <= t_min (a sample that is not needed: t_max = -1) accepts nothing.
    template <class SHAPE>
    __device__ __forceinline__ void shadow2_small(float3 o, float3 dA, float tA, float3 dB, float tB, bool &occA, bool &occB) {
        const float t_min = 0.001f;
        const int n_sph = SHAPE::kFixed ? SHAPE::NS : c_small.n_sph;
        const int n_pl = SHAPE::kFixed ? SHAPE::NP : c_small.n_pl;
        const int n_tri = SHAPE::kFixed ? SHAPE::NT : c_small.n_tri;
        bool hA = false, hB = false;
        const f2_t DX = f2_pack(dA.x, dB.x), DY = f2_pack(dA.y, dB.y), DZ = f2_pack(dA.z, dB.z);
#pragma unroll
        for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NS : TRT_STATIC_S); j++) {
            if (j < n_sph) {
                const float4 sp = c_small.sph[j];
                const float r2 = c_small.sph_r2[j].x;
                const float3 tc = V::sub(xyz(sp), o);
                const float tcc = V::dot(tc, tc);
                const f2_t tca = f2z_fma(ss2(tc.z), DZ, f2z_fma(ss2(tc.y), DY, f2z_mul(ss2(tc.x), DX)));
                const f2_t d2 = f2z_sub(ss2(tcc), f2z_mul(tca, tca));
                const f2_t rr = ss2(sp.w);
                float a0, a1, e0, e1, f0, f1, n0, n1;
                f2_unpack(tca, a0, a1); f2_unpack(d2, e0, e1);
                f2_unpack(f2z_add(tca, rr), f0, f1); f2_unpack(f2z_sub(tca, rr), n0, n1);
                if (sphere_tail(a0, tcc, e0, f0, n0, r2, t_min, tA) > 0.0f) hA = true;
                if (sphere_tail(a1, tcc, e1, f1, n1, r2, t_min, tB) > 0.0f) hB = true;
            }
        }
        // bounds as the bits of t_max - t_min (scan_small_native: TRT_V_SHIFT_T); 0 accepts nothing
        const unsigned cA = tA > t_min ? __float_as_uint(tA - t_min) : 0u, cB = tB > t_min ? __float_as_uint(tB - t_min) : 0u;
        if (n_pl > 0) {
            const float nan = __int_as_float(0x7fc00000);
            const float3 iA = f3(fabsf(dA.x) >= 0.0001f ? rcp_native(dA.x) : nan, fabsf(dA.y) >= 0.0001f ? rcp_native(dA.y) : nan,
                                 fabsf(dA.z) >= 0.0001f ? rcp_native(dA.z) : nan);
            const float3 iB = f3(fabsf(dB.x) >= 0.0001f ? rcp_native(dB.x) : nan, fabsf(dB.y) >= 0.0001f ? rcp_native(dB.y) : nan,
                                 fabsf(dB.z) >= 0.0001f ? rcp_native(dB.z) : nan);
            const f2_t IX = f2_pack(iA.x, iB.x), IY = f2_pack(iA.y, iB.y), IZ = f2_pack(iA.z, iB.z);
            // -(o * inv + t_min), the offset of every axis-aligned plane's t'
            const f2_t NX = f2z_sub(f2z_mul(ss2(-o.x), IX), ss2(t_min)), NY = f2z_sub(f2z_mul(ss2(-o.y), IY), ss2(t_min)),
                       NZ = f2z_sub(f2z_mul(ss2(-o.z), IZ), ss2(t_min));
#pragma unroll
            for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NP : TRT_STATIC_P); j++) {
                if (j < n_pl) {
                    const float4 p = c_small.pl_p[j];
                    const int kind = SHAPE::kFixed ? SHAPE::kind(j) : __float_as_int(p.w);
                    float t0, t1;
                    if (kind == 0) f2_unpack(f2z_fma(ss2(p.x), IX, NX), t0, t1);
                    else if (kind == 1) f2_unpack(f2z_fma(ss2(p.y), IY, NY), t0, t1);
                    else if (kind == 2) f2_unpack(f2z_fma(ss2(p.z), IZ, NZ), t0, t1);
                    else {
                        const float3 n = xyz(c_small.pl_n[j]);
                        const float num = V::dot(V::sub(xyz(p), o), n);
                        const float dnA = V::dot(n, dA), dnB = V::dot(n, dB);
                        t0 = num * (fabsf(dnA) >= 0.0001f ? rcp_native(dnA) : nan) - t_min;
                        t1 = num * (fabsf(dnB) >= 0.0001f ? rcp_native(dnB) : nan) - t_min;
                    }
                    if (__float_as_uint(t0) < cA) hA = true;
                    if (__float_as_uint(t1) < cB) hB = true;
                }
            }
        }
#pragma unroll
        for (int j = 0; j < (SHAPE::kFixed ? SHAPE::NT : TRT_STATIC_T); j++) {
            if (j < n_tri) {
                // hit_triangle (shader.wgsl:192-223) per ray, the zero-folded form of scan_small_native
                const unsigned z = SHAPE::tri_zero(j);
                const bool e1x0 = z & 1u, e1y0 = z & 2u, e1z0 = z & 4u, e2x0 = z & 8u, e2y0 = z & 16u, e2z0 = z & 32u;
                const float3 e1 = xyz(c_small.tri_e1[j]), e2 = xyz(c_small.tri_e2[j]);
                const float3 s = V::sub(o, xyz(c_small.tri_v0[j]));
                const float3 q = f3(dpz(s.y, e1.z, e1z0, s.z, e1.y, e1y0), dpz(s.z, e1.x, e1x0, s.x, e1.z, e1z0),
                                    dpz(s.x, e1.y, e1y0, s.y, e1.x, e1x0));                   // cross(s, e1): the same for both rays
                const bool qx0 = e1z0 && e1y0, qy0 = e1x0 && e1z0, qz0 = e1y0 && e1x0;
                const float e2q = dotz(e2, q, e2x0 || qx0, e2y0 || qy0, e2z0 || qz0);
                const bool hx0 = e2z0 && e2y0, hy0 = e2x0 && e2z0, hz0 = e2y0 && e2x0;
                auto one = [&](float3 rd, unsigned cb) -> bool {
                    const float3 h = f3(dpz(rd.y, e2.z, e2z0, rd.z, e2.y, e2y0), dpz(rd.z, e2.x, e2x0, rd.x, e2.z, e2z0),
                                        dpz(rd.x, e2.y, e2y0, rd.y, e2.x, e2x0));                       // cross(rd, e2)
                    const float a = dotz(e1, h, e1x0 || hx0, e1y0 || hy0, e1z0 || hz0);
                    if (a > -0.00001f && a < 0.00001f) return false;
                    const float f = rcp_native(a);
                    const float u = f * dotz(s, h, hx0, hy0, hz0);
                    if (u < 0.0f || u > 1.0f) return false;
                    const float v = f * dotz(rd, q, qx0, qy0, qz0);
                    if (v < 0.0f || u + v > 1.0f) return false;
                    const float t = f * e2q;
                    return __float_as_uint(t - t_min) < cb;
                };
                if (one(dA, cA)) hA = true;
                if (one(dB, cB)) hB = true;
            }
        }
        occA = hA;
        occB = hB;
    }
#endif

    // any-hit form of the same scan for shadow rays: only `hit.t < 0` is read by the caller
    // (shader.wgsl:400-402), and before the first accepted hit the bound is still t_max, so stopping at
    // the first accepted primitive returns the same truth value.
    __device__ __forceinline__ bool scan_any(float3 ro, float3 rd, float t_min, float t_max) {
        for (int r = 0; r < sc.n_runs; r++) {
            const int4 run = src.run(r);
#pragma unroll 1                        // one copy of each test: the loop body must stay inside the instruction cache
            for (int i = run.y; i < run.z; i++) {
                uint32_t cost;
                const float t = test_any(run.x, i, ro, rd, t_min, t_max, cost);
                if (t > 0.0f && t < t_max) return true;
            }
        }
        return false;
    }

    // shader.wgsl:234-245 (plane / triangle normals precomputed with the same operations); the index is
    // lane-divergent here, so these come from HBM/L1 through the read-only path
    __device__ __forceinline__ float3 normal_at(int i, int type, float3 p) const {
        if (type == 0) {
            float4 a = __ldg(&sc.ga[i]);
            if (CONV == 2) {                                    // (p - c) / r normalised: only the sign of r survives
                const float3 d = V::sub(p, xyz(a));
                return V::scale(d, copysignf(rsqrt_native(V::dot(d, d)), a.w));
            }
            return V::normalize(V::divs(V::sub(p, xyz(a)), a.w));
        }
        if (type >= 1 && type <= 8) return xyz(__ldg(&sc.gn[i]));
        return f3(0.0f, 1.0f, 0.0f);
    }

    // shader.wgsl:297-304
    __device__ __forceinline__ float3 sky(float3 d) const {
        float t = 0.5f * (d.y + 1.0f);
        float3 c = V::add(V::scale(f3(1.0f, 1.0f, 1.0f), 1.0f - t), V::scale(f3(0.5f, 0.7f, 1.0f), t));
        return V::scale(c, 0.8f);
    }

    // shader.wgsl:312-335
    __device__ __forceinline__ float3 cosine_hemisphere(uint32_t &rng, float3 normal) {
        float r1 = rnd(rng), r2 = rnd(rng);
        if (COUNT) cn.add(K_RNG_DRAWS, 2);
        float cos_theta = sqrtf(r1);
        float sin_theta = sqrtf(1.0f - r1);
        float sp, cp;
        sincos_turn<CONV>(r2, sp, cp);                      // phi = 2 pi r2 (:317)
        float x = sin_theta * cp, y = sin_theta * sp, z = cos_theta;
        float3 w = V::normalize(normal);
        // one cross + normalize for both helper axes (:327-331): same operations, no divergent branch
        const bool use_y = fabsf(w.x) > 0.1f;
        float3 u = V::normalize(V::cross(f3(use_y ? 0.0f : 1.0f, use_y ? 1.0f : 0.0f, 0.0f), w));
        float3 v = V::cross(w, u);
        return V::normalize(V::lin3(u, x, v, y, w, z));
    }

    // conv 2 form of the same function for a UNIT normal (the hit record's): the re-normalisation of w is a
    // rounding-level no-op, the helper axis is (0,1,0) or (1,0,0), so u = helper x w has one zero component and
    // v = w x u one component equal to |helper x w|: written out (the x <-> y mirror image serves the second
    // helper), the frame costs one rsqrt instead of two normalisations and two cross products.
    __device__ __forceinline__ float3 cosine_hemisphere_unit(uint32_t &rng, float3 w) {
        const float r1 = rnd(rng), r2 = rnd(rng);
        const float cos_theta = sqrt_native(r1);
        const float sin_theta = sqrt_native(1.0f - r1);
        float sp, cp;
        sincos_turn<2>(r2, sp, cp);
        const float y = sin_theta * sp, z = cos_theta;
        const bool use_y = fabsf(w.x) > 0.1f;
        const float x = use_y ? sin_theta * cp : -(sin_theta * cp);
        const float wx = use_y ? w.x : w.y, wy = use_y ? w.y : w.x;        // mirrored frame for the x helper
        const float n2 = __fmaf_rn(w.z, w.z, wx * wx);
        const float inv = rsqrt_native(n2);
        const float a = w.z * inv, b = wx * inv, nn = n2 * inv, t = wy * y;
        const float dx = __fmaf_rn(wx, z, __fmaf_rn(-b, t, a * x));
        const float dy = __fmaf_rn(wy, z, nn * y);
        const float dz = __fmaf_rn(w.z, z, __fmaf_rn(-a, t, -(b * x)));
#if TRT_V_NO_RENORM
        return f3(use_y ? dx : dy, use_y ? dy : dx, dz);
#else
        return V::normalize(f3(use_y ? dx : dy, use_y ? dy : dx, dz));
#endif
    }

    // shader.wgsl:338-436.  The loop runs over the precomputed list of emissive spheres/triangles
    // (same order; plane emitters and dark primitives consume no random numbers, :346-347,:390-392).
    __device__ __forceinline__ float3 direct_lighting(uint32_t &rng, float3 hit_point, float3 normal, float3 color,
                                                      float3 attenuation) {
        float3 direct = f3(0.0f, 0.0f, 0.0f);
        for (int k = 0; k < sc.n_lights; k++) {
            const float4 a = src.la(k);
            const float4 nl = src.ln(k);
            const float4 em = src.le(k);
            const int type = __float_as_int(em.w);
            float3 light_point, light_normal;
            float r1 = rnd(rng), r2 = rnd(rng);
            if (COUNT) cn.add(K_RNG_DRAWS, 2);
            if (type == 0) {                                   // uniform sphere sampling :352-369
                float cos_theta = 1.0f - 2.0f * r1;
                float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
                float sp, cp;
                sincos_turn<CONV>(r2, sp, cp);              // phi = 2 pi r2 (:355)
                float3 local = f3(sin_theta * cp, sin_theta * sp, cos_theta);
                light_point = V::axpy(xyz(a), local, a.w);
                light_normal = local;
                if (COUNT) cn.flops(20);
            } else {                                           // uniform triangle sampling :370-389
                float sqrt_r1 = sqrtf(r1);
                float u = 1.0f - sqrt_r1;
                float v = r2 * sqrt_r1;
                light_point = V::lin3(xyz(a), 1.0f - u - v, xyz(src.lb(k)), u, xyz(src.lc(k)), v);
                light_normal = xyz(nl);
                if (COUNT) cn.flops(60);
            }
            float light_area = nl.w;
            float3 light_vec = V::sub(light_point, hit_point);
            float light_distance = V::length(light_vec);
            float3 light_dir = V::divs(light_vec, light_distance);
            float3 so = V::axpy(hit_point, normal, 0.001f);
            if (COUNT) cn.flops(19);
            // the two cosines do not depend on the shadow test (:403-404), so they are evaluated first
            float cos_surface = fmaxf(0.0f, V::dot(normal, light_dir));
            float cos_light = fmaxf(0.0f, V::dot(light_normal, V::scale(light_dir, -1.0f)));
            bool needed = cos_surface > 0.0f && cos_light > 0.0f;
            bool unoccluded;
            if (COUNT) {
                cn.add(K_SHADOW_SCANS, 1);
                cn.add(K_SHADOW_NEEDED, needed ? 1u : 0u);
                int dummy;
                unoccluded = scan_closest(so, light_dir, 0.001f, light_distance - 0.001f, dummy, needed ? 1 : 0) < 0.0f;
                if (unoccluded) { cn.add(K_SHADOW_UNOCCLUDED, 1); cn.flops(13); }
            } else {
                // a scan whose result cannot change the sum is skipped (no RNG inside hit_scene)
                unoccluded = needed && !scan_any(so, light_dir, 0.001f, light_distance - 0.001f);
            }
            if (unoccluded && needed) {
                if (COUNT) { cn.add(K_SHADOW_LIT, 1); cn.flops(19); }
                float distance_squared = light_distance * light_distance;
                float geometric = (cos_surface * cos_light) / distance_squared;
                float3 brdf = V::divs(color, TRT_PI_F);
                float3 contrib = V::mulv(V::mulv(brdf, xyz(em)), V::scale(attenuation, geometric * light_area));
                direct = V::add(direct, f3(fminf(contrib.x, 10.0f), fminf(contrib.y, 10.0f), fminf(contrib.z, 10.0f)));
            }
        }
        return direct;
    }

    // One iteration of the bounce loop shader.wgsl:444-493 (i = bounce index).  Returns true when the
    // path ends here (miss or roulette); the caller also stops at i + 1 == max_depth.
    __device__ __forceinline__ bool bounce(uint32_t &rng, float3 &ro, float3 &rd, float3 &att, float3 &acc, uint32_t i) {
        int idx;
        n_bounces++;
        if (COUNT) cn.add(K_CLOSEST_SCANS, 1);
        const float t = scan_closest(ro, rd, 0.001f, 1e10f, idx, -1);
        if (t < 0.0f) {
            if (COUNT) { cn.add(K_MISSES, 1); cn.flops(21); }
            acc = V::vmad(acc, sky(rd), att);
            return true;
        }
        if (COUNT) { cn.add(K_HITS, 1); cn.flops(9); }
        const float4 mc = __ldg(&sc.mc[idx]);
        const float4 me = __ldg(&sc.me[idx]);
        const int type = __float_as_int(me.w);
        const float3 p = V::axpy(ro, rd, t);
        float3 n = normal_at(idx, type, p);
        if (!(V::dot(rd, n) < 0.0f)) n = V::scale(n, -1.0f);
        acc = V::vmad(acc, xyz(me), att);
        acc = V::add(acc, direct_lighting(rng, p, n, xyz(mc), att));
        float3 scatter;
        if (COUNT) cn.add(K_RNG_DRAWS, 1);
        if (mc.w > rnd(rng)) {
            if (COUNT) { cn.add(K_MIRROR, 1); cn.flops(12); }
            scatter = V::reflect(rd, n);
        } else {
            if (COUNT) { cn.add(K_DIFFUSE, 1); cn.flops(75); }
            scatter = cosine_hemisphere(rng, n);
        }
        att = V::mulv(att, xyz(mc));
        if (COUNT) cn.flops(9);
        ro = V::axpy(p, scatter, 0.001f);
        rd = scatter;
        if (i > 3u) {
            float survival = fminf(0.95f, fmaxf(att.x, fmaxf(att.y, att.z)));
            if (COUNT) { cn.add(K_ROULETTE_TESTS, 1); cn.add(K_RNG_DRAWS, 1); }
            if (survival < rnd(rng)) { if (COUNT) cn.add(K_ROULETTE_KILLS, 1); return true; }
            if (COUNT) cn.flops(3);
            att = V::divs(att, survival);
        }
        return false;
    }

    // shader.wgsl:439-495
    __device__ __forceinline__ float3 ray_color(uint32_t &rng, float3 ro, float3 rd) {
        float3 acc = f3(0.0f, 0.0f, 0.0f);
        float3 att = f3(1.0f, 1.0f, 1.0f);
        for (uint32_t i = 0; i < fp.max_depth; i++)
            if (bounce(rng, ro, rd, att, acc, i)) break;
        return acc;
    }

    // shader.wgsl:519-546
    __device__ __forceinline__ void camera_ray(uint32_t &rng, uint32_t x, uint32_t y, float3 &ro, float3 &rd) {
        if (COUNT) cn.add(K_RNG_DRAWS, 2);
#if TRT_V_CAMFOLD
        if (CONV == 2) {
            // the same affine map with the frame-uniform divisions and products folded into three host constants
            // (FrameParams::cam_kx/ky/cy): two fused multiply-adds per axis instead of :520-531's chain
            const float px = __uint2float_rn(x) + rnd(rng);
            const float py = __uint2float_rn(fp.height - 1u - y) + rnd(rng);
            const float vx = __fmaf_rn(px, fp.cam_kx, -fp.half_w);
            const float vy = __fmaf_rn(py, fp.cam_ky, -fp.cam_cy);
            rd = V::normalize(V::axpy(V::axpy(fp.cam_fwd, fp.cam_up, vy), fp.cam_right, vx));
            ro = fp.cam_pos;
            return;
        }
#endif
        float u = (__uint2float_rn(x) + rnd(rng)) / fp.wm1;
        float v = (__uint2float_rn(fp.height - 1u - y) + rnd(rng)) / fp.hm1;
        float ndc_x = 2.0f * u - 1.0f;
        float ndc_y = (2.0f * v - 1.0f) / fp.char_aspect;
        float vx = fp.half_w * ndc_x;
        float vy = fp.half_h * ndc_y;
        rd = V::normalize(V::lin3(fp.cam_right, vx, fp.cam_up, vy, fp.cam_fwd, 1.0f));
        ro = fp.cam_pos;
    }

    // shader.wgsl:555-562: variance of the base samples and the number of adaptive extra samples
    __device__ __forceinline__ uint32_t adaptive_extra(float3 color_sum, float3 sq_sum, float &variance_value) const {
        float3 mean = V::divs(color_sum, fp.base_f);
        float3 var_vec = V::sub(V::divs(sq_sum, fp.base_f), V::mulv(mean, mean));
        variance_value = var_vec.x + var_vec.y + var_vec.z;
        if (variance_value > 10.0f && fp.base_samples < fp.spp)
            return min(fp.spp - fp.base_samples, __float2uint_rz(variance_value * 50.0f));
        return 0u;
    }

    // shader.wgsl:499-621 for one pixel
    __device__ __forceinline__ void pixel(uint32_t x, uint32_t y) {
        uint32_t rng = (y * fp.width + x) * 1973u + fp.seed * 9277u + fp.frame_number * 12345u;
        float3 pixel_color = f3(0, 0, 0), sq_sum = f3(0, 0, 0);
        const uint32_t base = fp.base_samples;
        if (COUNT) { cn.add(K_PIXELS, 1); cn.flops(30); }
        for (uint32_t i = 0; i < base; i++) {
            rng = pcg_hash(rng + i * 5096u);
            float3 ro, rd;
            camera_ray(rng, x, y, ro, rd);
            float3 c = ray_color(rng, ro, rd);
            n_samples++;
            pixel_color = V::add(pixel_color, c);        // color_sum == pixel_color during the base samples (:549-550)
            sq_sum = V::add(sq_sum, V::mulv(c, c));
            if (COUNT) { cn.add(K_SAMPLES_BASE, 1); cn.flops(50); }
        }
        float variance_value;
        const uint32_t additional = adaptive_extra(pixel_color, sq_sum, variance_value);
        const bool adaptive = variance_value > 10.0f && base < fp.spp;
        for (uint32_t i = 0; i < additional; i++) {
            rng = pcg_hash(rng + (base + i) * 5096u);
            float3 ro, rd;
            camera_ray(rng, x, y, ro, rd);
            pixel_color = V::add(pixel_color, ray_color(rng, ro, rd));
            n_samples++;
            if (COUNT) { cn.add(K_SAMPLES_EXTRA, 1); cn.flops(41); }
        }
        if (adaptive) pixel_color = V::divs(V::scale(pixel_color, fp.spp_f), __uint2float_rn(base + additional));
        finish_pixel(x, y, pixel_color, variance_value);
    }

    // shader.wgsl:605-621: variance store, temporal accumulation, pixel store (one float4 per buffer).
    // In a frame batch (fp.batch) the pixel's mean goes to the batch buffer instead (fslot = frame * n_local_pixels
    // + local pixel) and only the last frame of the batch leaves its variance behind, as sequential frames would.
    __device__ __forceinline__ void finish_pixel(uint32_t x, uint32_t y, float3 pixel_color, float variance_value, uint32_t fslot = 0u) {
        size_t index = (size_t)y * fp.width + x;
        trt_variance vb;
        vb.variance = variance_value;
        vb.sample_count = fp.base_samples;
        float3 cur = V::divs(pixel_color, fp.spp_f);
        if (fp.batch) {
            fp.batch_cur[fslot] = make_float4(cur.x, cur.y, cur.z, 0.0f);
            if (fslot >= (fp.n_batch - 1u) * fp.n_local_pixels) fp.variance[index] = vb;
            return;
        }
        fp.variance[index] = vb;
        float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (fp.frame_number != 0u) a = fp.accumulation[index];
        const float4 out = blend_pixel<CONV>(a, cur, fp.frame_number);
        fp.accumulation[index] = out;
        if (fp.pixels != fp.accumulation) fp.pixels[index] = out;       // (the library aliases the two: one store)
    }
};

}  // namespace trt
