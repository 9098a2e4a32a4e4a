// trt_kernels.cuh -- __global__ entry points of the trace path (sm_100a).
#pragma once
#include "trt_trace.cuh"

namespace trt {

// ---- K0: scene repack, AoS 180 B (primitive.rs:4-34) -> SoA float4 arrays --------------------------
// Derived values use the same helper functions (and therefore the same roundings, in the context's
// arithmetic convention) as the shader expressions they replace:
//   r*r (shader.wgsl:264), edge1/edge2 (:193-194,:240-241,:386-387), normalize(plane_normal) (:238),
//   normalize(cross(edge1,edge2)) (:242,:388), 0.5*length(cross) (:389), 4*pi*r*r (:369).
struct RepackParams {
    const trt_primitive *aos;
    int n_prims;
    float4 *ga, *gb, *gc, *gn, *mc, *me;
    const int *light_prim;       // prim index of each light (host-built, scan order)
    int n_lights;
    float4 *la, *lb, *lc, *ln, *le;
};

__device__ __forceinline__ float3 ld3(const trt_vec3 &v) { return make_float3(v.x, v.y, v.z); }
__device__ __forceinline__ float4 f4(float3 v, float w) { return make_float4(v.x, v.y, v.z, w); }

template <int CONV>
__global__ void repack_kernel(RepackParams rp) {
    typedef M<CONV> V;
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < rp.n_prims) {
        const trt_primitive &p = rp.aos[i];
        float4 a = make_float4(0, 0, 0, 0), b = a, c = a, n = make_float4(0.0f, 1.0f, 0.0f, 0.0f);
        if (p.primitive_type == 0u) {
            a = f4(ld3(p.sphere_center), p.sphere_radius);
            b.x = p.sphere_radius * p.sphere_radius;
        } else if (p.primitive_type == 1u) {
            a = f4(ld3(p.plane_point), 0.0f);
            b = f4(ld3(p.plane_normal), 0.0f);
            n = f4(V::normalize(ld3(p.plane_normal)), 0.0f);
        } else if (p.primitive_type == 2u) {
            float3 e1 = V::sub(ld3(p.triangle_v1), ld3(p.triangle_v0));
            float3 e2 = V::sub(ld3(p.triangle_v2), ld3(p.triangle_v0));
            // .w: squared radius of a sphere around v0 that holds the triangle, 10 % to spare (scan_run's cull, conv 2)
            a = f4(ld3(p.triangle_v0), 1.21f * fmaxf(V::dot(e1, e1), V::dot(e2, e2)));
            b = f4(e1, 0.0f);
            c = f4(e2, 0.0f);
            n = f4(V::normalize(V::cross(e1, e2)), 0.0f);
        }
        rp.ga[i] = a; rp.gb[i] = b; rp.gc[i] = c; rp.gn[i] = n;
        rp.mc[i] = f4(ld3(p.color), p.reflectivity);
        rp.me[i] = f4(ld3(p.emission), __int_as_float((int)p.primitive_type));
    }
    if (i < rp.n_lights) {
        const trt_primitive &p = rp.aos[rp.light_prim[i]];
        float4 a, b = make_float4(0, 0, 0, 0), c = b, n;
        if (p.primitive_type == 0u) {
            a = f4(ld3(p.sphere_center), p.sphere_radius);
            n = make_float4(0, 0, 0, TRT_FOUR_PI_F * p.sphere_radius * p.sphere_radius);
            c.w = fabsf(a.x) + fabsf(a.y) + fabsf(a.z) + a.w;      // magnitude scale for the reach test (trt_native.cuh)
        } else {
            float3 e1 = V::sub(ld3(p.triangle_v1), ld3(p.triangle_v0));
            float3 e2 = V::sub(ld3(p.triangle_v2), ld3(p.triangle_v0));
            float3 cr = V::cross(e1, e2);
            a = f4(ld3(p.triangle_v0), 0.0f);
            b = f4(ld3(p.triangle_v1), 0.0f);
            c = f4(ld3(p.triangle_v2), 0.0f);
            n = f4(V::normalize(cr), 0.5f * V::length(cr));
            b.w = V::dot(xyz(n), xyz(a));                          // plane offset of the light (reach test, trt_native.cuh)
            c.w = fabsf(b.w);
        }
        rp.la[i] = a; rp.lb[i] = b; rp.lc[i] = c; rp.ln[i] = n;
        rp.le[i] = f4(ld3(p.emission), __int_as_float((int)p.primitive_type));
    }
}

// Cull data of the triangle and sphere tiles, two primitives per entry pair (SceneView::gd): one block per tile.
// Per primitive a point and a squared distance R such that a ray whose line stays further from the point than that
// cannot hit: triangle (v0, ga.w); sphere (centre, r^2 with a little to spare).
static __global__ void pack_cull_kernel(const float4 *ga, const int4 *tiles, int n_tiles, float4 *gd) {
    const int k = blockIdx.x;
    if (k >= n_tiles) return;
    const int4 td = tiles[k];
    if (td.x != 2 && td.x != 0) return;
    for (int j = threadIdx.x; 2 * j < td.z; j += blockDim.x) {
        float4 a = ga[td.y + 2 * j];
        // (an odd tile's last entry: a partner that is never near -- the scan tests that primitive on its own anyway)
        float4 b = 2 * j + 1 < td.z ? ga[td.y + 2 * j + 1] : make_float4(0.0f, 0.0f, 0.0f, __int_as_float(0xff800000));
        if (td.x == 0) {
            a.w = a.w * a.w * 1.0001f;
            if (2 * j + 1 < td.z) b.w = b.w * b.w * 1.0001f;
        }
        gd[td.w + 2 * j] = make_float4(a.x, b.x, a.y, b.y);
        gd[td.w + 2 * j + 1] = make_float4(a.z, b.z, -a.w, -b.w);
    }
}

// ---- K2 (v0): one thread per pixel, loops exactly as shader.wgsl main() ----------------------------
// Kept as the plain device restatement (and the instrumented build); the production kernel is
// trace_persistent_kernel below.
template <int CONV, bool COUNT>
__global__ void __launch_bounds__(128) trace_pixel_kernel(SceneView sc, LightView lv, FrameParams fp) {
    Counters<COUNT> cn;
    Tracer<CONV, COUNT, SrcGlobal> tr(sc, lv, fp, cn);
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp < fp.n_local_pixels) {
        const uint32_t lr = lp / fp.width, x = lp - lr * fp.width;
        const uint32_t y = fp.row_map ? fp.row_map[lr] : lr;
        tr.pixel(x, y);
    }
    tr.flush_stats();
    if (COUNT) cn.flush(fp.counters);
}

// ---- hand-out order: longest pixels first ------------------------------------------------------------
// A pixel is a serial chain (one RNG state, shader.wgsl:507-565) of base samples and, when the variance of
// the base samples exceeds 10, up to 3x as many extra ones (:555-565): at 1024 spp the longest chain runs
// for tens of milliseconds, and the frame's makespan is bounded by the pixels started last.  The variance
// buffer of the previous frame (:608) predicts every pixel's sample count by the shader's own rule (:559-562);
// a counting sort over TRT_ORDER_BUCKETS classes of predicted length hands the longest pixels out first and
// lets the shortest fill the tail.  Scheduling only: a pixel's result is independent of the order
// (bit-identical frames).  order_count_kernel builds the histogram, order_scatter_kernel places the pixels
// (order inside a class is arbitrary).
#define TRT_ORDER_BUCKETS 32
// With `cost` (the bounce count the production kernel measured for every pixel of the previous frame) the classes are
// eighth-octaves of the measured chain length below max_cost = spp * max_depth: what bounds a short frame is its
// longest chains (deep paths x many samples), which the sample count alone does not tell.
__device__ __forceinline__ uint32_t order_bucket(const trt_variance *variance, const uint32_t *row_map, uint32_t width, uint32_t lp,
                                                 uint32_t base_samples, uint32_t spp, const uint32_t *cost, float max_cost) {
    const uint32_t lr = lp / width, x = lp - lr * width;
    const uint32_t y = row_map ? row_map[lr] : lr;
    if (cost) {
        const float c = fmaxf(1.0f, (float)cost[(size_t)y * width + x]);
        const float b = 8.0f * log2f(fmaxf(1.0f, max_cost / c));
        return min((uint32_t)b, (uint32_t)TRT_ORDER_BUCKETS - 1u);                               // bucket 0 = longest
    }
    const float v = variance[(size_t)y * width + x].variance;
    uint32_t predicted = base_samples;
    if (v > 10.0f && base_samples < spp) predicted += min(spp - base_samples, __float2uint_rz(v * 50.0f));
    const uint32_t cls = (uint32_t)(((uint64_t)predicted * (TRT_ORDER_BUCKETS - 1)) / (spp ? spp : 1u));
    return (TRT_ORDER_BUCKETS - 1u) - min(cls, (uint32_t)TRT_ORDER_BUCKETS - 1u);        // bucket 0 = longest
}
static __global__ void order_count_kernel(const trt_variance *variance, const uint32_t *row_map, uint32_t width,
                                          uint32_t n_local_pixels, uint32_t base_samples, uint32_t spp, unsigned int *hist,
                                          const uint32_t *cost, float max_cost) {
    __shared__ unsigned int sh[TRT_ORDER_BUCKETS];
    if (threadIdx.x < TRT_ORDER_BUCKETS) sh[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp < n_local_pixels) atomicAdd(&sh[order_bucket(variance, row_map, width, lp, base_samples, spp, cost, max_cost)], 1u);
    __syncthreads();
    if (threadIdx.x < TRT_ORDER_BUCKETS && sh[threadIdx.x]) atomicAdd(&hist[threadIdx.x], sh[threadIdx.x]);
}
// hist[0..B) = class sizes, hist[B..2B) = cursors (zeroed by the host)
static __global__ void order_scatter_kernel(const trt_variance *variance, const uint32_t *row_map, uint32_t width,
                                            uint32_t n_local_pixels, uint32_t base_samples, uint32_t spp, unsigned int *hist,
                                            uint32_t *order, const uint32_t *cost, float max_cost) {
    __shared__ unsigned int cnt[TRT_ORDER_BUCKETS], start[TRT_ORDER_BUCKETS], base[TRT_ORDER_BUCKETS];
    if (threadIdx.x < TRT_ORDER_BUCKETS) cnt[threadIdx.x] = 0u;
    __syncthreads();
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t b = 0, rank = 0;
    const bool live = lp < n_local_pixels;
    if (live) {
        b = order_bucket(variance, row_map, width, lp, base_samples, spp, cost, max_cost);
        rank = atomicAdd(&cnt[b], 1u);
    }
    __syncthreads();
    if (threadIdx.x < TRT_ORDER_BUCKETS) {
        unsigned int off = 0;
        for (uint32_t i = 0; i < threadIdx.x; i++) off += hist[i];
        start[threadIdx.x] = off;
        base[threadIdx.x] = cnt[threadIdx.x] ? atomicAdd(&hist[TRT_ORDER_BUCKETS + threadIdx.x], cnt[threadIdx.x]) : 0u;
    }
    __syncthreads();
    if (live) order[start[b] + base[b] + rank] = lp;
}

// ---- frame batch: temporal blend in frame order (shader.wgsl:610-619) ------------------------------------
// cur[f][lp] holds frame f's pixel mean (finish_pixel); one thread per local pixel walks the frames in order and
// applies exactly the blend a sequence of single-frame launches applies, so the accumulation after the batch is
// bit-identical to theirs.  The running value is written back over cur[f][lp]: frame f's displayed pixels (:620),
// which is what the quantiser reads when every frame's byte stream is wanted.
template <int CONV>
static __global__ void blend_frames_kernel(float4 *cur, const uint2 *batch, uint32_t n_batch, uint32_t n_local_pixels,
                                           const uint32_t *row_map, uint32_t width, float4 *accumulation, float4 *pixels) {
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp >= n_local_pixels) return;
    const uint32_t lr = lp / width, x = lp - lr * width;
    const uint32_t y = row_map ? row_map[lr] : lr;
    const size_t index = (size_t)y * width + x;
    float4 a = accumulation[index];
    for (uint32_t f = 0; f < n_batch; f++) {
        const size_t s = (size_t)f * n_local_pixels + lp;
        const float4 c = cur[s];
        a = blend_pixel<CONV>(a, make_float3(c.x, c.y, c.z), __ldg(&batch[f]).y);
        cur[s] = a;
    }
    accumulation[index] = a;
    if (pixels != accumulation) pixels[index] = a;
}

// pack the rows this shard owns contiguously (local row order) for the NCCL gather
static __global__ void pack_rows_kernel(const float4 *src, float4 *dst, const uint32_t *row_map, uint32_t width,
                                 uint32_t n_local_pixels) {
    const uint32_t lp = blockIdx.x * blockDim.x + threadIdx.x;
    if (lp < n_local_pixels) {
        const uint32_t lr = lp / width, x = lp - lr * width;
        const uint32_t y = row_map ? row_map[lr] : lr;
        dst[lp] = src[(size_t)y * width + x];
    }
}

}  // namespace trt
