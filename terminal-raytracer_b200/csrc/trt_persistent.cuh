// trt_persistent.cuh -- the production trace kernel: persistent warps, one path state per lane,
// lane-level path regeneration.
//
// Why: in shader.wgsl every invocation runs `for sample { for bounce { ... } }` for one pixel, so a
// warp waits for its longest path (1..32 bounces) on every sample and for its most-sampled pixel
// (adaptive sampling: base .. 4 x base samples).  Here each lane is a small state machine
//     NEED_PIXEL -> NEED_SAMPLE -> (bounce)* -> sample done -> ... -> pixel done -> NEED_PIXEL
// and one loop iteration executes ONE bounce for every lane that has a live path; a lane whose path
// ended starts its next sample (or fetches its next pixel from a global counter) in the same
// iteration.  The per-pixel order of random numbers, samples and bounces is untouched (a pixel still
// belongs to one lane from its first to its last sample: the RNG chain of shader.wgsl:507,517,565 is
// serial per pixel), so results are bit-identical to the per-pixel loop nest.
//
// Grid: one CTA of TRT_PT_THREADS threads per resident slot, SM count x TRT_PT_CTAS_PER_SM (or fewer
// CTAs when the frame has fewer pixels than lanes).
#pragma once
#include "trt_kernels.cuh"

namespace trt {

#ifndef TRT_PT_THREADS
#define TRT_PT_THREADS 128
#endif
#ifndef TRT_PT_CTAS_PER_SM
#define TRT_PT_CTAS_PER_SM 4
#endif

enum { ST_NEED_PIXEL = 0, ST_NEED_SAMPLE = 1, ST_TRACE = 2, ST_DONE = 3 };

template <int CONV, class SRC>
__global__ void __launch_bounds__(TRT_PT_THREADS, TRT_PT_CTAS_PER_SM)
trace_persistent_kernel(SceneView sc, LightView lv, FrameParams fp) {
    typedef M<CONV> V;
    Counters<false> cn;
    Tracer<CONV, false, SRC> tr(sc, lv, fp, cn);
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;

    int state = ST_NEED_PIXEL;
    uint32_t x = 0, y = 0, rng = 0;
    uint32_t si = 0, target = 0;            // samples done / samples to do in the current phase
    bool extra_phase = false;
    float variance_value = 0.0f;
    float3 pixel_color = f3(0, 0, 0), sq_sum = f3(0, 0, 0);
    float3 ro = f3(0, 0, 0), rd = f3(0, 0, 1), att = f3(1, 1, 1), acc = f3(0, 0, 0);
    uint32_t depth = 0;

    for (;;) {
        // ---- hand out pixels: one atomic per warp per round (all lanes converge here) ------------
        const unsigned need = __ballot_sync(0xffffffffu, state == ST_NEED_PIXEL);
        if (need) {
            const int leader = __ffs(need) - 1;
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(fp.work_counter, __popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (state == ST_NEED_PIXEL) {
                const uint32_t lp = base + __popc(need & lt_mask);
                if (lp < fp.n_local_pixels) {
                    const uint32_t lr = lp / fp.width;
                    x = lp - lr * fp.width;
                    y = fp.row_map ? fp.row_map[lr] : lr;
                    rng = (y * fp.width + x) * 1973u + fp.seed * 9277u + fp.frame_number * 12345u;   // shader.wgsl:507
                    pixel_color = f3(0, 0, 0);
                    sq_sum = f3(0, 0, 0);
                    si = 0;
                    target = fp.base_samples;
                    extra_phase = false;
                    state = ST_NEED_SAMPLE;
                } else {
                    state = ST_DONE;
                }
            }
        }
        if (__all_sync(0xffffffffu, state == ST_DONE)) break;

        // ---- start a sample: reseed + camera ray (shader.wgsl:517-546 / :565-594) ---------------------
        if (state == ST_NEED_SAMPLE) {
            rng = pcg_hash(rng + si * 5096u);          // si counts base and extra samples alike: (base + i)
            tr.camera_ray(rng, x, y, ro, rd);
            att = f3(1.0f, 1.0f, 1.0f);
            acc = f3(0.0f, 0.0f, 0.0f);
            depth = 0;
            state = fp.max_depth ? ST_TRACE : ST_NEED_SAMPLE;
        }

        // ---- one bounce for every live path ---------------------------------------------------------------
        bool path_end = (state == ST_NEED_SAMPLE);     // max_depth == 0: the sample is black
        if (state == ST_TRACE) {
            path_end = tr.bounce(rng, ro, rd, att, acc, depth);
            depth++;
            if (depth >= fp.max_depth) path_end = true;
        }

        // ---- sample / pixel bookkeeping (shader.wgsl:548-551, :555-565, :596-621) -----------------------
        if (path_end) {
            tr.n_samples++;
            pixel_color = V::add(pixel_color, acc);
            if (!extra_phase) sq_sum = V::add(sq_sum, V::mulv(acc, acc));
            si++;
            state = ST_NEED_SAMPLE;
            if (si == target) {
                bool finished = true;
                if (!extra_phase) {
                    const uint32_t additional = tr.adaptive_extra(pixel_color, sq_sum, variance_value);
                    if (variance_value > 10.0f && fp.base_samples < fp.spp) {
                        extra_phase = true;
                        target = fp.base_samples + additional;
                        finished = (additional == 0u);
                    }
                }
                if (finished) {
                    if (extra_phase) pixel_color = V::divs(V::scale(pixel_color, fp.spp_f), __uint2float_rn(target));
                    tr.finish_pixel(x, y, pixel_color, variance_value);
                    state = ST_NEED_PIXEL;
                }
            }
        }
    }
    tr.flush_stats();
}

// ---- host side ------------------------------------------------------------------------------------------
struct ConstOwner { const void *key = nullptr; unsigned long long serial = 0; };
static ConstOwner g_const_owner[64];

static inline bool scene_fits_constant(const SceneView &sv) {
    return sv.n_prims <= TRT_CONST_PRIMS && sv.n_lights <= TRT_CONST_LIGHTS && sv.n_runs <= TRT_CONST_RUNS;
}

// copy the SoA arrays (device) into the __constant__ mirror; D2D, ordered on `stream`
static inline cudaError_t persistent_scene_upload(const SceneView &sv, const LightView &lv, cudaStream_t stream) {
    if (!scene_fits_constant(sv)) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    auto put = [&](const void *src, size_t off, size_t n) {
        if (e == cudaSuccess && n) e = cudaMemcpyToSymbolAsync(c_scene, src, n, off, cudaMemcpyDeviceToDevice, stream);
    };
    const size_t np = (size_t)sv.n_prims * sizeof(float4), nl = (size_t)sv.n_lights * sizeof(float4);
    put(sv.ga, offsetof(ConstScene, ga), np);
    put(sv.gb, offsetof(ConstScene, gb), np);
    put(sv.gc, offsetof(ConstScene, gc), np);
    put(lv.la, offsetof(ConstScene, la), nl);
    put(lv.lb, offsetof(ConstScene, lb), nl);
    put(lv.lc, offsetof(ConstScene, lc), nl);
    put(lv.ln, offsetof(ConstScene, ln), nl);
    put(lv.le, offsetof(ConstScene, le), nl);
    put(sv.runs, offsetof(ConstScene, runs), (size_t)sv.n_runs * sizeof(int4));
    return e;
}

// The constant mirror is per device, contexts are not: re-upload when another context (or another
// scene of this context) used it last.
static inline cudaError_t persistent_bind_scene(int device, const void *ctx_key, unsigned long long serial,
                                                const SceneView &sv, const LightView &lv, cudaStream_t stream) {
    if (!scene_fits_constant(sv) || device < 0 || device >= 64) return cudaSuccess;
    ConstOwner &o = g_const_owner[device];
    if (o.key == ctx_key && o.serial == serial) return cudaSuccess;
    cudaError_t e = cudaDeviceSynchronize();            // nobody may still be reading the old mirror
    if (e != cudaSuccess) return e;
    e = persistent_scene_upload(sv, lv, stream);
    if (e == cudaSuccess) { o.key = ctx_key; o.serial = serial; }
    return e;
}

#ifndef TRT_NATIVE_TU      // the strict launchers instantiate conv 0 / 1 only in trt_api.cu
static inline cudaError_t persistent_trace_launch(const SceneView &sv, const LightView &lv, const FrameParams &fp, int conv,
                                                  int sm_count, cudaStream_t stream, uint32_t *launches) {
    cudaError_t e = cudaMemsetAsync(fp.work_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const uint32_t want = (fp.n_local_pixels + TRT_PT_THREADS - 1u) / TRT_PT_THREADS;
    const uint32_t cap = (uint32_t)sm_count * TRT_PT_CTAS_PER_SM;
    const uint32_t blocks = want < cap ? want : cap;
    const bool small = scene_fits_constant(sv);
    if (small) {
        if (conv) trace_persistent_kernel<1, SrcConst><<<blocks, TRT_PT_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_persistent_kernel<0, SrcConst><<<blocks, TRT_PT_THREADS, 0, stream>>>(sv, lv, fp);
    } else {
        if (conv) trace_persistent_kernel<1, SrcGlobal><<<blocks, TRT_PT_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_persistent_kernel<0, SrcGlobal><<<blocks, TRT_PT_THREADS, 0, stream>>>(sv, lv, fp);
    }
    *launches += 1;
    return cudaGetLastError();
}

#endif

}  // namespace trt
