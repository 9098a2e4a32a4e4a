// trt_wave.cuh -- production trace kernel (v2): persistent warps, one path state per lane, and ONE
// shared scan slot per loop iteration.
//
// v1 (trt_persistent.cuh) ran, per bounce, the closest-hit scan and then one shadow scan per light
// inside the same iteration.  ncu (profiles/r1_trace_v1_*.txt) showed the consequence: the shadow
// scans are needed by few lanes (Cornell: 0.38 per bounce) but were executed by the whole warp every
// bounce, so two thirds of all scan instructions issued with ~5 of 32 lanes active (13.7 active
// lanes per instruction overall).
//
// Here every lane owns exactly one "scan ray" per iteration -- either its path's closest-hit ray or a
// pending shadow ray -- and the warp runs a single linear scan over the primitives for all 32 rays
// (hit_scene, shader.wgsl:248-294, serves both: for a shadow ray only `t < 0` is read, :400-402, and
// narrowing the bound never changes that truth value).  After the scan a lane continues where it was:
//   closest hit  -> emission, then the light loop from light 0
//   shadow ray k -> add its (precomputed, clamped) contribution if unoccluded, light loop from k + 1
//   light loop   -> sample light k (2 draws, in order); if both cosines are positive the lane posts a
//                   shadow ray and waits for the next scan, else it goes on to light k + 1
//   lights done  -> scatter, roulette, next closest ray (or next sample / next pixel)
// A bounce therefore takes 1 + (#needed shadow scans) iterations of that lane, and every scan runs
// with all live lanes.  The per-pixel order of random draws, of samples and of floating-point
// additions is exactly the shader's, so the result is bit-identical to the per-pixel loop nest.
#pragma once
#include "trt_kernels.cuh"

namespace trt {

#ifndef TRT_WV_THREADS
#define TRT_WV_THREADS 128
#endif
#ifndef TRT_WV_CTAS_PER_SM
#define TRT_WV_CTAS_PER_SM 6
#endif

enum { WM_CLOSEST = 0, WM_SHADOW = 1, WM_NEED_PIXEL = 2, WM_DONE = 3 };

// BATCH: the frame-batch form (work item = (frame, pixel), trt_render_frames); the single-frame instantiation carries
// neither the extra state nor the branches
template <int CONV, class SRC, bool BATCH>
__global__ void __launch_bounds__(TRT_WV_THREADS, TRT_WV_CTAS_PER_SM)
trace_wave_kernel(SceneView sc, LightView lv, FrameParams fp) {
    typedef M<CONV> V;
    Counters<false> cn;
    Tracer<CONV, false, SRC> tr(sc, lv, fp, cn);
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;

    int mode = WM_NEED_PIXEL;
    uint32_t xy = 0, rng = 0, si = 0, target = 0, depth = 0, fslot = 0;
    const uint32_t n_work = BATCH ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    bool extra_phase = false;
    float variance_value = 0.0f;
    float3 pixel_color = f3(0, 0, 0), sq_sum = f3(0, 0, 0);
    float3 att = f3(1, 1, 1), acc = f3(0, 0, 0);
    // the scan ray of this iteration
    float3 ro = f3(0, 0, 0), rd = f3(0, 0, 1);
    float t_max = 0.0f;
    // shading context kept while shadow rays of this bounce are pending
    float3 hp = f3(0, 0, 0), hn = f3(0, 1, 0), rd_in = f3(0, 0, 1), direct = f3(0, 0, 0), contrib = f3(0, 0, 0);
    float4 mc = make_float4(0, 0, 0, 0);
    int k = 0;

    for (;;) {
        // ---- hand out pixels: one atomic per warp per round (all lanes converge here) ----------------
        const unsigned need = __ballot_sync(0xffffffffu, mode == WM_NEED_PIXEL);
        if (need) {
            const int leader = __ffs(need) - 1;
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(fp.work_counter, __popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (mode == WM_NEED_PIXEL) {
                const uint32_t slot = base + __popc(need & lt_mask);
                if (slot < n_work) {
                    uint32_t ls = slot, seed = fp.seed, frame_number = fp.frame_number;
                    if (BATCH) {                                 // frame batch: work item = (frame, local pixel)
                        const uint32_t f = slot / fp.n_local_pixels;
                        ls = slot - f * fp.n_local_pixels;
                        const uint2 b = __ldg(&fp.batch[f]);
                        seed = b.x; frame_number = b.y;
                    }
                    const uint32_t lp = fp.order ? fp.order[ls] : ls;
                    const uint32_t lr = lp / fp.width;
                    const uint32_t x = lp - lr * fp.width;
                    const uint32_t y = fp.row_map ? fp.row_map[lr] : lr;
                    xy = x | (y << 16);
                    if (BATCH) fslot = slot - ls + lp;
                    rng = (y * fp.width + x) * 1973u + seed * 9277u + frame_number * 12345u;   // shader.wgsl:507
                    pixel_color = f3(0, 0, 0);
                    sq_sum = f3(0, 0, 0);
                    si = 0;
                    target = fp.base_samples;
                    extra_phase = false;
                    // first sample of the pixel (shader.wgsl:517-546)
                    rng = pcg_hash(rng);
                    tr.camera_ray(rng, x, y, ro, rd);
                    t_max = 1e10f;
                    att = f3(1.0f, 1.0f, 1.0f);
                    acc = f3(0.0f, 0.0f, 0.0f);
                    depth = 0;
                    mode = WM_CLOSEST;
                } else {
                    mode = WM_DONE;
                }
            }
        }
        if (__all_sync(0xffffffffu, mode == WM_DONE)) break;

        // ---- the scan slot: hit_scene for this lane's ray (closest-hit or shadow alike) -------------------
        int best = -1;
        float t_hit = -1.0f;
        if (CONV == 2) {
            // every lane scans (finished lanes keep a stale ray and ignore the result): uniform control flow
            if (fp.max_depth != 0u)
                t_hit = SRC::kGrid ? tr.scan_grid(ro, rd, t_max, best)
                        : (SRC::kSmall && c_small.ok) ? tr.template scan_small_native<ShapeAny>(ro, rd, t_max, best) : tr.scan_native(ro, rd, t_max, best);
        } else if (mode != WM_DONE && fp.max_depth != 0u) {
            t_hit = tr.scan_lean(ro, rd, t_max, best);
        }

        __syncwarp();
        bool path_end = false;
        bool lights = false;
        if (mode == WM_CLOSEST) {
            if (fp.max_depth == 0u) {
                path_end = true;                                  // no bounce at all: the sample is black
            } else {
                tr.n_bounces++;
                if (t_hit < 0.0f) {                               // miss: sky (shader.wgsl:447-452)
                    acc = V::vmad(acc, tr.sky(rd), att);
                    path_end = true;
                } else {                                          // hit: complete the record (shader.wgsl:280-290)
                    mc = __ldg(&sc.mc[best]);
                    const float4 me = __ldg(&sc.me[best]);
                    hp = V::axpy(ro, rd, t_hit);
                    hn = tr.normal_at(best, __float_as_int(me.w), hp);
                    if (!(V::dot(rd, hn) < 0.0f)) hn = V::scale(hn, -1.0f);
                    acc = V::vmad(acc, xyz(me), att);             // shader.wgsl:455
                    rd_in = rd;
                    direct = f3(0.0f, 0.0f, 0.0f);
                    k = 0;
                    lights = true;
                }
            }
        } else if (mode == WM_SHADOW) {                           // shader.wgsl:402-431 for light k
            if (t_hit < 0.0f) direct = V::add(direct, contrib);
            k++;
            lights = true;
        }

        // ---- light loop (shader.wgsl:342-433), resumable at light k --------------------------------------------
        // Warp-uniform trip structure: lanes sit at different k, every trip serves all lanes that still
        // have a light to sample, and the warp reconverges before each trip and before the scatter step
        // (without the votes the compiler lets early leavers run ahead and the scatter code then
        // executes once per leaving group: profiles/r1_trace_wave2_by_line.txt).
        if (lights) mode = WM_CLOSEST;
        for (;;) {
            const bool in_loop = lights && mode == WM_CLOSEST && k < sc.n_lights;
            if (!__any_sync(0xffffffffu, in_loop)) break;
            if (in_loop) {
                const float4 a = tr.src.la(k);
                const float4 nl = tr.src.ln(k);
                const float4 em = tr.src.le(k);
                float3 light_point, light_normal;
                const float r1 = rnd(rng), r2 = rnd(rng);
                if (__float_as_int(em.w) == 0) {                  // sphere light :352-369
                    const float cos_theta = 1.0f - 2.0f * r1;
                    const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
                    float sp, cp;
                    sincos_turn<CONV>(r2, sp, cp);
                    const float3 local = f3(sin_theta * cp, sin_theta * sp, cos_theta);
                    light_point = V::axpy(xyz(a), local, a.w);
                    light_normal = local;
                } else {                                          // triangle light :370-389
                    const float sqrt_r1 = sqrtf(r1);
                    const float u = 1.0f - sqrt_r1;
                    const float v = r2 * sqrt_r1;
                    light_point = V::lin3(xyz(a), 1.0f - u - v, xyz(tr.src.lb(k)), u, xyz(tr.src.lc(k)), v);
                    light_normal = xyz(nl);
                }
                const float3 light_vec = V::sub(light_point, hp);
                // Early out, exact: if dot(n_light, light_vec) is positive (or dot(n_surface, light_vec)
                // negative) by more than the rounding of the normalisation that follows, the matching
                // cosine (:403-404) is certainly 0 and this light adds nothing (DESIGN.md "Light loop").
                const float d2 = V::dot(light_vec, light_vec);
                const float sl = V::dot(light_normal, light_vec), ss = V::dot(hn, light_vec);
                float cos_surface = 0.0f, cos_light = 0.0f, light_distance = 0.0f;
                float3 light_dir = f3(0.0f, 0.0f, 0.0f);
                if (CONV == 2) {                                  // native: the cosines' signs are those of ss and -sl
                    if (ss > 0.0f && sl < 0.0f) {
                        const float inv_len = rsqrt_native(d2);
                        light_distance = d2 * inv_len;
                        light_dir = V::scale(light_vec, inv_len);
                        cos_surface = ss * inv_len;
                        cos_light = -(sl * inv_len);
                    }
                } else if (!((sl > 0.0f && sl * sl > 1e-11f * d2) || (ss < 0.0f && ss * ss > 1e-11f * d2))) {
                    light_distance = sqrtf(d2);
                    light_dir = V::divs(light_vec, light_distance);
                    cos_surface = fmaxf(0.0f, V::dot(hn, light_dir));
                    cos_light = fmaxf(0.0f, V::dot(light_normal, V::scale(light_dir, -1.0f)));
                }
                if (cos_surface > 0.0f && cos_light > 0.0f) {
                    // contribution if unoccluded (:408-427); the shadow ray goes to the next scan (:399-400)
                    const float geometric = (cos_surface * cos_light) / (light_distance * light_distance);
                    const float3 brdf = V::divs(xyz(mc), TRT_PI_F);
                    const float3 c = V::mulv(V::mulv(brdf, xyz(em)), V::scale(att, geometric * nl.w));
                    contrib = f3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
                    ro = V::axpy(hp, hn, 0.001f);
                    rd = light_dir;
                    t_max = light_distance - 0.001f;
                    mode = WM_SHADOW;
                } else {
                    k++;                                          // a scan that cannot change the sum is skipped
                }
            }
        }
        const bool scatter = lights && mode == WM_CLOSEST;
        __syncwarp();

        // ---- scatter + roulette (shader.wgsl:459-492) ---------------------------------------------------------------
        if (scatter) {
            acc = V::add(acc, direct);
            float3 s;
            if (mc.w > rnd(rng)) s = V::reflect(rd_in, hn);
            else s = tr.cosine_hemisphere(rng, hn);
            att = V::mulv(att, xyz(mc));
            ro = V::axpy(hp, s, 0.001f);
            rd = s;
            t_max = 1e10f;
            if (depth > 3u) {
                const float survival = fminf(0.95f, fmaxf(att.x, fmaxf(att.y, att.z)));
                if (survival < rnd(rng)) path_end = true;
                else att = V::divs(att, survival);
            }
            depth++;
            if (depth >= fp.max_depth) path_end = true;
        }

        // ---- sample / pixel bookkeeping (shader.wgsl:548-551, :555-565, :596-621) --------------------------------------
        __syncwarp();
        if (path_end) {
            tr.n_samples++;
            pixel_color = V::add(pixel_color, acc);
            if (!extra_phase) sq_sum = V::add(sq_sum, V::mulv(acc, acc));
            si++;
            bool finished = false;
            if (si == target) {
                finished = true;
                if (!extra_phase) {
                    const uint32_t additional = tr.adaptive_extra(pixel_color, sq_sum, variance_value);
                    if (variance_value > 10.0f && fp.base_samples < fp.spp) {
                        extra_phase = true;
                        target = fp.base_samples + additional;
                        finished = (additional == 0u);
                    }
                }
            }
            if (finished) {
                if (extra_phase) pixel_color = V::divs(V::scale(pixel_color, fp.spp_f), __uint2float_rn(target));
                tr.finish_pixel(xy & 0xffffu, xy >> 16, pixel_color, variance_value, BATCH ? fslot : 0u);
                mode = WM_NEED_PIXEL;
            } else {                                              // next sample (shader.wgsl:517 / :565)
                rng = pcg_hash(rng + si * 5096u);
                tr.camera_ray(rng, xy & 0xffffu, xy >> 16, ro, rd);
                t_max = 1e10f;
                att = f3(1.0f, 1.0f, 1.0f);
                acc = f3(0.0f, 0.0f, 0.0f);
                depth = 0;
                mode = WM_CLOSEST;
            }
        }
    }
    tr.flush_stats();
}

#ifndef TRT_NATIVE_TU      // the strict launchers instantiate conv 0 / 1 only in trt_api.cu
static inline cudaError_t wave_trace_launch(const SceneView &sv, const LightView &lv, const FrameParams &fp, int conv,
                                            int sm_count, bool small, cudaStream_t stream, uint32_t *launches) {
    cudaError_t e = cudaMemsetAsync(fp.work_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    const uint32_t n_work = fp.batch ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    const uint32_t want = (n_work + TRT_WV_THREADS - 1u) / TRT_WV_THREADS;
    // resident CTAs per SM: the caller's hint for short frames (trt_api.cu frame_ctas_hint), else all that fit
    const uint32_t ctas = fp.ctas_hint >= 1u && fp.ctas_hint < TRT_WV_CTAS_PER_SM ? fp.ctas_hint : TRT_WV_CTAS_PER_SM;
    const uint32_t cap = (uint32_t)sm_count * ctas;
    const uint32_t blocks = want < cap ? want : cap;
#define TRT_WV_LAUNCH(CONV_, SRC_)                                                                                  \
    do {                                                                                                            \
        if (fp.batch) trace_wave_kernel<CONV_, SRC_, true><<<blocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);      \
        else trace_wave_kernel<CONV_, SRC_, false><<<blocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);              \
    } while (0)
    if (small) {
        if (conv) TRT_WV_LAUNCH(1, SrcConst);
        else TRT_WV_LAUNCH(0, SrcConst);
    } else {
        if (conv) TRT_WV_LAUNCH(1, SrcGlobal);
        else TRT_WV_LAUNCH(0, SrcGlobal);
    }
#undef TRT_WV_LAUNCH
    *launches += 1;
    return cudaGetLastError();
}

#endif

}  // namespace trt
