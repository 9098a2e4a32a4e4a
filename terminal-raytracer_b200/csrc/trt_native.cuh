// trt_native.cuh -- production trace kernel of the device-native convention (conv 2): persistent warps,
// one path per lane, and DEFERRED shadow rays.
//
// The wave kernel (trt_wave.cuh) gives every lane one scan slot per iteration; a lane whose bounce
// needs a shadow ray spends its next slot on that ray and only then goes on, so that 28 % of all
// lane-iterations of the Cornell box are shadow iterations in which the lane skips most of the shading
// code the warp executes anyway (ncu: 20.6 of 28 live lanes in the hit / light / scatter blocks,
// profiles/r1_native_wave_regions.txt).
//
// Visibility only gates an addition into the sample's radiance (shader.wgsl:402-431): it influences
// neither the random-number stream nor the path.  conv 2 is not bound to the order of floating-point
// additions, so here a lane does NOT wait for its shadow rays:
//   * light loop: a needed shadow ray is pushed, with its clamped contribution and its owner, onto a
//     per-warp FIFO in shared memory; the lane goes straight on to the scatter step and traces its next
//     closest-hit ray in the next iteration -- one full bounce per iteration for every live lane;
//   * when 32 rays have queued up the warp spends ONE iteration on them: every lane scans one queued ray
//     (full SIMT width) and adds the contribution of an unoccluded ray to its owner's "late" sum;
//   * a sample is finished (shader.wgsl:548-551) only when all the shadow rays it posted are resolved;
//     a lane whose path ends earlier waits (it is counted towards flushing the queue sooner).
//   * triangle lights are one-sided (cos_light = max(0, dot(n_light, -dir)), :404): a shading point behind
//     the light's plane receives nothing whatever point is sampled, so such a lane only advances its
//     random state by the two draws; a point in front of the plane queues the WHOLE light sample (its two
//     random numbers, the hit point and normal, colour x attenuation) and the sample point, cosines,
//     contribution and shadow ray are computed in the queue pass at full SIMT width instead of by the few
//     lanes of a warp that happen to see the light (Cornell box: 19 % of the hits).
// Per pixel the order of random draws, of samples and of bounces is the shader's; only the order in
// which a sample's light contributions are summed differs (rounding-level, inside conv 2's tolerance).
#pragma once
#include "trt_kernels.cuh"

namespace trt {

#ifndef TRT_NV_THREADS
#define TRT_NV_THREADS 128
#endif
#ifndef TRT_NV_CTAS_PER_SM
#define TRT_NV_CTAS_PER_SM 8
#endif
#define TRT_NV_RING 128          // queued shadow rays per warp (power of two)
#define TRT_NV_MAX_LIGHTS 3      // 31 left over + 32 lanes x lights per iteration must fit the ring

enum { NM_ACTIVE = 0, NM_WAIT = 1, NM_NEED_PIXEL = 2, NM_DONE = 3 };
// queue entry, 12 words.  kind 0 (shadow ray ready): origin, direction, t_max, contribution.
//                         kind 1 (triangle-light sample to evaluate): hit point, normal, r1, colour x attenuation, r2.
// meta = owner thread | light << 8 | kind << 31
enum { RQ_W0, RQ_W1, RQ_W2, RQ_W3, RQ_W4, RQ_W5, RQ_W6, RQ_W7, RQ_W8, RQ_W9, RQ_W10, RQ_META, RQ_FIELDS };

template <class SRC>
__global__ void __launch_bounds__(TRT_NV_THREADS, TRT_NV_CTAS_PER_SM)
trace_native_kernel(SceneView sc, LightView lv, FrameParams fp) {
    typedef M<2> V;
    Counters<false> cn;
    Tracer<2, false, SRC> tr(sc, lv, fp, cn);
    __shared__ float s_ring[TRT_NV_THREADS / 32][RQ_FIELDS][TRT_NV_RING];
    __shared__ float s_late[3][TRT_NV_THREADS];
    __shared__ unsigned s_res[TRT_NV_THREADS];
    const unsigned tid = threadIdx.x, lane = tid & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    float (*ring)[TRT_NV_RING] = s_ring[tid >> 5];
    volatile float *late_x = &s_late[0][tid], *late_y = &s_late[1][tid], *late_z = &s_late[2][tid];
    volatile unsigned *res = &s_res[tid];
    *late_x = 0.0f; *late_y = 0.0f; *late_z = 0.0f; *res = 0u;
    __syncwarp();

    uint32_t head = 0, tail = 0;                 // FIFO counters, identical in every lane of the warp
    uint32_t pend = 0;                           // shadow rays this lane has posted so far
    int mode = NM_NEED_PIXEL;
    uint32_t xy = 0, rng = 0, si = 0, target = 0, depth = 0;
    bool extra_phase = false;
    float3 pixel_color = f3(0, 0, 0), sq_sum = f3(0, 0, 0);
    float3 att = f3(1, 1, 1), acc = f3(0, 0, 0);
    float3 ro = f3(0, 0, 0), rd = f3(0, 0, 1);   // this lane's closest-hit ray
    const bool small = SRC::kSmall && c_small.ok;

    for (;;) {
        // ---- finish the sample of a lane whose path has ended and whose shadow rays are all resolved ------
        // (shader.wgsl:548-551, :555-565, :596-621)
        if (mode == NM_WAIT && *res == pend) {
            acc = V::add(acc, f3(*late_x, *late_y, *late_z));
            *late_x = 0.0f; *late_y = 0.0f; *late_z = 0.0f;
            tr.n_samples++;
            pixel_color = V::add(pixel_color, acc);
            if (!extra_phase) sq_sum = V::add(sq_sum, V::mulv(acc, acc));
            si++;
            bool finished = false;
            float variance_value = 0.0f;
            if (si == target) {
                finished = true;
                if (!extra_phase) {
                    const uint32_t additional = tr.adaptive_extra(pixel_color, sq_sum, variance_value);
                    sq_sum.x = variance_value;                 // kept for the variance store of an adaptive pixel
                    if (variance_value > 10.0f && fp.base_samples < fp.spp) {
                        extra_phase = true;
                        target = fp.base_samples + additional;
                        finished = (additional == 0u);
                    }
                } else {
                    variance_value = sq_sum.x;
                }
            }
            if (finished) {
                if (extra_phase) pixel_color = V::divs(V::scale(pixel_color, fp.spp_f), __uint2float_rn(target));
                tr.finish_pixel(xy & 0xffffu, xy >> 16, pixel_color, variance_value);
                mode = NM_NEED_PIXEL;
            } else {                                              // next sample (shader.wgsl:517 / :565)
                rng = pcg_hash(rng + si * 5096u);
                tr.camera_ray(rng, xy & 0xffffu, xy >> 16, ro, rd);
                att = f3(1.0f, 1.0f, 1.0f);
                acc = f3(0.0f, 0.0f, 0.0f);
                depth = 0;
                mode = NM_ACTIVE;
            }
        }

        // ---- hand out pixels: one atomic per warp per round --------------------------------------------------
        const unsigned need = __ballot_sync(0xffffffffu, mode == NM_NEED_PIXEL);
        if (need) {
            const int leader = __ffs(need) - 1;
            unsigned base = 0;
            if ((int)lane == leader) base = atomicAdd(fp.work_counter, __popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (mode == NM_NEED_PIXEL) {
                const uint32_t slot = base + __popc(need & lt_mask);
                if (slot < fp.n_local_pixels) {
                    const uint32_t lp = fp.order ? fp.order[slot] : slot;
                    const uint32_t lr = lp / fp.width;
                    const uint32_t x = lp - lr * fp.width;
                    const uint32_t y = fp.row_map ? fp.row_map[lr] : lr;
                    xy = x | (y << 16);
                    rng = (y * fp.width + x) * 1973u + fp.seed * 9277u + fp.frame_number * 12345u;   // shader.wgsl:507
                    pixel_color = f3(0, 0, 0);
                    sq_sum = f3(0, 0, 0);
                    si = 0;
                    target = fp.base_samples;
                    extra_phase = false;
                    rng = pcg_hash(rng);                         // first sample of the pixel (shader.wgsl:517-546)
                    tr.camera_ray(rng, x, y, ro, rd);
                    att = f3(1.0f, 1.0f, 1.0f);
                    acc = f3(0.0f, 0.0f, 0.0f);
                    depth = 0;
                    mode = NM_ACTIVE;
                } else {
                    mode = NM_DONE;
                }
            }
        }
        if (__all_sync(0xffffffffu, mode == NM_DONE)) break;

        // ---- this iteration's scan: the queued shadow rays (when enough have queued up, or lanes are waiting
        // for them) or every lane's own closest-hit ray ------------------------------------------------------------
        const unsigned m_active = __ballot_sync(0xffffffffu, mode == NM_ACTIVE);
        const unsigned m_wait = __ballot_sync(0xffffffffu, mode == NM_WAIT);
        const uint32_t cnt = tail - head;
        const bool spass = cnt >= 32u || (cnt > 0u && (cnt + 2u * __popc(m_wait) >= 32u || !m_active));
        float3 qo = ro, qd = rd, qc = f3(0, 0, 0);
        float qt = 1e10f;
        unsigned qowner = 0;
        bool qneed = false;
        const uint32_t qn = cnt < 32u ? cnt : 32u;
        const uint32_t qe = (head + lane) & (TRT_NV_RING - 1u);
        const bool qvalid = spass && lane < qn;
        if (qvalid) {
            const unsigned meta = __float_as_uint(ring[RQ_META][qe]);
            qowner = meta & 0xffu;
            const float3 w0 = f3(ring[RQ_W0][qe], ring[RQ_W1][qe], ring[RQ_W2][qe]);
            const float3 w3 = f3(ring[RQ_W3][qe], ring[RQ_W4][qe], ring[RQ_W5][qe]);
            const float w6 = ring[RQ_W6][qe];
            const float3 w7 = f3(ring[RQ_W7][qe], ring[RQ_W8][qe], ring[RQ_W9][qe]);
            if (meta >> 31) {                                     // triangle-light sample :370-431, evaluated here
                const int k = (int)((meta >> 8) & 0xffu);
                const float r2 = ring[RQ_W10][qe];
                const float4 nl = tr.src.ln(k);
                const float4 em = tr.src.le(k);
                const float sqrt_r1 = sqrt_native(w6);
                const float u = 1.0f - sqrt_r1;
                const float v = r2 * sqrt_r1;
                const float3 light_point = V::lin3(xyz(tr.src.la(k)), 1.0f - u - v, xyz(tr.src.lb(k)), u, xyz(tr.src.lc(k)), v);
                const float3 light_vec = V::sub(light_point, w0);
                const float sl = V::dot(xyz(nl), light_vec), ss = V::dot(w3, light_vec);
                qo = V::axpy(w0, w3, 0.001f);                     // :394-399
                qt = -1.0f;                                       // nothing to find unless the sample is needed
                if (ss > 0.0f && sl < 0.0f) {                     // both cosines positive (:403-404)
                    qneed = true;
                    const float d2 = V::dot(light_vec, light_vec);
                    const float inv_len = rsqrt_native(d2);
                    qd = V::scale(light_vec, inv_len);
                    qt = d2 * inv_len - 0.001f;
                    const float geometric = ((ss * inv_len) * -(sl * inv_len)) * rcp_native(d2);
                    const float3 cc = V::mulv(V::scale(w7, 0.31830988618f), V::scale(xyz(em), geometric * nl.w));
                    qc = f3(fminf(cc.x, 10.0f), fminf(cc.y, 10.0f), fminf(cc.z, 10.0f));
                }
            } else {
                qo = w0; qd = w3; qt = w6; qc = w7;
                qneed = true;
            }
        }
        int best = -1;
        float t_hit = -1.0f;
        if (fp.max_depth != 0u) t_hit = small ? tr.scan_small_native(qo, qd, qt, best) : tr.scan_native(qo, qd, qt, best);

        if (spass) {                                              // shader.wgsl:402-431 for the queued rays
            if (qvalid) {
                if (t_hit < 0.0f && qneed) {
                    atomicAdd(&s_late[0][qowner], qc.x);
                    atomicAdd(&s_late[1][qowner], qc.y);
                    atomicAdd(&s_late[2][qowner], qc.z);
                }
                atomicAdd(&s_res[qowner], 1u);
            }
            head += qn;
            __syncwarp();
            continue;
        }

        // ---- closest hit: miss -> sky, hit -> record (shader.wgsl:447-459, :280-290) ----------------------------
        bool path_end = false, hit = false;
        float3 hp = f3(0, 0, 0), hn = f3(0, 1, 0);
        float4 mc = make_float4(0, 0, 0, 0);
        if (mode == NM_ACTIVE) {
            if (fp.max_depth == 0u) {
                path_end = true;
            } else {
                tr.n_bounces++;
                if (t_hit < 0.0f) {
                    acc = V::vmad(acc, tr.sky(rd), att);
                    path_end = true;
                } else {
                    hit = true;
                    mc = __ldg(&sc.mc[best]);
                    const float4 me = __ldg(&sc.me[best]);
                    hp = V::axpy(ro, rd, t_hit);
                    hn = tr.normal_at(best, __float_as_int(me.w), hp);
                    if (!(V::dot(rd, hn) < 0.0f)) hn = V::scale(hn, -1.0f);
                    acc = V::vmad(acc, xyz(me), att);             // shader.wgsl:455
                }
            }
        }

        // ---- next-event estimation (shader.wgsl:342-433): every light in order, two draws each; a needed
        // shadow ray goes to the queue with its clamped contribution ----------------------------------------------
        for (int k = 0; k < sc.n_lights; k++) {
            const float4 a = tr.src.la(k);
            const float4 nl = tr.src.ln(k);
            const float4 em = tr.src.le(k);
            bool post = false;
            float3 w0 = f3(0, 0, 0), w3 = f3(0, 0, 1), w7 = f3(0, 0, 0);
            float w6 = 0.0f, w10 = 0.0f;
            unsigned kind = 0;
            if (__float_as_int(em.w) == 0) {                      // sphere light :352-369 -> a ready shadow ray
                if (hit) {
                    const float r1 = rnd(rng), r2 = rnd(rng);
                    const float cos_theta = 1.0f - 2.0f * r1;
                    const float sin_theta = sqrt_native(1.0f - cos_theta * cos_theta);
                    float sp, cp;
                    sincos_turn<2>(r2, sp, cp);
                    const float3 local = f3(sin_theta * cp, sin_theta * sp, cos_theta);
                    const float3 light_vec = V::sub(V::axpy(xyz(a), local, a.w), hp);
                    const float sl = V::dot(local, light_vec), ss = V::dot(hn, light_vec);
                    if (ss > 0.0f && sl < 0.0f) {                 // both cosines positive (:403-404)
                        const float d2 = V::dot(light_vec, light_vec);
                        const float inv_len = rsqrt_native(d2);
                        w0 = V::axpy(hp, hn, 0.001f);             // :394-399
                        w3 = V::scale(light_vec, inv_len);
                        w6 = d2 * inv_len - 0.001f;
                        const float geometric = ((ss * inv_len) * -(sl * inv_len)) * rcp_native(d2);
                        const float3 brdf = V::scale(xyz(mc), 0.31830988618f);
                        const float3 cc = V::mulv(V::mulv(brdf, xyz(em)), V::scale(att, geometric * nl.w));
                        w7 = f3(fminf(cc.x, 10.0f), fminf(cc.y, 10.0f), fminf(cc.z, 10.0f));
                        post = true;
                    }
                }
            } else {                                              // triangle light :370-389 -> queue the sample
                // plane offset of the light and a margin for the roundings of the sample point (a convex
                // combination of the vertices) and of the two dot products
                const float c_l = V::dot(xyz(nl), xyz(a));
                if (hit) {
                    const uint32_t s1 = pcg_hash(rng);
                    rng = pcg_hash(s1);                           // the two draws of :371-372, in order
                    const float side = V::dot(xyz(nl), hp) - c_l;
                    const float eps = 1e-5f * (1.0f + fabsf(c_l) + fabsf(hp.x) + fabsf(hp.y) + fabsf(hp.z));
                    if (side > -eps) {                            // in front of the light's plane: it may shine here
                        w0 = hp; w3 = hn;
                        w6 = __uint2float_rn(s1) * 2.3283064365386963e-10f;
                        w10 = __uint2float_rn(rng) * 2.3283064365386963e-10f;
                        w7 = V::mulv(xyz(mc), att);
                        kind = 1u;
                        post = true;
                    }
                }
            }
            const unsigned mp = __ballot_sync(0xffffffffu, post);
            if (mp) {
                if (post) {
                    const uint32_t e = (tail + __popc(mp & lt_mask)) & (TRT_NV_RING - 1u);
                    ring[RQ_W0][e] = w0.x; ring[RQ_W1][e] = w0.y; ring[RQ_W2][e] = w0.z;
                    ring[RQ_W3][e] = w3.x; ring[RQ_W4][e] = w3.y; ring[RQ_W5][e] = w3.z;
                    ring[RQ_W6][e] = w6;
                    ring[RQ_W7][e] = w7.x; ring[RQ_W8][e] = w7.y; ring[RQ_W9][e] = w7.z;
                    ring[RQ_W10][e] = w10;
                    ring[RQ_META][e] = __uint_as_float(tid | ((unsigned)k << 8) | (kind << 31));
                    pend++;
                }
                tail += __popc(mp);
            }
        }
        __syncwarp();

        // ---- scatter + roulette (shader.wgsl:459-492) --------------------------------------------------------------
        if (hit) {
            float3 s;
            if (mc.w > rnd(rng)) s = V::reflect(rd, hn);
            else s = tr.cosine_hemisphere(rng, hn);
            att = V::mulv(att, xyz(mc));
            ro = V::axpy(hp, s, 0.001f);
            rd = s;
            if (depth > 3u) {
                const float survival = fminf(0.95f, fmaxf(att.x, fmaxf(att.y, att.z)));
                if (survival < rnd(rng)) path_end = true;
                else att = V::divs(att, survival);
            }
            depth++;
            if (depth >= fp.max_depth) path_end = true;
        }
        if (path_end) mode = NM_WAIT;
    }
    tr.flush_stats();
}

}  // namespace trt
