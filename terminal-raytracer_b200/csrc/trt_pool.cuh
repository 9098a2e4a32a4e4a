// trt_pool.cuh -- the deferred-NEE machine of trt_native.cuh with REGENERATION AT SIMT WIDTH: a pool of pixel
// chains per warp.
//
// In trace_native_kernel a lane is a pixel: when its path ends the lane closes the sample, draws the next camera ray
// (or takes the next pixel) and goes on.  Paths end at different times, so that block -- 104 of the ~690 warp
// instructions of an iteration -- runs in almost every iteration with 2.4-4.6 of 32 lanes (profiles/
// r2_native_regions.txt: 15 % of all issued instructions).  A sample cannot be prepared ahead of time either: its
// seed is the random state its predecessor ENDS with (shader.wgsl:517, :565).
//
// Here a pixel is a CHAIN -- the serial sequence of its samples -- and a warp owns more chains than lanes
// (TRT_PL_CHAINS, in shared memory).  A lane whose path ends drops its chain on the warp's DEAD list (radiance so
// far, random state: 5 stores) and picks up a chain from the READY list (a prepared camera ray: 4 loads), so it
// traces again in the very next iteration.  When the dead list holds a warp's width of closable chains -- or lanes
// would go without a ray -- ONE pass closes their samples, runs the adaptive-sampling / pixel bookkeeping
// (shader.wgsl:548-621), hands out pixels and generates the camera rays with every lane busy, and the chains move to
// the ready list.  Nothing but the order in which INDEPENDENT pixels advance changes: per pixel the samples, the
// random draws and every floating-point operation are those of trace_native_kernel in the same convention (strict
// builds: bit-exact against the oracle; conv 2: every sample is closed whole, in the strict machine's order).
//
//   * a dead chain is closable when the light samples it posted are resolved: the queue is a FIFO, so that is
//     `head >= the queue's tail when the path ended` -- no per-owner counters, no parking of samples;
//   * a chain is live about (lanes / chains) of the time, so a pixel takes longer in wall-clock terms: once the work
//     list is dry the pass runs as soon as any chain can be closed, and frames with fewer pixels than the pool
//     would hold go to trace_native_kernel (host side: pool_wanted).
#pragma once
#include "trt_native.cuh"

namespace trt {

#ifndef TRT_PL_CHAINS
#define TRT_PL_CHAINS 64          // chains per warp (32 live + the dead and ready lists)
#endif
#ifndef TRT_PL_CTAS_PER_SM
#define TRT_PL_CTAS_PER_SM 7      // conv 2: ring 11.3 KB + chains 19 words x 64 x 4 warps = 19.5 KB per CTA
#endif
#ifndef TRT_PL_STRICT_CTAS
#define TRT_PL_STRICT_CTAS 5
#endif

enum { PM_ACTIVE = 0, PM_IDLE = 1, PM_ENDED = 2 };   // ENDED: the path ended in the last iteration, the chain is still the lane's
// chain record (words; odd stride)
enum { CH_L0, CH_L1, CH_L2,      // the late sum of the open sample (zeroed by the regeneration pass)
       CH_A0, CH_A1, CH_A2,      // dead: the ended sample's radiance (the lane's acc); ready: direction of the prepared camera ray
       CH_TAD,                   // dead: the queue's tail when the path ended
       CH_RNG,                   // dead: state the path ended with; ready: state after the camera ray's draws
       CH_PC0, CH_PC1, CH_PC2,   // the pixel's colour sum (:549)
       CH_SQ0, CH_SQ1, CH_SQ2,   // the pixel's sum of squares (:550-551); .x = the variance value in the extra phase
       CH_XY, CH_SI,             // pixel, index of the open sample
       CH_TGT,                   // samples of the current phase | CHF_* flags
       CH_FSLOT, CH_NB,          // output slot; bounces of this pixel so far (-> fp.cost)
       CH_WORDS };
static_assert(CH_WORDS % 2 == 1, "odd stride");
#define CHF_EXTRA 0x80000000u    // adaptive extra phase (:565-602)
#define CHF_NEW 0x40000000u      // no pixel yet

template <int CONV, class SRC, class SHAPE>
__device__ __forceinline__ void trace_pool_body(const SceneView &sc, const LightView &lv, const FrameParams &fp) {
    typedef M<CONV> V;
    static_assert(!SRC::kTiled && !SRC::kGrid, "small scenes only");
    constexpr bool kStrict = CONV != 2;
    constexpr int kFields = kStrict ? (int)DQ_FIELDS_STRICT : (int)DQ_FIELDS;
    constexpr unsigned NCH = TRT_PL_CHAINS, NW = TRT_NV_THREADS / 32;
    static_assert(NCH >= 32 && NCH <= 256, "chain ids travel in 8 bits");
    Counters<false> cn;
    Tracer<CONV, false, SRC> tr(sc, lv, fp, cn);
    __shared__ float s_ring[NW][kFields][TRT_NV_RING];
    __shared__ float s_ch[NW * NCH * CH_WORDS];
    __shared__ unsigned char s_dead[NW][NCH], s_ready[NW][NCH];
    const unsigned tid = threadIdx.x, lane = tid & 31u, warp = tid >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    float (*ring)[TRT_NV_RING] = s_ring[warp];
    unsigned char *dead = s_dead[warp], *ready = s_ready[warp];
    const unsigned ch_base = (unsigned)__cvta_generic_to_shared(s_ch) + 4u * warp * NCH * CH_WORDS;
#define TRT_CH(c, f) (ch_base + 4u * ((c) * (unsigned)CH_WORDS + (unsigned)(f)))
    for (unsigned c = lane; c < NCH; c += 32u) {
#pragma unroll
        for (unsigned f = 0; f < CH_WORDS; f++) sts_u(TRT_CH(c, f), 0u);
        sts_u(TRT_CH(c, CH_TGT), CHF_NEW);
        dead[c] = (unsigned char)c;
    }
    __syncwarp();

    // warp-uniform state: the queue, the two lists (ring indices < NCH and counts), whether the work list is dry
    uint32_t head = 0, tail = 0;
    uint32_t dhead = 0, dcount = NCH, rhead = 0, rcount = 0;
    uint32_t closable = NCH;                     // dead chains (a prefix of the list) known to be closable; exact after every queue / regeneration pass
    bool dry = false;
    int sp_k = -1;
    const uint32_t n_work = fp.batch ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    // per-lane state: the live chain and its path
    int mode = PM_IDLE;
    unsigned cid = 0;
    uint32_t rng = 0, depth = 0, nb0 = 0;
    float3 att = f3(1, 1, 1), acc = f3(0, 0, 0);
    float3 ro = f3(0, 0, 0), rd = f3(0, 0, 1);
    float3 qsum = f3(0, 0, 0);
    const bool small = !kStrict && (SHAPE::kFixed || (SRC::kSmall && c_small.ok));

    // The dead list is in order of death and the queue is a FIFO, so the closable chains -- those whose light samples
    // are resolved: head has passed the queue's tail at their death -- are a prefix of the list (counted up to 32).
    auto count_closable = [&]() {
        const uint32_t n = dcount < 32u ? dcount : 32u;
        bool ok = false;
        if (lane < n) {
            uint32_t pos = dhead + lane;
            if (pos >= NCH) pos -= NCH;
            ok = (int)(head - lds_u(TRT_CH(dead[pos], CH_TAD))) >= 0;
        }
        const unsigned okm = __ballot_sync(0xffffffffu, ok);
        closable = okm == 0xffffffffu ? 32u : (uint32_t)__ffs((int)~okm) - 1u;
    };

    for (;;) {
        // ---- regeneration passes: close the ended samples of up to 32 dead chains, finish / hand out pixels, prepare
        // the next camera rays (shader.wgsl:548-565, :596-621, :507-546) -------------------------------------------
        const unsigned m_na = __ballot_sync(0xffffffffu, mode != PM_ACTIVE);     // lanes without a ray
        while (closable >= (dry ? 1u : 32u) || (closable != 0u && (uint32_t)__popc(m_na) > rcount)) {
            const uint32_t n_in = closable < 32u ? closable : 32u;
            unsigned c = 0;
            if (lane < n_in) {
                uint32_t pos = dhead + lane;
                if (pos >= NCH) pos -= NCH;
                c = dead[pos];
            }
            const bool in = lane < n_in;
            float3 pc = f3(0, 0, 0), sq = f3(0, 0, 0);
            uint32_t xy = 0, si = 0, tgt = CHF_NEW, crng = 0, fslot = 0, nb = 0;
            bool need_pixel = false, cont = false;
            if (in) {
                tgt = lds_u(TRT_CH(c, CH_TGT));
                need_pixel = (tgt & CHF_NEW) != 0u;
            }
            if (in && !need_pixel) {
                const float3 a0 = f3(lds_f(TRT_CH(c, CH_A0)), lds_f(TRT_CH(c, CH_A1)), lds_f(TRT_CH(c, CH_A2)));
                const float3 late = f3(lds_f(TRT_CH(c, CH_L0)), lds_f(TRT_CH(c, CH_L1)), lds_f(TRT_CH(c, CH_L2)));
                pc = f3(lds_f(TRT_CH(c, CH_PC0)), lds_f(TRT_CH(c, CH_PC1)), lds_f(TRT_CH(c, CH_PC2)));
                sq = f3(lds_f(TRT_CH(c, CH_SQ0)), lds_f(TRT_CH(c, CH_SQ1)), lds_f(TRT_CH(c, CH_SQ2)));
                xy = lds_u(TRT_CH(c, CH_XY));
                si = lds_u(TRT_CH(c, CH_SI));
                crng = lds_u(TRT_CH(c, CH_RNG));
                fslot = lds_u(TRT_CH(c, CH_FSLOT));
                nb = lds_u(TRT_CH(c, CH_NB));
                bool extra = (tgt & CHF_EXTRA) != 0u;
                uint32_t target = tgt & 0x00ffffffu;
                const float3 a = V::add(a0, late);                      // the sample's radiance (:548)
                pc = V::add(pc, a);
                if (!extra) sq = V::add(sq, V::mulv(a, a));
                si++;
                bool finished = false;
                float variance_value = 0.0f;
                if (si == target) {
                    finished = true;
                    if (!extra) {
                        const uint32_t additional = tr.adaptive_extra(pc, sq, variance_value);
                        sq.x = variance_value;                          // kept for the variance store of an adaptive pixel
                        if (variance_value > 10.0f && fp.base_samples < fp.spp) {
                            extra = true;
                            target = fp.base_samples + additional;
                            finished = (additional == 0u);
                        }
                    } else {
                        variance_value = sq.x;
                    }
                }
                if (finished) {
                    if (extra) pc = V::divs(V::scale(pc, fp.spp_f), __uint2float_rn(target));
                    tr.finish_pixel(xy & 0xffffu, xy >> 16, pc, variance_value, fslot);
                    tr.n_samples += target;
                    if (fp.cost) fp.cost[(size_t)(xy >> 16) * fp.width + (xy & 0xffffu)] = nb;
                    need_pixel = true;
                } else {
                    cont = true;
                    tgt = target | (extra ? CHF_EXTRA : 0u);
                }
            }
            // pixels: one atomic per pass
            const unsigned need = __ballot_sync(0xffffffffu, need_pixel);
            if (need) {
                const int leader = __ffs((int)need) - 1;
                unsigned base = 0;
                if ((int)lane == leader) base = atomicAdd(fp.work_counter, (unsigned)__popc(need));
                base = __shfl_sync(0xffffffffu, base, leader);
                if (base + (uint32_t)__popc(need) > n_work) dry = true;
                if (need_pixel) {
                    const uint32_t slot = base + (uint32_t)__popc(need & lt_mask);
                    if (slot < n_work) {
                        uint32_t ls = slot, seed = fp.seed, frame_number = fp.frame_number;
                        if (fp.batch) {                              // frame batch: work item = (frame, local pixel)
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
                        fslot = slot - ls + lp;
                        crng = (y * fp.width + x) * 1973u + seed * 9277u + frame_number * 12345u;   // shader.wgsl:507
                        pc = f3(0, 0, 0);
                        sq = f3(0, 0, 0);
                        si = 0;
                        nb = 0;
                        tgt = fp.base_samples;
                        cont = true;
                    }
                }
            }
            // next sample of the chain: reseed (:517 / :565), camera ray; the chain is ready
            const unsigned m_cont = __ballot_sync(0xffffffffu, cont);
            if (cont) {
                crng = pcg_hash(crng + si * 5096u);
                float3 o, d;
                tr.camera_ray(crng, xy & 0xffffu, xy >> 16, o, d);
                sts_f(TRT_CH(c, CH_A0), d.x); sts_f(TRT_CH(c, CH_A1), d.y); sts_f(TRT_CH(c, CH_A2), d.z);
                sts_f(TRT_CH(c, CH_L0), 0.0f); sts_f(TRT_CH(c, CH_L1), 0.0f); sts_f(TRT_CH(c, CH_L2), 0.0f);
                sts_u(TRT_CH(c, CH_RNG), crng);
                sts_f(TRT_CH(c, CH_PC0), pc.x); sts_f(TRT_CH(c, CH_PC1), pc.y); sts_f(TRT_CH(c, CH_PC2), pc.z);
                sts_f(TRT_CH(c, CH_SQ0), sq.x); sts_f(TRT_CH(c, CH_SQ1), sq.y); sts_f(TRT_CH(c, CH_SQ2), sq.z);
                sts_u(TRT_CH(c, CH_XY), xy);
                sts_u(TRT_CH(c, CH_SI), si);
                sts_u(TRT_CH(c, CH_TGT), tgt);
                sts_u(TRT_CH(c, CH_FSLOT), fslot);
                sts_u(TRT_CH(c, CH_NB), nb);
                uint32_t pos = rhead + rcount + (uint32_t)__popc(m_cont & lt_mask);
                if (pos >= NCH) pos -= NCH;
                if (pos >= NCH) pos -= NCH;
                ready[pos] = (unsigned char)c;
            }
            rcount += (uint32_t)__popc(m_cont);
            dhead += n_in;
            if (dhead >= NCH) dhead -= NCH;
            dcount -= n_in;
            __syncwarp();
            count_closable();
        }

        // ---- lanes without a ray: an ended path's chain joins the dead list, the lane takes a ready chain -----------
        if (m_na != 0u) {
            const unsigned m_end = __ballot_sync(0xffffffffu, mode == PM_ENDED);
            if (mode == PM_ENDED) {
                uint32_t pos = dhead + dcount + (uint32_t)__popc(m_end & lt_mask);
                if (pos >= NCH) pos -= NCH;
                if (pos >= NCH) pos -= NCH;
                dead[pos] = (unsigned char)cid;
                sts_f(TRT_CH(cid, CH_A0), acc.x); sts_f(TRT_CH(cid, CH_A1), acc.y); sts_f(TRT_CH(cid, CH_A2), acc.z);
                sts_u(TRT_CH(cid, CH_RNG), rng);
                sts_u(TRT_CH(cid, CH_TAD), tail);                 // closable once the queue's head has passed this point
                red_u(TRT_CH(cid, CH_NB), tr.n_bounces - nb0);
                mode = PM_IDLE;
            }
            dcount += (uint32_t)__popc(m_end);
            if (tail == head) closable = dcount;                  // with an empty queue a chain is closable the moment it dies
            const uint32_t k = (uint32_t)__popc(m_na & lt_mask);
            if (mode == PM_IDLE && k < rcount) {
                uint32_t pos = rhead + k;
                if (pos >= NCH) pos -= NCH;
                cid = ready[pos];
                rd = f3(lds_f(TRT_CH(cid, CH_A0)), lds_f(TRT_CH(cid, CH_A1)), lds_f(TRT_CH(cid, CH_A2)));
                rng = lds_u(TRT_CH(cid, CH_RNG));
                ro = fp.cam_pos;                                       // camera_ray: every ray starts at the eye (:546)
                att = f3(1.0f, 1.0f, 1.0f);
                acc = f3(0.0f, 0.0f, 0.0f);
                depth = 0;
                nb0 = tr.n_bounces;
                mode = PM_ACTIVE;
            }
            const uint32_t take = (uint32_t)__popc(m_na) < rcount ? (uint32_t)__popc(m_na) : rcount;
            rhead += take;
            if (rhead >= NCH) rhead -= NCH;
            rcount -= take;
            __syncwarp();
        }
        const unsigned m_active = __ballot_sync(0xffffffffu, mode == PM_ACTIVE);
        // every chain has left the pool (no pixel left for it) and nothing is queued
        if (m_active == 0u && tail == head && dcount == 0u && rcount == 0u) break;

        // ---- this iteration's scan: queued light samples (a warp's width of them, or lanes are out of work) or
        // every lane's own closest-hit ray ------------------------------------------------------------------------
        bool spass = false, qvalid = false, qneed = false;
        float3 qo = ro, qd = rd, qc = f3(0, 0, 0);
        float qt = 1e10f;                                         // (an idle lane scans its stale ray; nothing reads the result)
        unsigned qowner = 0, qrest = 0;
        uint32_t qrng = 0, qn = 0;
        const uint32_t qe = (head + lane) & (TRT_NV_RING - 1u);
        {
            const uint32_t cnt = tail - head;
            if (sp_k < 0 && cnt > 0u) {
                if (cnt >= 32u || cnt + TRT_NV_WAIT_W * (uint32_t)__popc(~m_active) >= 32u || !m_active) { sp_k = 0; qsum = f3(0, 0, 0); }
            }
            spass = sp_k >= 0;
            qn = cnt < 32u ? cnt : 32u;
            qvalid = spass && lane < qn;
            if (spass) {
                unsigned lmask = 0;
                if (qvalid) {
                    const unsigned meta = __float_as_uint(ring[DQ_META][qe]);
                    qowner = meta & 0xffu;
                    lmask = (meta >> 8) & 0xffffu;
                    qrng = __float_as_uint(ring[DQ_RNG][qe]);
                }
                const unsigned uni = __reduce_or_sync(0xffffffffu, lmask);
                int k = sp_k;
                while (uni != 0u && !((uni >> k) & 1u)) {         // lights no queued hit can see: only their draws
                    qrng = pcg_hash(pcg_hash(qrng));
                    k++;
                }
                sp_k = k;
                qrest = uni ? uni >> (k + 1) : 0u;                // (strict: a batch may hold emission-only entries)
                const uint32_t s1 = pcg_hash(qrng);               // the two draws of :350-351 / :371-372, in order
                qrng = pcg_hash(s1);
                qt = -1.0f;                                       // nothing to find unless the sample is needed
                if ((lmask >> k) & 1u) {
                    const float3 P = f3(ring[DQ_PX][qe], ring[DQ_PY][qe], ring[DQ_PZ][qe]);
                    const float3 N = f3(ring[DQ_NX][qe], ring[DQ_NY][qe], ring[DQ_NZ][qe]);
                    const float3 C = f3(ring[DQ_CX][qe], ring[DQ_CY][qe], ring[DQ_CZ][qe]);
                    const float3 A = kStrict ? f3(ring[kStrict ? DQ_AX : 0][qe], ring[kStrict ? DQ_AY : 0][qe], ring[kStrict ? DQ_AZ : 0][qe]) : C;
                    qneed = nee_sample<CONV, SHAPE>(tr, k, P, N, C, A, __uint2float_rn(s1) * 2.3283064365386963e-10f,
                                       __uint2float_rn(qrng) * 2.3283064365386963e-10f, qo, qd, qt, qc);
                }
            }
        }
        int best = -1;
        float t_hit = -1.0f;
        if (fp.max_depth != 0u) {
            if (small) t_hit = tr.template scan_small_native<SHAPE>(qo, qd, qt, best);
            else t_hit = kStrict ? tr.template scan_lean<typename SHAPE::Runs>(qo, qd, qt, best) : tr.scan_native(qo, qd, qt, best);
        }

        if (spass) {                                              // shader.wgsl:402-431 for the queued samples
            const bool lit = qvalid && t_hit < 0.0f && qneed;
            // the entry's lights are summed in light order in a register; the entry's sum joins its chain's late sum
            // once, at its last light, in posting order (ordered_late_add)
            if (lit) qsum = V::add(qsum, qc);
            if (qrest) {
                if (qvalid) ring[DQ_RNG][qe] = __uint_as_float(qrng);
                sp_k++;
            } else {
                __syncwarp();
                // strict: the hit's own emission term (:455) first, then its light samples (:459), entry by entry
                const float3 E = kStrict && qvalid ? f3(ring[kStrict ? DQ_EX : 0][qe], ring[kStrict ? DQ_EY : 0][qe], ring[kStrict ? DQ_EZ : 0][qe])
                                                   : f3(0.0f, 0.0f, 0.0f);
                const float3 EA = kStrict && qvalid ? f3(ring[kStrict ? DQ_AX : 0][qe], ring[kStrict ? DQ_AY : 0][qe], ring[kStrict ? DQ_AZ : 0][qe])
                                                    : f3(0.0f, 0.0f, 0.0f);
                const bool any = qsum.x != 0.0f || qsum.y != 0.0f || qsum.z != 0.0f || E.x != 0.0f || E.y != 0.0f || E.z != 0.0f;
                ordered_late_add<CONV>(qvalid && any, qowner, TRT_CH(qowner, CH_L0), E, EA, qsum);
                sp_k = -1;
                head += qn;
                count_closable();
            }
            __syncwarp();
            continue;
        }

        // ---- closest hit: miss -> sky, hit -> record (shader.wgsl:447-459, :280-290) ----------------------------
        bool path_end = false, hit = false;
        float3 hp = f3(0, 0, 0), hn = f3(0, 1, 0), emis = f3(0, 0, 0);
        float4 mc = make_float4(0, 0, 0, 0);
        if (mode == PM_ACTIVE) {
            if (fp.max_depth == 0u) {
                path_end = true;
            } else {
                tr.n_bounces++;
                if (t_hit < 0.0f) {
                    if (kStrict) emis = tr.sky(rd);               // :450, posted like an emission term (order of additions)
                    else acc = V::vmad(acc, tr.sky(rd), att);
                    path_end = true;
                } else {
                    hit = true;
                    mc = __ldg(&sc.mc[best]);
                    const float4 me = __ldg(&sc.me[best]);
                    hp = V::axpy(ro, rd, t_hit);
                    hn = tr.normal_at(best, __float_as_int(me.w), hp);
                    if (!(V::dot(rd, hn) < 0.0f)) hn = V::scale(hn, -1.0f);
                    if (kStrict) emis = xyz(me);                  // :455, posted with the hit (order of additions)
                    else acc = V::vmad(acc, xyz(me), att);        // shader.wgsl:455
                }
            }
        }

        // ---- next-event estimation, bounce side (shader.wgsl:342-433): every light in order consumes its two
        // draws; the samples themselves are evaluated when the queue is resolved ------------------------------------
        const uint32_t rng0 = rng;
        unsigned lmask = 0;
        const float hp_mag = 1.0f + fabsf(hp.x) + fabsf(hp.y) + fabsf(hp.z);
        const int n_lights = SHAPE::kFixedLights ? SHAPE::NL : sc.n_lights;
#pragma unroll
        for (int k = 0; k < n_lights; k++) {
            const uint32_t adv = pcg_hash(pcg_hash(rng));
            float side;
            if (SHAPE::light_is_sphere(k, __float_as_int(tr.src.le(k).w))) {   // sphere: some of it above the tangent plane
                const float4 a = tr.src.la(k);
                side = V::dot(hn, V::sub(xyz(a), hp)) + a.w;
            } else {                                              // triangle: the point in front of its plane
                side = V::dot(xyz(tr.src.ln(k)), hp) - tr.src.lb(k).w;
            }
            if (hit) {
                rng = adv;
                if (side > -1e-5f * (hp_mag + tr.src.lc(k).w)) lmask |= 1u << k;
            }
        }
        const bool post = lmask != 0u || (kStrict && (emis.x != 0.0f || emis.y != 0.0f || emis.z != 0.0f));
        const unsigned mp = __ballot_sync(0xffffffffu, post);
        if (mp) {
            if (post) {
                const uint32_t e = (tail + __popc(mp & lt_mask)) & (TRT_NV_RING - 1u);
                ring[DQ_PX][e] = hp.x; ring[DQ_PY][e] = hp.y; ring[DQ_PZ][e] = hp.z;
                ring[DQ_NX][e] = hn.x; ring[DQ_NY][e] = hn.y; ring[DQ_NZ][e] = hn.z;
                if (kStrict) {
                    ring[DQ_CX][e] = mc.x; ring[DQ_CY][e] = mc.y; ring[DQ_CZ][e] = mc.z;
                    ring[kStrict ? DQ_AX : 0][e] = att.x; ring[kStrict ? DQ_AY : 0][e] = att.y; ring[kStrict ? DQ_AZ : 0][e] = att.z;
                    ring[kStrict ? DQ_EX : 0][e] = emis.x; ring[kStrict ? DQ_EY : 0][e] = emis.y; ring[kStrict ? DQ_EZ : 0][e] = emis.z;
                } else {
                    ring[DQ_CX][e] = mc.x * att.x; ring[DQ_CY][e] = mc.y * att.y; ring[DQ_CZ][e] = mc.z * att.z;
                }
                ring[DQ_RNG][e] = __uint_as_float(rng0);
                ring[DQ_META][e] = __uint_as_float(cid | (lmask << 8));      // owner chain | light mask << 8
            }
            tail += __popc(mp);
        }

        // ---- scatter + roulette (shader.wgsl:459-492) --------------------------------------------------------------
        if (hit) {
            float3 s;
            if (mc.w > rnd(rng)) s = V::reflect(rd, hn);
            else s = kStrict ? tr.cosine_hemisphere(rng, hn) : tr.cosine_hemisphere_unit(rng, hn);
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

        if (path_end) mode = PM_ENDED;                           // its chain is handed in at the top of the loop
        __syncwarp();
    }
    tr.flush_stats();
#undef TRT_CH
}

template <int CONV, class SRC, class SHAPE>
__global__ void __launch_bounds__(TRT_NV_THREADS, CONV != 2 ? TRT_PL_STRICT_CTAS : TRT_PL_CTAS_PER_SM)
trace_pool_kernel(SceneView sc, LightView lv, FrameParams fp) {
    trace_pool_body<CONV, SRC, SHAPE>(sc, lv, fp);
}

}  // namespace trt
