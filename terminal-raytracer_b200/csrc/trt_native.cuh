// trt_native.cuh -- production trace kernel of the device-native convention (conv 2): persistent warps,
// one path per lane, and DEFERRED next-event estimation.
//
// The wave kernel (trt_wave.cuh) gives every lane one scan slot per iteration; a lane whose bounce
// needs a shadow ray spends its next slot on that ray and only then goes on, so that 28 % of all
// lane-iterations of the Cornell box are shadow iterations in which the lane skips most of the shading
// code the warp executes anyway (profiles/r1_native_wave_regions.txt).
//
// Visibility only gates an addition into the sample's radiance (shader.wgsl:402-431): it influences
// neither the random-number stream nor the path.  conv 2 is not bound to the order of floating-point
// additions, so here a lane does NOT wait for its light samples:
//   * light loop (:342-433) in the bounce iteration: per light only the two random draws are consumed
//     (the state moves on exactly as in the shader) and a cheap conservative test decides whether the
//     light can reach the shading point at all -- a triangle light is one-sided (cos_light = max(0,
//     dot(n_light, -dir)), :404), so a point behind its plane receives nothing whatever point is sampled;
//     a sphere light entirely below the point's tangent plane likewise (cos_surface <= 0, :403);
//   * a hit that some light may reach posts ONE entry on a per-warp FIFO in shared memory: hit point,
//     normal, colour x attenuation, the random state before the light loop and the mask of those lights;
//     the lane goes straight on to the scatter step -- one full bounce per iteration for every live lane;
//   * when 32 entries have queued up the warp resolves them at full SIMT width: every lane takes one
//     entry and, light by light (warp-uniform index: no divergence on the light type), re-derives the two
//     draws, evaluates the light sample (:352-389), the cosines and the clamped contribution (:403-429) and
//     scans its shadow ray (:394-400); an unoccluded contribution is added to the owner's "late" sum;
//   * a base sample is closed (:548-551) only when every entry it posted is resolved, but the lane does not wait
//     for that: the sample is parked (TRT_NV_SLOTS below) and the next one starts; an adaptive extra sample is
//     finished at once and its light samples join the pixel's sum (:596) any time before the pixel ends.  Only the
//     last sample of a phase waits (waiting lanes make the queue flush sooner).
// Per pixel the order of random draws, of samples and of bounces is the shader's; only the order in which light
// contributions are summed differs (rounding-level, inside conv 2's tolerance).
// There is no limit on the number of lights beyond the 16-bit mask (TRT_NV_MAX_LIGHTS).
#pragma once
#include "trt_trace.cuh"

namespace trt {

#ifndef TRT_NV_THREADS
#define TRT_NV_THREADS 128
#endif
#ifndef TRT_NV_CTAS_PER_SM
#define TRT_NV_CTAS_PER_SM 8
#endif
// Tuning switches (A/B through TRT_B200_JIT_DEFS / the Makefile's VARIANTS; measured in profiles/r1_ab_variants_v6.txt):
// TRT_NV_WAIT_W: weight of a waiting lane in the flush rule of the queue;
// TRT_NV_SLOTS (power of two): a BASE sample whose light samples are still queued is parked in shared memory
//   (its radiance so far, and the number of entries the lane had posted when it ended) and the lane starts the next
//   sample; queue entries carry the sample's slot, so every parked sample has its own late sum.  Parked samples are
//   closed in sample order as soon as their entries are resolved: the sums of :549-551 see the same values in the
//   same order as without parking.  1 = no parking.
// An adaptive EXTRA sample only adds to the pixel's sum (shader.wgsl:596): it is finished without waiting, its light
// samples land in slot 0's late sum until the pixel's last sample, which waits for all of them.
#ifndef TRT_NV_WAIT_W
#define TRT_NV_WAIT_W 1
#endif
#ifndef TRT_NV_SLOTS
#define TRT_NV_SLOTS 4
#endif
// TRT_V_ORDERED_LATE: light-sample sums join the owner's late sum in posting order (ordered_late_add) instead of by
// `red.shared.add` in whatever order the batches and the hardware produce: conv 2 frames become run-to-run and
// shard-union bit-identical.  0 = the round-1 behaviour (A/B).
#ifndef TRT_V_ORDERED_LATE
#define TRT_V_ORDERED_LATE 1
#endif
// TRT_V_DUAL_SHADOW: a queue pass of the small-scene kernel takes the samples of TWO lights and sends both shadow rays
// through one any-hit scan on packed pairs (Tracer::shadow2_small).  Measured (profiles/r2_ab_packed_fp32.txt): the queue
// passes halve (16.7 M -> 8.3 M on the A/B frame) and a pass for two lights costs ~640 instead of ~1 030 instructions,
// but at 64 registers the second ray's state spills into the bounce path (scatter 77 -> 107 instructions, post 26 -> 40):
// 37.13 G instead of 37.05 G warp-instructions, 36.47 -> 39.78 ms.  With 72 registers (7 CTAs) 36.61 -> 36.39 ms
// against 36.50 ms for the default: not adopted, 0 = one light per pass through the closest-hit scan.
#ifndef TRT_V_DUAL_SHADOW
#define TRT_V_DUAL_SHADOW 0
#endif
#define TRT_NV_RING 64           // queued hits per warp: at most 31 left over + 32 new ones per iteration
#define TRT_NV_MAX_LIGHTS 16     // light mask bits in the entry's meta word

enum { NM_ACTIVE = 0, NM_WAIT = 1, NM_NEED_PIXEL = 2, NM_DONE = 3 };
// queue entry, 11 words: hit point, unit normal (facing the ray), colour x attenuation, random state before
// the light loop (advanced light by light while the entry is being resolved), meta = owner thread | mask << 8
enum { DQ_PX, DQ_PY, DQ_PZ, DQ_NX, DQ_NY, DQ_NZ, DQ_CX, DQ_CY, DQ_CZ, DQ_RNG, DQ_META, DQ_FIELDS,
       // strict builds (conv 0 / 1) keep colour and attenuation apart -- the shader multiplies (colour / pi) * emission
       // and attenuation * (G * area) separately (:416-422), the product colour * attenuation would round differently --
       // and carry the hit's own emission term (:455), which must join the sample's sum BEFORE its light samples:
       // DQ_C* = colour, DQ_A* = attenuation, DQ_E* = the hit's emission (or the sky colour of a path that left the scene)
       DQ_AX = DQ_FIELDS, DQ_AY, DQ_AZ, DQ_EX, DQ_EY, DQ_EZ, DQ_FIELDS_STRICT };

// Explicit shared-space accesses of the per-lane record (32-bit shared addresses: base + immediate, no generic
// pointer kept in a register across the loop); "memory": ordered against the warp barriers around them.
__device__ __forceinline__ float lds_f(unsigned a) { float v; asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ unsigned lds_u(unsigned a) { unsigned v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a) : "memory"); return v; }
__device__ __forceinline__ void sts_f(unsigned a, float v) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void sts_u(unsigned a, unsigned v) { asm volatile("st.shared.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ void red_f(unsigned a, float v) { asm volatile("red.shared.add.f32 [%0], %1;" ::"r"(a), "f"(v) : "memory"); }
__device__ __forceinline__ void red_u(unsigned a, unsigned v) { asm volatile("red.shared.add.u32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }

// Ordered addition of a batch of light-sample sums into the owners' late sums.  Several lanes of one queue pass may
// hold entries of the SAME (owner, slot); `red.shared.add` would add them in an order the hardware chooses, and how
// entries fall into batches depends on which pixels share a warp (dynamic hand-out, row shards).  Here the lowest lane
// of every such group adds the group's values one after the other, in lane order = posting order, with plain loads and
// stores (different groups touch different words): late = ((late + e1) + e2) ... whatever the batches were, so a conv 2
// frame is run-to-run and shard-union bit-identical.  All 32 lanes call this; `active` lanes carry a value.
// CONV != 2 (strict builds): every lane first adds its hit's emission term the way the shader does -- acc = acc +
// emission * attenuation, :455, one fused multiply-add in conv 1 -- and then the sum of its light samples (:459);
// a zero emission changes nothing.  (A path that leaves the scene posts its sky term the same way, :450.)
template <int CONV>
__device__ __forceinline__ void ordered_late_add(bool active, unsigned key, unsigned addr, float3 e, float3 a, float3 v) {
    typedef M<CONV> V;
    constexpr bool PAIR = CONV != 2;
    const unsigned lane = threadIdx.x & 31u;
    const unsigned peers = __match_any_sync(0xffffffffu, active ? key : (0x80000000u | lane));
    const unsigned others = peers & ~(1u << lane);
    if (!__any_sync(0xffffffffu, active && others != 0u)) {      // the common case: every owner appears once
        if (active) {
            float3 t = f3(lds_f(addr), lds_f(addr + 4u), lds_f(addr + 8u));
            if (PAIR) t = V::vmad(t, e, a);
            t = V::add(t, v);
            sts_f(addr, t.x); sts_f(addr + 4u, t.y); sts_f(addr + 8u, t.z);
        }
        return;
    }
    const bool leader = active && (peers & ((1u << lane) - 1u)) == 0u;
    float3 t = f3(0.0f, 0.0f, 0.0f);
    if (leader) {
        t = f3(lds_f(addr), lds_f(addr + 4u), lds_f(addr + 8u));
        if (PAIR) t = V::vmad(t, e, a);
        t = V::add(t, v);
    }
    unsigned rest = leader ? others : 0u;
    const int n = (int)__reduce_max_sync(0xffffffffu, active ? (unsigned)__popc(peers) : 1u) - 1;
    for (int i = 0; i < n; i++) {
        const int src = rest ? __ffs(rest) - 1 : (int)lane;
        if (PAIR) {
            const float3 pe = f3(__shfl_sync(0xffffffffu, e.x, src), __shfl_sync(0xffffffffu, e.y, src), __shfl_sync(0xffffffffu, e.z, src));
            const float3 pa = f3(__shfl_sync(0xffffffffu, a.x, src), __shfl_sync(0xffffffffu, a.y, src), __shfl_sync(0xffffffffu, a.z, src));
            if (rest) t = V::vmad(t, pe, pa);
        }
        const float3 pv = f3(__shfl_sync(0xffffffffu, v.x, src), __shfl_sync(0xffffffffu, v.y, src), __shfl_sync(0xffffffffu, v.z, src));
        if (rest) { t = V::add(t, pv); rest &= rest - 1u; }
    }
    if (leader) { sts_f(addr, t.x); sts_f(addr + 4u, t.y); sts_f(addr + 8u, t.z); }
}

// The sample of light k for a queued hit (shader.wgsl:350-429): sample point from the two draws (:352-389), the
// two cosines (:403-404) and, when both are positive, the shadow ray (:394-399) and the clamped contribution
// (:408-429).  Returns whether the shadow ray is needed.
template <int CONV, class SHAPE, class TR>
__device__ __forceinline__ bool nee_sample(TR &tr, int k, float3 P, float3 N, float3 C, float3 A, float r1, float r2, float3 &qo,
                                           float3 &qd, float &qt, float3 &qc) {
    typedef M<CONV> V;
    const float4 a = tr.src.la(k);
    const float4 nl = tr.src.ln(k);
    const float4 em = tr.src.le(k);
    if (CONV != 2) {
        // as written, operation for operation (the sequence of trace_wave_kernel's light loop, trt_wave.cuh); C = the
        // hit's colour, A = the attenuation at the hit
        float3 light_point, light_normal;
        if (SHAPE::light_is_sphere(k, __float_as_int(em.w))) {        // :352-369
            const float cos_theta = 1.0f - 2.0f * r1;
            const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
            float sp, cp;
            sincos_turn<CONV>(r2, sp, cp);
            const float3 local = f3(sin_theta * cp, sin_theta * sp, cos_theta);
            light_point = V::axpy(xyz(a), local, a.w);
            light_normal = local;
        } else {                                                      // :370-389
            const float sqrt_r1 = sqrtf(r1);
            const float u = 1.0f - sqrt_r1;
            const float v = r2 * sqrt_r1;
            light_point = V::lin3(xyz(a), 1.0f - u - v, xyz(tr.src.lb(k)), u, xyz(tr.src.lc(k)), v);
            light_normal = xyz(nl);
        }
        const float3 light_vec = V::sub(light_point, P);
        const float d2 = V::dot(light_vec, light_vec);
        const float sl = V::dot(light_normal, light_vec), ss = V::dot(N, light_vec);
        // exact early out: a cosine that is certainly <= 0 (margin for the rounding of the normalisation)
        if ((sl > 0.0f && sl * sl > 1e-11f * d2) || (ss < 0.0f && ss * ss > 1e-11f * d2)) return false;
        const float light_distance = sqrtf(d2);
        const float3 light_dir = V::divs(light_vec, light_distance);
        const float cos_surface = fmaxf(0.0f, V::dot(N, light_dir));                                   // :403
        const float cos_light = fmaxf(0.0f, V::dot(light_normal, V::scale(light_dir, -1.0f)));         // :404
        if (!(cos_surface > 0.0f && cos_light > 0.0f)) return false;
        const float geometric = (cos_surface * cos_light) / (light_distance * light_distance);         // :408-410
        const float3 brdf = V::divs(C, TRT_PI_F);
        const float3 c = V::mulv(V::mulv(brdf, xyz(em)), V::scale(A, geometric * nl.w));
        qc = f3(fminf(c.x, 10.0f), fminf(c.y, 10.0f), fminf(c.z, 10.0f));
        qo = V::axpy(P, N, 0.001f);                                                                    // :399
        qd = light_dir;
        qt = light_distance - 0.001f;
        return true;
    }
    float3 light_vec;
    float sl;
    if (SHAPE::light_is_sphere(k, __float_as_int(em.w))) {   // uniform sphere sampling :352-369
        const float cos_theta = 1.0f - 2.0f * r1;
        const float sin_theta = sqrt_native(1.0f - cos_theta * cos_theta);
        float sp, cp;
        sincos_turn<2>(r2, sp, cp);
        const float3 local = f3(sin_theta * cp, sin_theta * sp, cos_theta);
        light_vec = V::sub(V::axpy(xyz(a), local, a.w), P);
        sl = V::dot(local, light_vec);
    } else {                                          // uniform triangle sampling :370-389
        const float sqrt_r1 = sqrt_native(r1);
        const float u = 1.0f - sqrt_r1;
        const float v = r2 * sqrt_r1;
        const float3 light_point = V::lin3(xyz(a), 1.0f - u - v, xyz(tr.src.lb(k)), u, xyz(tr.src.lc(k)), v);
        light_vec = V::sub(light_point, P);
        sl = V::dot(xyz(nl), light_vec);
    }
    const float ss = V::dot(N, light_vec);
    if (!(ss > 0.0f && sl < 0.0f)) return false;      // both cosines must be positive (:403-404)
    const float d2 = V::dot(light_vec, light_vec);
    const float inv_len = rsqrt_native(d2);
    qo = V::axpy(P, N, 0.001f);                       // :394-399
    qd = V::scale(light_vec, inv_len);
    qt = d2 * inv_len - 0.001f;
    const float geometric = ((ss * inv_len) * -(sl * inv_len)) * rcp_native(d2);
    const float3 cc = V::mulv(V::scale(C, 0.31830988618f), V::scale(xyz(em), geometric * nl.w));
    qc = f3(fminf(cc.x, 10.0f), fminf(cc.y, 10.0f), fminf(cc.z, 10.0f));
    return true;
}

// the kernel body; the __global__ entry points are trace_native_kernel below and, for a scene-specialised build
// compiled at run time, trt_jit_trace (trt_jit.cpp)
// CONV 2: the production arithmetic.  CONV 0 / 1: the same machine in the strict conventions, BIT-EXACT against the
// oracle: every term of a sample's radiance (the emission of a hit :455, its light samples :458-459) goes through the
// sample's late sum in posting order = the shader's order of additions (ordered_late_add), nothing is added ahead of an
// older pending term, and adaptive extra samples are closed whole like base samples (:596 adds complete values).
template <int CONV, class SRC, class SHAPE>
__device__ __forceinline__ void trace_native_body(const SceneView &sc, const LightView &lv, const FrameParams &fp) {
    typedef M<CONV> V;
    static_assert(CONV == 2 || !SRC::kTiled, "the tiled cooperative scan exists in the native convention only");
    constexpr bool kStrict = CONV != 2;
    constexpr int kFields = kStrict ? (int)DQ_FIELDS_STRICT : (int)DQ_FIELDS;
    Counters<false> cn;
    Tracer<CONV, false, SRC> tr(sc, lv, fp, cn);
    __shared__ float s_ring[TRT_NV_THREADS / 32][kFields][TRT_NV_RING];
    // Parking needs the queue to resolve a lane's entries in the order they were posted (the small-scene queue does:
    // an entry is resolved whole, at its last light).  The two-stage queue of SrcTiled resolves a light sample that
    // is not needed on the spot, ahead of older samples still in the ray FIFO: no parking there (one slot).
    constexpr unsigned kSlots = SRC::kTiled ? 1u : (unsigned)TRT_NV_SLOTS;
    // per-lane record (words): late sum [slot][3], parked radiance [slot][3], entries posted when the parked sample
    // ended [slot], entries resolved so far, the lane's bounce counter when its pixel began (-> fp.cost); odd stride
    // (one slot: nothing is ever parked, the record is late sum, resolved count, bounce mark = 5 words)
    constexpr unsigned kLate = 0u, kPark = kSlots > 1u ? 3u * kSlots : 0u, kMark = kSlots > 1u ? 6u * kSlots : 0u;
    constexpr unsigned kRes = kSlots > 1u ? 7u * kSlots : 3u, kB0 = kRes + 1u;
    constexpr unsigned kRec = (kB0 + 1u) | 1u;
    __shared__ float s_lane[TRT_NV_THREADS * kRec];
    __shared__ __align__(128) float4 s_tile[SRC::kTileSlots];   // staged geometry of the CTA-cooperative scan (SrcTiled)
    __shared__ __align__(8) unsigned long long s_mbar[2];       // SrcTiled + TRT_V_TMA_TILES: one mbarrier per tile buffer
    unsigned tile_phase = 0;
#ifndef __CUDACC_RTC__
    if (SRC::kTiled && TRT_V_TMA_TILES) {
        if (threadIdx.x == 0) {
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&s_mbar[0])) : "memory");
            asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&s_mbar[1])) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
#endif
    // SrcTiled: second FIFO, of ready shadow rays (origin, direction, contribution, t_max in DQ_RNG, owner in DQ_META)
    __shared__ float s_rays[SRC::kTiled ? TRT_NV_THREADS / 32 : 1][DQ_FIELDS][SRC::kTiled ? TRT_NV_RING : 1];
    const unsigned tid = threadIdx.x, lane = tid & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    float (*ring)[TRT_NV_RING] = s_ring[tid >> 5];
    float (*rays)[SRC::kTiled ? TRT_NV_RING : 1] = s_rays[SRC::kTiled ? tid >> 5 : 0];
    const unsigned lane_base = (unsigned)__cvta_generic_to_shared(s_lane);
#define TRT_REC(thread, word) (lane_base + 4u * ((thread) * kRec + (word)))
#pragma unroll
    for (unsigned i = 0; i < kRec; i++) sts_u(TRT_REC(tid, i), 0u);
    __syncwarp();

    uint32_t head = 0, tail = 0;                 // FIFO counters, identical in every lane of the warp
    uint32_t rhead = 0, rtail = 0;               // SrcTiled: the FIFO of ready shadow rays
    // small scenes: >= 0 while the warp resolves the oldest entries against light sp_k; SrcTiled: the next light the
    // oldest batch of hits is expanded against
    int sp_k = SRC::kTiled ? 0 : -1;
    uint32_t sp_n = 0;                           // SrcTiled: number of hits in the batch being expanded
    uint32_t pend = 0;                           // entries this lane has posted so far
    int mode = NM_NEED_PIXEL;
    uint32_t xy = 0, rng = 0, si = 0, target = 0, depth = 0, fslot = 0;
    uint32_t fin = 0;                            // samples of the pixel closed so far; fin .. si - 1 are parked
    const uint32_t n_work = fp.batch ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    bool extra_phase = false;
    float3 pixel_color = f3(0, 0, 0), sq_sum = f3(0, 0, 0);
    float3 att = f3(1, 1, 1), acc = f3(0, 0, 0);
    float3 ro = f3(0, 0, 0), rd = f3(0, 0, 1);   // this lane's closest-hit ray
    float3 qsum = f3(0, 0, 0);                   // small scenes: the sum over the lights of the entry this lane is resolving
    const bool small = !kStrict && (SHAPE::kFixed || (SRC::kSmall && c_small.ok));

#if TRT_TAIL_DIAG
    // measurement build: when (0.5 ms bins from the first warp's start) this warp first found the work list empty
    // -> stats[520 + bin], and when it left -> stats[8 + bin]
    unsigned long long t_base;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_base));
    {
        const unsigned long long old = atomicCAS(&fp.stats[2], 0ull, t_base);
        if (old) t_base = old;
        t_base = __shfl_sync(0xffffffffu, t_base, 0);
    }
    bool saw_dry = false;
    unsigned long long t_pix = 0;
#endif
    for (;;) {
        // ---- finish the sample of a lane whose path has ended and whose queue entries are all resolved ------
        // (shader.wgsl:548-551, :555-565, :596-621)
        bool new_ray = false;
        bool retire = false, closed = false;
        if (mode == NM_WAIT) {
            const unsigned r = lds_u(TRT_REC(tid, kRes));
            // parked samples whose entries have all been resolved, oldest first
            while (kSlots > 1u && fin != si) {
                const unsigned ps = fin & (kSlots - 1u);
                if ((int)(r - lds_u(TRT_REC(tid, kMark + ps))) < 0) break;
                const unsigned al = TRT_REC(tid, kLate + 3u * ps), ap = TRT_REC(tid, kPark + 3u * ps);
                const float3 a = f3(lds_f(ap) + lds_f(al), lds_f(ap + 4u) + lds_f(al + 4u), lds_f(ap + 8u) + lds_f(al + 8u));
                sts_f(al, 0.0f); sts_f(al + 4u, 0.0f); sts_f(al + 8u, 0.0f);
                pixel_color = V::add(pixel_color, a);
                if (!extra_phase) sq_sum = V::add(sq_sum, V::mulv(a, a));      // (strict: extra samples are parked too)
                fin++;
            }
            // the sample that has just ended: an extra sample only adds to the pixel's sum (:596) and its light samples
            // may arrive until the pixel's last sample (all in slot 0); a base sample is closed if nothing is
            // outstanding, else parked if a slot is free for the next one; the last sample of a phase waits
            const unsigned cs = (extra_phase && !kStrict) ? 0u : si & (kSlots - 1u);
            const bool last = si + 1u == target;
            if (!kStrict && extra_phase && !last) {
                retire = closed = true;
            } else if (fin == si && r == pend) {
                const unsigned al = TRT_REC(tid, kLate + 3u * cs);
                acc = V::add(acc, f3(lds_f(al), lds_f(al + 4u), lds_f(al + 8u)));
                sts_f(al, 0.0f); sts_f(al + 4u, 0.0f); sts_f(al + 8u, 0.0f);
                retire = closed = true;
            } else if (kSlots > 1u && !last && si + 1u - fin < kSlots) {
                const unsigned ap = TRT_REC(tid, kPark + 3u * cs);
                sts_f(ap, acc.x); sts_f(ap + 4u, acc.y); sts_f(ap + 8u, acc.z);
                sts_u(TRT_REC(tid, kMark + cs), pend);         // its value joins the sums when it is closed
                retire = true;
            }
        }
        if (retire) {
            if (closed) {
                fin++;
                pixel_color = V::add(pixel_color, acc);
                if (!extra_phase) sq_sum = V::add(sq_sum, V::mulv(acc, acc));
            }
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
                tr.finish_pixel(xy & 0xffffu, xy >> 16, pixel_color, variance_value, fslot);
                tr.n_samples += target;                        // the samples this pixel traced
                if (fp.cost) fp.cost[(size_t)(xy >> 16) * fp.width + (xy & 0xffffu)] = tr.n_bounces - lds_u(TRT_REC(tid, kB0));
#if TRT_TAIL_DIAG >= 2
                if (saw_dry) {                                 // the last 250 pixels to finish: who, how long, from when to when
                    unsigned long long t;
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                    const unsigned long long i = atomicAdd(&fp.stats[3], 1ull) % 250ull;
                    fp.stats[520 + 2 * i] = (unsigned long long)xy | ((unsigned long long)target << 32);
                    fp.stats[521 + 2 * i] = (((t_pix - t_base) / 100000ull) << 32) | ((t - t_base) / 100000ull);
                }
#endif
                mode = NM_NEED_PIXEL;
            } else {
                new_ray = true;
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
                    rng = (y * fp.width + x) * 1973u + seed * 9277u + frame_number * 12345u;   // shader.wgsl:507
                    pixel_color = f3(0, 0, 0);
                    sq_sum = f3(0, 0, 0);
                    si = 0;
                    fin = 0;
                    target = fp.base_samples;
                    extra_phase = false;
                    new_ray = true;
                    sts_u(TRT_REC(tid, kB0), tr.n_bounces);
#if TRT_TAIL_DIAG >= 2
                    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_pix));
#endif
                } else {
                    mode = NM_DONE;
                }
            }
        }
        // ---- next sample of the pixel: reseed (shader.wgsl:517 / :565; si = 0 for a new pixel), camera ray ----
        if (new_ray) {
            rng = pcg_hash(rng + si * 5096u);
            tr.camera_ray(rng, xy & 0xffffu, xy >> 16, ro, rd);
            att = f3(1.0f, 1.0f, 1.0f);
            acc = f3(0.0f, 0.0f, 0.0f);
            depth = 0;
            mode = NM_ACTIVE;
        }
        // (the cooperative scan of SrcTiled needs every warp of the CTA in every iteration: leave together)
#if TRT_TAIL_DIAG
        if (!saw_dry && __any_sync(0xffffffffu, mode == NM_DONE)) {
            saw_dry = true;
            unsigned long long t;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            const unsigned long long b = (t - t_base) / 500000ull;
            if (lane == 0 && TRT_TAIL_DIAG < 2) atomicAdd(&fp.stats[520 + (b < 511ull ? b : 511ull)], 1ull);
        }
#endif
        if (SRC::kTiled ? __syncthreads_and(mode == NM_DONE) : __all_sync(0xffffffffu, mode == NM_DONE)) break;

        // ---- this iteration's scan: queued light samples (when enough have queued up, or lanes are waiting for
        // them) or every lane's own closest-hit ray -------------------------------------------------------------------
        bool spass = false, qvalid = false, qneed = false;
        float3 qo = ro, qd = rd, qc = f3(0, 0, 0);
        float qt = 1e10f;
        unsigned qowner = 0, qrest = 0, qslot = 0;
        uint32_t qrng = 0, qn = 0;
        // TRT_V_DUAL_SHADOW (small scenes, conv 2): the samples of TWO lights per queue pass, both shadow rays in one scan
        bool dual = false, qneedB = false;
        float3 qdB = f3(0, 0, 1), qcB = f3(0, 0, 0);
        float qtB = -1.0f;
        const uint32_t qe = (head + lane) & (TRT_NV_RING - 1u);
        if (SRC::kTiled) {
            // Large scenes: a scan costs ~0.5 M instructions per warp, so only NEEDED shadow rays may occupy a scan
            // slot.  Stage 1 (no scan): the queued hits are expanded light by light into a second FIFO of ready
            // shadow rays; a sample that is not needed (a cosine <= 0) is resolved on the spot.  Stage 2: 32 queued
            // rays share one scan.
            const unsigned m_active = __ballot_sync(0xffffffffu, mode == NM_ACTIVE);
            const uint32_t waiters2 = 2u * __popc(__ballot_sync(0xffffffffu, mode == NM_WAIT));
            uint32_t cnt = tail - head, rcnt = rtail - rhead;
            while (cnt > 0u && rcnt < 32u && (cnt >= 32u || !m_active || cnt + waiters2 >= 32u)) {
                if (sp_k == 0) sp_n = cnt < 32u ? cnt : 32u;      // a batch keeps its size while hits queue up behind it
                const uint32_t en = sp_n;
                const uint32_t ee = (head + lane) & (TRT_NV_RING - 1u);
                unsigned lmask = 0, owner = 0;
                uint32_t erng = 0;
                if (lane < en) {
                    const unsigned meta = __float_as_uint(ring[DQ_META][ee]);
                    owner = meta & 0xffu;                         // owner | slot << 8 travels on to the ray FIFO
                    lmask = (meta >> 8) & 0xffffu;
                    owner |= (meta >> 24) << 8;
                    erng = __float_as_uint(ring[DQ_RNG][ee]);
                }
                const unsigned uni = __reduce_or_sync(0xffffffffu, lmask);
                int k = sp_k;
                while (!((uni >> k) & 1u)) {                      // lights no queued hit can see: only their draws
                    erng = pcg_hash(pcg_hash(erng));
                    k++;
                }
                const uint32_t s1 = pcg_hash(erng);               // the two draws of :350-351 / :371-372, in order
                erng = pcg_hash(s1);
                bool needed = false;
                float3 so = f3(0, 0, 0), sd = f3(0, 0, 1), scn = f3(0, 0, 0);
                float st = 0.0f;
                if ((lmask >> k) & 1u) {
                    const float3 P = f3(ring[DQ_PX][ee], ring[DQ_PY][ee], ring[DQ_PZ][ee]);
                    const float3 N = f3(ring[DQ_NX][ee], ring[DQ_NY][ee], ring[DQ_NZ][ee]);
                    const float3 C = f3(ring[DQ_CX][ee], ring[DQ_CY][ee], ring[DQ_CZ][ee]);
                    needed = nee_sample<CONV, SHAPE>(tr, k, P, N, C, C, __uint2float_rn(s1) * 2.3283064365386963e-10f,
                                        __uint2float_rn(erng) * 2.3283064365386963e-10f, so, sd, st, scn);
                    if (!needed) red_u(TRT_REC(owner & 0xffu, kRes), 1u);
                }
                const unsigned mn = __ballot_sync(0xffffffffu, needed);
                if (needed) {
                    const uint32_t e = (rtail + __popc(mn & lt_mask)) & (TRT_NV_RING - 1u);
                    rays[DQ_PX][e] = so.x; rays[DQ_PY][e] = so.y; rays[DQ_PZ][e] = so.z;
                    rays[DQ_NX][e] = sd.x; rays[DQ_NY][e] = sd.y; rays[DQ_NZ][e] = sd.z;
                    rays[DQ_CX][e] = scn.x; rays[DQ_CY][e] = scn.y; rays[DQ_CZ][e] = scn.z;
                    rays[DQ_RNG][e] = st;
                    rays[DQ_META][e] = __uint_as_float(owner);
                }
                rtail += __popc(mn);
                if (uni >> (k + 1)) {                             // more lights for this batch of hits
                    sp_k = k + 1;
                    if (lane < en) ring[DQ_RNG][ee] = __uint_as_float(erng);
                } else {
                    sp_k = 0;
                    head += en;
                }
                __syncwarp();
                cnt = tail - head;
                rcnt = rtail - rhead;
            }
            spass = rcnt >= 32u || (rcnt > 0u && (!m_active || rcnt + waiters2 >= 32u));
            if (spass) {
                qn = rcnt < 32u ? rcnt : 32u;
                qvalid = lane < qn;
                qt = -1.0f;
                if (qvalid) {
                    const uint32_t e = (rhead + lane) & (TRT_NV_RING - 1u);
                    qo = f3(rays[DQ_PX][e], rays[DQ_PY][e], rays[DQ_PZ][e]);
                    qd = f3(rays[DQ_NX][e], rays[DQ_NY][e], rays[DQ_NZ][e]);
                    qc = f3(rays[DQ_CX][e], rays[DQ_CY][e], rays[DQ_CZ][e]);
                    qt = rays[DQ_RNG][e];
                    qowner = __float_as_uint(rays[DQ_META][e]);
                    qslot = qowner >> 8;
                    qowner &= 0xffu;
                    qneed = true;
                }
            }
        } else {
            // Small scenes: a scan is a few hundred instructions; the queued hits are resolved in place, one light
            // per iteration (warp-uniform light index), the lanes whose sample is not needed idle through the scan.
            const uint32_t cnt = tail - head;
            if (sp_k < 0 && cnt > 0u) {
                const unsigned m_active = __ballot_sync(0xffffffffu, mode == NM_ACTIVE);
                const unsigned m_wait = __ballot_sync(0xffffffffu, mode == NM_WAIT);
                if (cnt >= 32u || cnt + TRT_NV_WAIT_W * __popc(m_wait) >= 32u || !m_active) { sp_k = 0; qsum = f3(0, 0, 0); }
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
                    qslot = meta >> 24;
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
#if TRT_V_DUAL_SHADOW && TRT_V_PACKED && TRT_V_SHIFT_T && !TRT_V_KEY
                if (!kStrict && small && qrest != 0u) {           // the next light some queued hit can see, in the same pass
                    dual = true;
                    const int k2 = k + __ffs((int)qrest);
                    for (int i = k + 1; i < k2; i++) qrng = pcg_hash(pcg_hash(qrng));      // lights in between: only their draws
                    const uint32_t s1b = pcg_hash(qrng);
                    qrng = pcg_hash(s1b);
                    if ((lmask >> k2) & 1u) {
                        const float3 P = f3(ring[DQ_PX][qe], ring[DQ_PY][qe], ring[DQ_PZ][qe]);
                        const float3 N = f3(ring[DQ_NX][qe], ring[DQ_NY][qe], ring[DQ_NZ][qe]);
                        const float3 C = f3(ring[DQ_CX][qe], ring[DQ_CY][qe], ring[DQ_CZ][qe]);
                        float3 qoB;
                        qneedB = nee_sample<CONV, SHAPE>(tr, k2, P, N, C, C, __uint2float_rn(s1b) * 2.3283064365386963e-10f,
                                            __uint2float_rn(qrng) * 2.3283064365386963e-10f, qoB, qdB, qtB, qcB);
                        if (qneedB) qo = qoB;                     // (the same point, P + 0.001 N: set when only the second sample is needed)
                    }
                    sp_k = k2;
                    qrest = uni >> (k2 + 1);
                }
#endif
            }
        }
        int best = -1;
        float t_hit = -1.0f;
        bool occB = true;
        if (fp.max_depth != 0u) {
#if TRT_V_DUAL_SHADOW && TRT_V_PACKED && TRT_V_SHIFT_T && !TRT_V_KEY
            if (dual) {
                bool occA;
                tr.template shadow2_small<SHAPE>(qo, qd, qt, qdB, qtB, occA, occB);
                t_hit = occA ? 1.0f : -1.0f;
            } else
#endif
            if (small) t_hit = tr.template scan_small_native<SHAPE>(qo, qd, qt, best);
#ifndef __CUDACC_RTC__
            else if (SRC::kTiled) t_hit = tr.scan_native_tiled(qo, qd, mode == NM_ACTIVE || spass ? qt : -1.0f, best, s_tile, spass, s_mbar, tile_phase);
#endif
            else t_hit = kStrict ? tr.template scan_lean<typename SHAPE::Runs>(qo, qd, qt, best) : tr.scan_native(qo, qd, qt, best);
        }

        if (spass) {                                              // shader.wgsl:402-431 for the queued samples
            const bool lit = qvalid && t_hit < 0.0f && qneed;
#if !TRT_V_ORDERED_LATE
            if (qvalid) {
                if (lit) {
                    const unsigned al = TRT_REC(qowner, kLate + 3u * qslot);
                    red_f(al, qc.x); red_f(al + 4u, qc.y); red_f(al + 8u, qc.z);
                }
                if (SRC::kTiled) red_u(TRT_REC(qowner, kRes), 1u);
                else if (qrest) ring[DQ_RNG][qe] = __uint_as_float(qrng);
                else red_u(TRT_REC(qowner, kRes), 1u);
            }
#else
            if (SRC::kTiled) {
                // one ready ray per lane: added in FIFO order (= the owner's posting order, light by light)
                __syncwarp();
                ordered_late_add<2>(lit, qowner << 2 | qslot, TRT_REC(qowner, kLate + 3u * qslot), qc, qc, qc);
                if (qvalid) red_u(TRT_REC(qowner, kRes), 1u);
            } else {
                // the entry's lights are summed in light order in a register, the entry's sum joins the owner's late
                // sum once, at its last light, in posting order
                if (lit) qsum = V::add(qsum, qc);
                if (dual && qvalid && qneedB && !occB) qsum = V::add(qsum, qcB);
                if (qrest) {
                    if (qvalid) ring[DQ_RNG][qe] = __uint_as_float(qrng);
                } else {
                    __syncwarp();
                    // strict: the hit's own emission term (:455) first, then its light samples (:459), entry by entry
                    const float3 E = kStrict && qvalid ? f3(ring[kStrict ? DQ_EX : 0][qe], ring[kStrict ? DQ_EY : 0][qe], ring[kStrict ? DQ_EZ : 0][qe])
                                                       : f3(0.0f, 0.0f, 0.0f);
                    const float3 EA = kStrict && qvalid ? f3(ring[kStrict ? DQ_AX : 0][qe], ring[kStrict ? DQ_AY : 0][qe], ring[kStrict ? DQ_AZ : 0][qe])
                                                        : f3(0.0f, 0.0f, 0.0f);
                    const bool any = qsum.x != 0.0f || qsum.y != 0.0f || qsum.z != 0.0f || E.x != 0.0f || E.y != 0.0f || E.z != 0.0f;
                    ordered_late_add<CONV>(qvalid && any, qowner << 2 | qslot, TRT_REC(qowner, kLate + 3u * qslot), E, EA, qsum);
                    if (qvalid) red_u(TRT_REC(qowner, kRes), 1u);
                }
            }
#endif
            if (SRC::kTiled) {
                rhead += qn;
            } else if (qrest) {
                sp_k++;
            } else {
                sp_k = -1;
                head += qn;
            }
            __syncwarp();
            continue;
        }

        // ---- closest hit: miss -> sky, hit -> record (shader.wgsl:447-459, :280-290) ----------------------------
        bool path_end = false, hit = false;
        float3 hp = f3(0, 0, 0), hn = f3(0, 1, 0), emis = f3(0, 0, 0);
        float4 mc = make_float4(0, 0, 0, 0);
        if (mode == NM_ACTIVE) {
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
        // all lights triangles, known at compile time: the reach tests of lights 2k and 2k + 1 on packed pairs (SmallScene::
        // lt_pair) -- 6 packed instructions and two compares for two lights instead of 7 each
        constexpr bool kPairLights = TRT_V_PACKED != 0 && !kStrict && SHAPE::kFixedLights && SHAPE::NL >= 2 && SHAPE::NL <= 2 * TRT_SMALL_LPAIRS && SHAPE::LSPH == 0u;
        unsigned reach = 0;
        if (kPairLights) {
#pragma unroll
            for (int k = 0; 2 * k + 1 < n_lights; k++) {
                const float4 q0 = c_small.lt_pair[3 * k], q1 = c_small.lt_pair[3 * k + 1], q2 = c_small.lt_pair[3 * k + 2];
                const f2_t side2 = f2z_sub(f2z_fma(f2_pack(q1.x, q1.y), ss2(hp.z), f2z_fma(f2_pack(q0.z, q0.w), ss2(hp.y), f2z_mul(f2_pack(q0.x, q0.y), ss2(hp.x)))),
                                           f2_pack(q1.z, q1.w));
                const f2_t thr2 = f2z_mul(f2z_add(f2_pack(q2.x, q2.y), ss2(hp_mag)), ss2(-1e-5f));
                float s0, s1, t0, t1;
                f2_unpack(side2, s0, s1); f2_unpack(thr2, t0, t1);
                if (s0 > t0) reach |= 1u << (2 * k);
                if (s1 > t1) reach |= 2u << (2 * k);
            }
        }
#pragma unroll
        for (int k = 0; k < n_lights; k++) {
            const uint32_t adv = pcg_hash(pcg_hash(rng));
            if (kPairLights && (k | 1) < n_lights) {              // (an odd last light takes the scalar test below)
                if (hit) { rng = adv; lmask |= reach & (1u << k); }
                continue;
            }
            // a quantity that must be positive for the light to contribute, with a margin for the roundings of
            // the sample point and of the dot products (lb.w, lc.w: repack_kernel)
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
                // meta: owner thread | light mask << 8 | slot of the sample << 24 (extra samples: slot 0)
                ring[DQ_META][e] = __uint_as_float(tid | (lmask << 8) | ((kSlots > 1u && (kStrict || !extra_phase) ? si & (kSlots - 1u) : 0u) << 24));
                pend += SRC::kTiled ? (uint32_t)__popc(lmask) : 1u;      // units to resolve: light samples / the entry
            }
            tail += __popc(mp);
        }
        __syncwarp();

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
        if (path_end) mode = NM_WAIT;
    }
#if TRT_TAIL_DIAG
    {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        const unsigned long long b = (t - t_base) / 500000ull;
        if (lane == 0) atomicAdd(&fp.stats[8 + (b < 511ull ? b : 511ull)], 1ull);
    }
#endif
    tr.flush_stats();
#undef TRT_REC
}

#ifndef TRT_NV_STRICT_CTAS
#define TRT_NV_STRICT_CTAS 5     // strict builds: 96 registers (IEEE division / square-root sequences; 80 spill), 34 KB of shared
                                 // memory.  Cornell 1920x1080 x 256 spp, conv 0: 6 CTAs 2164, 5 CTAs 2289, 4 CTAs 2080 Msamples/s
#endif
template <int CONV, class SRC, class SHAPE>
__global__ void __launch_bounds__(TRT_NV_THREADS, CONV != 2 ? TRT_NV_STRICT_CTAS : SRC::kCtasPerSm ? SRC::kCtasPerSm : TRT_NV_CTAS_PER_SM)
trace_native_kernel(SceneView sc, LightView lv, FrameParams fp) {
    trace_native_body<CONV, SRC, SHAPE>(sc, lv, fp);
}

}  // namespace trt
