// trt_native.cu -- the trace kernel in the "device native" arithmetic convention (conv 2).
//
// Same kernel source as the strict build (trt_wave.cuh: same state machine, same random-number
// order, same tests in the same scan order); what differs is the arithmetic: this translation unit
// is compiled with --fmad=true -use_fast_math, so a*b+c contracts to FFMA and `/`, sqrt, sin, cos
// map to the SM's MUFU units (1-2 ulp) -- exactly the latitude WGSL gives the reference's own shader
// compiler (naga -> driver; DESIGN.md "Arithmetic conventions").  Results track the strict build to
// rounding noise; the tolerance tests (tests/test_gpu_native.py) state the bounds.
#define TRT_NATIVE_TU 1
#ifndef TRT_WV_CTAS_PER_SM
#define TRT_WV_CTAS_PER_SM 8      // 64 registers per thread: 32 resident warps per SM (A/B: 6 -> 3083, 7 -> 3204, 8 -> 3242 Msamples/s)
#endif
#include "trt_kernels.cuh"
#include "trt_persistent.cuh"
#include "trt_wave.cuh"
#include "trt_native.cuh"
#include "trt_pool.cuh"
#include "trt_jit.cuh"
#include <cstdlib>
#include <cstring>

namespace trt {

// The chain-pool machine (trt_pool.cuh).  A pixel's wall-clock time grows with chains / lanes, so only frames that hold
// several pools' worth of pixels may use it (small frames are bounded by their longest pixel).  Measured on B200
// (profiles/r2_ab_pool.txt, C3): strict builds +3.4 % (2 412 -> 2 495 Msamples/s) -- on by default there; conv 2 -0.9 %
// (it needs 72 registers, i.e. 7 resident CTAs instead of 8) -- off by default.  TRT_B200_POOL=1 / 0 forces it on
// (wherever a pool kernel exists, whatever the frame size) / off; =auto applies the size rule to both.
bool pool_wanted(uint32_t n_work, int sm_count, int ctas_per_sm, bool strict) {
    static const char *env = getenv("TRT_B200_POOL");
    if (env && env[0] == '0') return false;
    if (env && env[0] == '1') return true;
    if (!strict && !(env && env[0] == 'a')) return false;
    const uint64_t chains = (uint64_t)sm_count * (uint64_t)ctas_per_sm * (TRT_NV_THREADS / 32) * TRT_PL_CHAINS;
    return (uint64_t)n_work >= 8u * chains;
}

cudaError_t native_trace_launch(int device, const void *ctx_key, unsigned long long serial, const SceneView &sv,
                                const LightView &lv, const SmallScene *small, const FrameParams &fp, int sm_count,
                                cudaStream_t stream, uint32_t *launches, const char **kernel_name) {
    // this translation unit has its own __constant__ mirrors of the scene; rebinding synchronises the device
    static ConstOwner small_owner[64];
    const bool rebind = device >= 0 && device < 64 && !(small_owner[device].key == ctx_key && small_owner[device].serial == serial);
    cudaError_t e = persistent_bind_scene(device, ctx_key, serial, sv, lv, stream);
    if (e != cudaSuccess) return e;
    if (rebind) {
        e = cudaDeviceSynchronize();
        if (e == cudaSuccess) e = cudaMemcpyToSymbol(c_small, small, sizeof(SmallScene));
        if (e != cudaSuccess) return e;
        small_owner[device].key = ctx_key;
        small_owner[device].serial = serial;
    }
    e = cudaMemsetAsync(fp.work_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    // deferred-NEE kernel (trt_native.cuh) when the light count fits its mask.  TRT_B200_NATIVE=wave forces the
    // wave kernel in the same convention, =generic the unspecialised scan
    // (A/B measurements only)
    static const char *force = getenv("TRT_B200_NATIVE");
    const bool force_wave = force && !strcmp(force, "wave");
    const bool force_generic = force && !strcmp(force, "generic");
    const bool force_untiled = force && !strcmp(force, "untiled");     // large scenes: every warp scans on its own
    const bool constant = scene_fits_constant(sv);
    static bool carveout_done[64] = {};          // the attribute applies per device (one process may drive several GPUs)
    bool &carveout_set = carveout_done[device >= 0 && device < 64 ? device : 0];
    if (!carveout_set) {
        cudaFuncSetAttribute(trace_native_kernel<2, SrcConst, ShapeCornell>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<2, SrcConst, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<2, SrcGlobal, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<2, SrcTiled, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_pool_kernel<2, SrcConst, ShapeCornell>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        carveout_set = true;
    }
    const uint32_t n_work = fp.batch ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    const uint32_t want = (n_work + TRT_NV_THREADS - 1u) / TRT_NV_THREADS;
    // resident CTAs per SM: the caller's hint (trt_api.cu frame_ctas_hint), else all that fit (8); TRT_B200_CTAS overrides (A/B)
    static const char *ctas_env = getenv("TRT_B200_CTAS");
    uint32_t ctas = fp.ctas_hint >= 1u && fp.ctas_hint < TRT_NV_CTAS_PER_SM ? fp.ctas_hint : TRT_NV_CTAS_PER_SM;
    if (ctas_env && atoi(ctas_env) >= 1 && atoi(ctas_env) <= TRT_NV_CTAS_PER_SM) ctas = (uint32_t)atoi(ctas_env);
    const uint32_t cap = (uint32_t)sm_count * ctas;
    const uint32_t blocks = want < cap ? want : cap;
    // the reference's default scene shape (3 spheres, 6 axis-aligned walls in y y x x z z order, 2 triangles)
    // has a compile-time specialised scan; any other scene takes the same kernel with run-time counts
    bool cornell = constant && !force_generic && small->ok && small->n_sph == ShapeCornell::NS &&
                   small->n_pl == ShapeCornell::NP && small->n_tri == ShapeCornell::NT;
    for (int j = 0; cornell && j < ShapeCornell::NP; j++) {
        int kind;
        memcpy(&kind, &small->pl_p[j].w, 4);
        cornell = kind == ShapeCornell::kind(j);
    }
    cornell = cornell && sv.n_lights == ShapeCornell::NL && small->n_light_spheres == 0 && small->tri_zero == ShapeCornell::TZ;
    // a canonical small scene of any other shape gets a kernel compiled for that shape at run time (trt_jit.cuh)
    JitKernel *jk = nullptr;
    const int jit = jit_mode(nullptr);
    if (!force_wave && !force_generic && constant && small->jit_ok && sv.n_lights <= TRT_NV_MAX_LIGHTS && jit && (!cornell || jit == 2)) {
        jk = jit_kernel_for(device, *small);
        if (jk->ok) {
            e = jit_bind_scene(*jk, ctx_key, serial, sv, lv, *small, stream);
            if (e != cudaSuccess) return e;
        } else {
            jk = nullptr;
        }
    }
    const char *name = "";
    if (sv.grid_off) {
        // opt-in acceleration structure: persistent lanes, every lane walks the grid for its own ray
        const uint32_t wwant = (n_work + TRT_WV_THREADS - 1u) / TRT_WV_THREADS;
        const uint32_t wcap = (uint32_t)sm_count * TRT_WV_CTAS_PER_SM;
        const uint32_t wblocks = wwant < wcap ? wwant : wcap;
        if (fp.batch) trace_wave_kernel<2, SrcGrid, true><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_wave_kernel<2, SrcGrid, false><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
        name = "trace_wave_kernel<2, SrcGrid> (uniform grid, opt-in)";
    } else if (jk) {
        void *args[] = {(void *)&sv, (void *)&lv, (void *)&fp};
        e = cudaLaunchKernel((const void *)jk->kernel, dim3(blocks), dim3(TRT_NV_THREADS), args, 0, stream);
        if (e != cudaSuccess) return e;
        name = jk->name.c_str();
    } else if (!force_wave && sv.n_lights <= TRT_NV_MAX_LIGHTS) {
        if (cornell && pool_wanted(n_work, sm_count, TRT_PL_CTAS_PER_SM, false)) {
            const uint32_t pcap = (uint32_t)sm_count * TRT_PL_CTAS_PER_SM;
            trace_pool_kernel<2, SrcConst, ShapeCornell><<<want < pcap ? want : pcap, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
            name = "trace_pool_kernel<SrcConst, ShapeCornell>";
        } else if (cornell) {
            trace_native_kernel<2, SrcConst, ShapeCornell><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
            name = "trace_native_kernel<SrcConst, ShapeCornell>";
        } else if (constant) {
            trace_native_kernel<2, SrcConst, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
            name = "trace_native_kernel<SrcConst, ShapeAny>";
        } else if (sv.n_tiles > 0 && !force_untiled) {
            const uint32_t tcap = (uint32_t)sm_count * SrcTiled::kCtasPerSm;
            trace_native_kernel<2, SrcTiled, ShapeAny><<<want < tcap ? want : tcap, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
            name = "trace_native_kernel<SrcTiled, ShapeAny>";
        } else {
            trace_native_kernel<2, SrcGlobal, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
            name = "trace_native_kernel<SrcGlobal, ShapeAny>";
        }
    } else {
        name = constant ? "trace_wave_kernel<2, SrcConst>" : "trace_wave_kernel<2, SrcGlobal>";
        const uint32_t wwant = (n_work + TRT_WV_THREADS - 1u) / TRT_WV_THREADS;
        const uint32_t wcap = (uint32_t)sm_count * TRT_WV_CTAS_PER_SM;
        const uint32_t wblocks = wwant < wcap ? wwant : wcap;
        if (constant) {
            if (fp.batch) trace_wave_kernel<2, SrcConst, true><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
            else trace_wave_kernel<2, SrcConst, false><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
        } else {
            if (fp.batch) trace_wave_kernel<2, SrcGlobal, true><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
            else trace_wave_kernel<2, SrcGlobal, false><<<wblocks, TRT_WV_THREADS, 0, stream>>>(sv, lv, fp);
        }
    }
    *launches += 1;
    if (kernel_name) *kernel_name = name;
    return cudaGetLastError();
}

// conv 2 instance of the frame-batch blend (same translation unit, hence same flags, as finish_pixel<2>)
cudaError_t native_blend_launch(float4 *cur, const uint2 *batch, uint32_t n_batch, uint32_t n_local_pixels, const uint32_t *row_map,
                                uint32_t width, float4 *accumulation, float4 *pixels, cudaStream_t stream) {
    blend_frames_kernel<2><<<(n_local_pixels + 255u) / 256u, 256, 0, stream>>>(cur, batch, n_batch, n_local_pixels, row_map, width,
                                                                              accumulation, pixels);
    return cudaGetLastError();
}

}  // namespace trt
