// trt_persistent.cuh -- production trace kernel (placeholder: dispatches the per-pixel kernel).
#pragma once
#include "trt_kernels.cuh"

namespace trt {

static inline cudaError_t persistent_scene_upload(const SceneView &, const LightView &, cudaStream_t) { return cudaSuccess; }

static inline cudaError_t persistent_trace_launch(const SceneView &sv, const LightView &lv, const FrameParams &fp, int conv,
                                                  int sm_count, cudaStream_t stream, uint32_t *launches) {
    (void)sm_count;
    const uint32_t blocks = (fp.n_local_pixels + 127u) / 128u;
    if (conv) trace_pixel_kernel<1, false><<<blocks, 128, 0, stream>>>(sv, lv, fp);
    else trace_pixel_kernel<0, false><<<blocks, 128, 0, stream>>>(sv, lv, fp);
    *launches += 1;
    return cudaGetLastError();
}

}  // namespace trt
