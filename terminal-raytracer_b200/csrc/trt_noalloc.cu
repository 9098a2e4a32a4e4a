// trt_noalloc.cu -- trt_quantise_noalloc: the one-pass terminal quantiser (quant_fused_kernel, trt_quantise.cuh) for
// callers that cannot allocate: code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
//
// No counterpart in the reference.  It changes no existing path: the
// context-based entry points (trt_quantise_async, trt_render_frame, ...) keep the scratch they allocate in trt_create.
//
// The kernel's scratch -- the tile ticket, one look-back state word per tile of 1024 pixels, the stream length -- lives
// in a file-scope __device__ array of fixed size (TRT_NOALLOC_SCRATCH_BYTES = 65 792 bytes for TRT_NOALLOC_MAX_PIXELS =
// 8 388 608 pixels: a 3840x2160 frame fits); every GPU that loads the library holds it until exit.  The entry point
// calls no cudaMalloc / cudaFree or allocating library routine; the caller passes the input and both outputs as
// device pointers.  ALL calls on a device share the array: it keeps the last call's contents, so the part that must
// start at zero is cleared by a kernel in front of every launch, and two calls that overlap on the GPU (different
// streams, no dependency) race on it -- serialise them.
#include <cstdio>
#include <cstdarg>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/trt_b200.h"
// This translation unit's own copy of the quantiser kernels, with internal linkage (trt_api.cu holds the copy the
// context-based entry points launch; the header's kernels are not templates, so two external copies would collide).
namespace {
#include "trt_quantise.cuh"
}

namespace {
using namespace trt;

static_assert(TRT_NOALLOC_MAX_PIXELS % TRT_QF_TILE == 0, "whole tiles");
constexpr uint32_t kNoallocMaxTiles = TRT_NOALLOC_MAX_PIXELS / TRT_QF_TILE;
// words: [0] ticket, [1 .. kNoallocMaxTiles] tile states, [kNoallocMaxTiles + 1] stream length; padded to 256 bytes
constexpr uint32_t kNoallocWords = ((kNoallocMaxTiles + 2u + 31u) / 32u) * 32u;
static_assert(kNoallocWords * sizeof(unsigned long long) == TRT_NOALLOC_SCRATCH_BYTES, "the size include/trt_b200.h states");

static __device__ __align__(256) unsigned long long g_noalloc_scratch[kNoallocWords];

// Device code names the array directly: ticket and the states of the tiles of this call back to zero.
__global__ void noalloc_clear_kernel(uint32_t n_words) {
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n_words; i += gridDim.x * blockDim.x) g_noalloc_scratch[i] = 0ull;
}
__global__ void noalloc_publish_kernel(uint32_t total_word, unsigned long long *d_len) { *d_len = g_noalloc_scratch[total_word]; }

// Host code needs the array's device address: the symbol's name is a host-side stand-in here, not a device pointer.
// Looked up once per device, with that device current.
static unsigned long long *g_scratch_ptr[64] = {};

static int say(char *err, size_t cap, int code, const char *fmt, ...) {
    if (err && cap) {
        va_list ap;
        va_start(ap, fmt);
        vsnprintf(err, cap, fmt, ap);
        va_end(ap);
    }
    return code;
}

}  // namespace

extern "C" int trt_quantise_noalloc(int device, const void *d_pixels, uint32_t width, uint32_t height, int mode, void *d_out,
                                    size_t out_cap, void *d_len, void *cuda_stream, char *err, size_t err_cap) {
    if (err && err_cap) err[0] = 0;
    if (device < 0 || device >= 64) return say(err, err_cap, TRT_ERR_INVALID, "trt_quantise_noalloc: device %d out of range", device);
    if (mode != TRT_MODE_ASCII && mode != TRT_MODE_TRUECOLOR) return say(err, err_cap, TRT_ERR_INVALID, "trt_quantise_noalloc: bad mode %d", mode);
    if (!d_pixels || !d_out || !d_len) return say(err, err_cap, TRT_ERR_INVALID, "trt_quantise_noalloc: null device pointer");
    if (((uintptr_t)d_pixels & 15u) || ((uintptr_t)d_len & 7u))
        return say(err, err_cap, TRT_ERR_INVALID, "trt_quantise_noalloc: d_pixels must be 16-byte and d_len 8-byte aligned");
    const uint64_t n = (uint64_t)width * height;
    if (width == 0u || height == 0u) return say(err, err_cap, TRT_ERR_INVALID, "trt_quantise_noalloc: empty frame");
    if (n > TRT_NOALLOC_MAX_PIXELS)
        return say(err, err_cap, TRT_ERR_INVALID,
                   "trt_quantise_noalloc: %ux%u = %llu pixels exceed the static scratch's limit of %u pixels (TRT_NOALLOC_MAX_PIXELS); "
                   "use a context (trt_create / trt_quantise_async) for larger frames",
                   width, height, (unsigned long long)n, (unsigned)TRT_NOALLOC_MAX_PIXELS);
    int prev = -1;
    cudaError_t e = cudaGetDevice(&prev);
    if (e == cudaSuccess && prev != device) e = cudaSetDevice(device);
    if (e != cudaSuccess) return say(err, err_cap, e == cudaErrorNoDevice ? TRT_ERR_NO_DEVICE : TRT_ERR_CUDA, "trt_quantise_noalloc: device %d: %s", device, cudaGetErrorString(e));
    int rc = TRT_OK;
    if (!g_scratch_ptr[device]) {
        unsigned long long *p = nullptr;
        e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) rc = say(err, err_cap, TRT_ERR_CUDA, "trt_quantise_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
        else g_scratch_ptr[device] = p;
    }
    if (rc == TRT_OK) {
        unsigned long long *scratch = g_scratch_ptr[device];
        cudaStream_t stream = (cudaStream_t)cuda_stream;
        QuantParams q{};
        q.src = (const float4 *)d_pixels; q.row_src = nullptr; q.width = width; q.height = height; q.mode = mode;
        q.n_blocks = (uint32_t)((n + TRT_Q_BLOCK - 1) / TRT_Q_BLOCK);
        q.out = (uint8_t *)d_out; q.out_cap = out_cap;
        q.n_tiles = (uint32_t)((n + TRT_QF_TILE - 1) / TRT_QF_TILE);
        q.ticket = (unsigned int *)scratch;
        q.tile_state = scratch + 1;
        q.row_start = nullptr;
        q.total = (uint64_t *)(scratch + 1 + kNoallocMaxTiles);
        const uint32_t clear_words = q.n_tiles + 1u;
        noalloc_clear_kernel<<<(clear_words + 255u) / 256u, 256, 0, stream>>>(clear_words);
        quant_fused_kernel<<<q.n_tiles, TRT_Q_BLOCK, 0, stream>>>(q);
        noalloc_publish_kernel<<<1, 1, 0, stream>>>(1u + kNoallocMaxTiles, (unsigned long long *)d_len);
        e = cudaGetLastError();
        if (e != cudaSuccess) rc = say(err, err_cap, TRT_ERR_CUDA, "trt_quantise_noalloc: launch: %s", cudaGetErrorString(e));
    }
    if (prev >= 0 && prev != device) cudaSetDevice(prev);
    return rc;
}
