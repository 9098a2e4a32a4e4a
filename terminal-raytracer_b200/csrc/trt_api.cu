// trt_api.cu -- the C ABI (include/trt_b200.h) over the CUDA kernels.  One context = one GPU.
// There is no CPU fallback: every compute entry point runs kernels or fails with a status code.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cmath>
#include <algorithm>
#include <atomic>
#include <mutex>
#include <string>
#include <vector>

#include "trt_kernels.cuh"
#include "trt_persistent.cuh"
#include "trt_wave.cuh"
#include "trt_native.cuh"
#include "trt_pool.cuh"
#include "trt_quantise.cuh"

using namespace trt;

// conv 2 lives in its own translation unit (trt_native.cu: --fmad=true -use_fast_math)
namespace trt {
cudaError_t native_blend_launch(float4 *cur, const uint2 *batch, uint32_t n_batch, uint32_t n_local_pixels, const uint32_t *row_map,
                                uint32_t width, float4 *accumulation, float4 *pixels, cudaStream_t stream);
bool pool_wanted(uint32_t n_work, int sm_count, int ctas_per_sm, bool strict);      // trt_native.cu
cudaError_t native_trace_launch(int device, const void *ctx_key, unsigned long long serial, const SceneView &sv,
                                const LightView &lv, const SmallScene *small, const FrameParams &fp, int sm_count,
                                cudaStream_t stream, uint32_t *launches, const char **kernel_name);
}

// The deferred-NEE machine (trt_native.cuh) in the strict conventions: conv 0 / 1, bit-exact against the oracle.
// TRT_B200_STRICT=wave selects the round-1 strict kernel (trace_wave_kernel) for A/B runs.
static cudaError_t strict_native_launch(int device, const SceneView &sv, const LightView &lv, const FrameParams &fp, int conv, int sm_count,
                                        bool constant, bool cornell, cudaStream_t stream, uint32_t *launches, const char **name) {
    cudaError_t e = cudaMemsetAsync(fp.work_counter, 0, sizeof(unsigned int), stream);
    if (e != cudaSuccess) return e;
    static bool carveout_set[64] = {};
    if (device >= 0 && device < 64 && !carveout_set[device]) {
        cudaFuncSetAttribute(trace_native_kernel<0, SrcConst, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<1, SrcConst, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<0, SrcGlobal, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<1, SrcGlobal, ShapeAny>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<0, SrcConst, ShapeCornellStrict>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_native_kernel<1, SrcConst, ShapeCornellStrict>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_pool_kernel<0, SrcConst, ShapeCornellStrict>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        cudaFuncSetAttribute(trace_pool_kernel<1, SrcConst, ShapeCornellStrict>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        carveout_set[device] = true;
    }
    const uint32_t n_work = fp.batch ? fp.n_local_pixels * fp.n_batch : fp.n_local_pixels;
    const uint32_t want = (n_work + TRT_NV_THREADS - 1u) / TRT_NV_THREADS;
    static const char *ctas_env = getenv("TRT_B200_STRICT_CTAS");
    uint32_t ctas = fp.ctas_hint >= 1u && fp.ctas_hint < TRT_NV_STRICT_CTAS ? fp.ctas_hint : TRT_NV_STRICT_CTAS;
    if (ctas_env && atoi(ctas_env) >= 1 && atoi(ctas_env) <= TRT_NV_STRICT_CTAS) ctas = (uint32_t)atoi(ctas_env);
    const uint32_t cap = (uint32_t)sm_count * ctas;
    const uint32_t blocks = want < cap ? want : cap;
    if (constant && cornell && pool_wanted(n_work, sm_count, TRT_PL_STRICT_CTAS, true)) {
        const uint32_t pcap = (uint32_t)sm_count * TRT_PL_STRICT_CTAS;
        const uint32_t pblocks = want < pcap ? want : pcap;
        if (conv) trace_pool_kernel<1, SrcConst, ShapeCornellStrict><<<pblocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_pool_kernel<0, SrcConst, ShapeCornellStrict><<<pblocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        *name = "trace_pool_kernel<CONV, SrcConst, ShapeCornell> (strict)";
    } else if (constant && cornell) {
        // the reference's default scene layout: the strict scan with its run list known to the compiler
        if (conv) trace_native_kernel<1, SrcConst, ShapeCornellStrict><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_native_kernel<0, SrcConst, ShapeCornellStrict><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        *name = "trace_native_kernel<CONV, SrcConst, ShapeCornell> (strict)";
    } else if (constant) {
        if (conv) trace_native_kernel<1, SrcConst, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_native_kernel<0, SrcConst, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        *name = "trace_native_kernel<CONV, SrcConst, ShapeAny> (strict)";
    } else {
        if (conv) trace_native_kernel<1, SrcGlobal, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        else trace_native_kernel<0, SrcGlobal, ShapeAny><<<blocks, TRT_NV_THREADS, 0, stream>>>(sv, lv, fp);
        *name = "trace_native_kernel<CONV, SrcGlobal, ShapeAny> (strict)";
    }
    *launches += 1;
    return cudaGetLastError();
}

struct trt_ctx {
    trt_config cfg{};
    int device = 0;
    int sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
    std::string err;

    // scene
    trt_primitive *d_aos = nullptr;
    float4 *d_geom = nullptr;          // ga|gb|gc|gn|mc|me, each cap_prims long
    uint8_t *d_type = nullptr;
    int4 *d_runs = nullptr;
    int4 *d_tiles = nullptr;           // (type, first, count, -) per tile of the cooperative scan
    size_t cap_tiles = 0;
    int4 *d_lights = nullptr;
    int *d_light_prim = nullptr;
    float4 *d_lgeom = nullptr;         // la|lb|lc|ln|le, each cap_lights long
    uint32_t cap_prims = 0, cap_lights = 0, cap_runs = 0;
    SceneView sv{};
    LightView lv{};
    SmallScene small{};                // canonical small-scene mirror for the conv 2 scan (host copy)
    // opt-in uniform grid (trt_config.accel)
    uint32_t *d_grid_off = nullptr, *d_grid_idx = nullptr;
    int *d_grid_planes = nullptr;
    bool have_scene = false;
    bool cornell_runs = false;         // the scene has the default Cornell layout incl. its run list (RunsCornell)
    uint64_t scene_serial = 0;

    // frame buffers (lib.rs:269-319: accumulation, output(pixels), variance; staging -> byte stream)
    size_t max_pixels = 0;
    float4 *d_pixels = nullptr, *d_accum = nullptr;
    trt_variance *d_variance = nullptr;
    unsigned long long *d_counters = nullptr;
    unsigned int *d_work = nullptr;
    unsigned long long *d_stats = nullptr;
    uint32_t *d_order = nullptr;       // hand-out order of the local pixels (longest first)
    uint32_t *d_cost = nullptr;        // bounces per pixel of the last frame (production kernel), valid when cost_valid
    bool cost_valid = false, cost_kernel = false;
    uint32_t order_w = 0, order_h = 0; // frame size the variance buffer was last written for (0: no history)
    uint32_t *d_row_map = nullptr;     // local row -> global row
    uint64_t *d_row_src = nullptr;     // quantiser source offsets per output row
    std::vector<uint32_t> rows;        // rows owned at `rows_height`
    uint32_t rows_height = 0, rows_width = 0;

    // quantiser
    uint32_t *d_packed = nullptr, *d_block_len = nullptr;
    uint64_t *d_block_off = nullptr;
    uint8_t *d_stream = nullptr;
    size_t stream_cap = 0;
    uint64_t *h_len = nullptr;         // pinned: [0] stream length, [2..3] lengths of the two pipelined frames, [4..5] frame statistics
    uint32_t last_q_blocks = 0;
    bool have_stream = false;
    int last_mode = TRT_MODE_NONE;
    // one-pass quantiser: tile states of the look-back (+ the ticket in front), row starts of the last stream
    unsigned long long *d_tile_state = nullptr;
    uint64_t *d_row_start = nullptr;
    bool fused_rows = false;           // d_row_start describes the last quantised stream

    // row-addressed stream: byte length of every row of the last quantised frame (local row order)
    uint32_t *d_row_len = nullptr;
    uint32_t *h_row_len = nullptr;     // pinned
    cudaEvent_t ev_row_len = nullptr;
    uint32_t row_len_rows = 0;
    bool have_row_len = false;
    // row table of trt_quantise_gathered (kept apart from this shard's own d_row_src), cached by layout
    uint64_t *d_row_src_g = nullptr;
    uint32_t g_w = 0, g_h = 0, g_shards = 0, g_stripe = 0;
    size_t g_stride = 0;

    // pipelined frames (trt_frame_submit / trt_frame_collect): a second device stream buffer, a copy stream
    cudaStream_t copy_stream = nullptr;
    uint8_t *d_stream2 = nullptr;
    cudaEvent_t ev_q[2] = {nullptr, nullptr};
    int pipe_mode[2] = {TRT_MODE_NONE, TRT_MODE_NONE};
    uint32_t pipe_submitted = 0, pipe_collected = 0;

    // frame batch (trt_render_frames)
    float4 *d_batch_cur = nullptr;     // [frames of one chunk][local pixels]
    uint2 *d_batch = nullptr;          // (seed, frame_number) per frame of the chunk
    size_t batch_cur_cap = 0;          // in float4
    uint32_t batch_cap = 0;
    uint8_t *d_batch_stream = nullptr; // [frames of one chunk][stride] byte streams when every frame is emitted
    size_t batch_stream_cap = 0;
    uint64_t *h_batch_len = nullptr;   // pinned, one length per frame of the chunk
    size_t batch_len_cap = 0;

    const char *kernel_name = "";      // trace kernel of the last frame (trt_trace_kernel_name)

    // work of the last single frame (samples, bounces: d_stats[0..1]) copied to h_len[4..5] behind the trace kernel;
    // read without waiting at the next launch to choose the number of resident CTAs (frame_ctas_hint)
    cudaEvent_t ev_stats = nullptr;
    bool stats_valid = false;
    uint32_t stats_w = 0, stats_h = 0, stats_pixels = 0, stats_spp = 0, stats_depth = 0;
    uint64_t stats_bounces = 0;        // bounce count of the latest completed frame of that description

    trt_uniforms last_u{};
    bool have_frame = false;
    trt_timings tm{};
    uint32_t launches = 0;
};

static thread_local std::string g_create_err;

#define CK(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) {                                                                      \
            ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                            \
            return TRT_ERR_CUDA;                                                                      \
        }                                                                                             \
    } while (0)

static int fail(trt_ctx *ctx, int code, const char *msg) {
    if (ctx) ctx->err = msg; else g_create_err = msg;
    return code;
}

extern "C" const char *trt_version(void) { return "terminal-raytracer_b200 0.1 (sm_100a)"; }

extern "C" const char *trt_last_error(const trt_ctx *ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

extern "C" void trt_destroy(trt_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    void *ptrs[] = {ctx->d_aos, ctx->d_geom, ctx->d_type, ctx->d_runs, ctx->d_tiles, ctx->d_lights, ctx->d_light_prim, ctx->d_lgeom,
                    ctx->d_accum, ctx->d_variance, ctx->d_counters, ctx->d_work, ctx->d_stats, ctx->d_row_map, ctx->d_order,
                    ctx->d_cost, ctx->d_row_src, ctx->d_packed, ctx->d_block_len, ctx->d_block_off, ctx->d_stream, ctx->d_batch_cur, ctx->d_batch, ctx->d_batch_stream,
                    ctx->d_row_len, ctx->d_row_src_g, ctx->d_stream2, ctx->d_tile_state, ctx->d_row_start, ctx->d_grid_off, ctx->d_grid_idx, ctx->d_grid_planes};
    for (void *p : ptrs) if (p) cudaFree(p);
    if (ctx->h_len) cudaFreeHost(ctx->h_len);
    if (ctx->h_row_len) cudaFreeHost(ctx->h_row_len);
    if (ctx->ev_row_len) cudaEventDestroy(ctx->ev_row_len);
    for (auto &e : ctx->ev_q) if (e) cudaEventDestroy(e);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->h_batch_len) cudaFreeHost(ctx->h_batch_len);
    for (auto &e : ctx->ev) if (e) cudaEventDestroy(e);
    if (ctx->ev_stats) cudaEventDestroy(ctx->ev_stats);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

extern "C" int trt_create(const trt_config *cfg, trt_ctx **out) {
    if (!cfg || !out) return fail(nullptr, TRT_ERR_INVALID, "trt_create: null argument");
    *out = nullptr;
    if (cfg->abi_version != TRT_ABI_VERSION) return fail(nullptr, TRT_ERR_INVALID, "trt_create: ABI version mismatch");
    if (cfg->max_width == 0 || cfg->max_height == 0) return fail(nullptr, TRT_ERR_INVALID, "trt_create: zero frame size");
    if (cfg->conv > 2u) return fail(nullptr, TRT_ERR_INVALID, "trt_create: unknown arithmetic convention");
    if (cfg->accel > TRT_ACCEL_GRID) return fail(nullptr, TRT_ERR_INVALID, "trt_create: unknown acceleration structure");
    if (cfg->accel != TRT_ACCEL_NONE && (cfg->conv != TRT_CONV_NATIVE || cfg->count_work))
        return fail(nullptr, TRT_ERR_INVALID, "trt_create: the acceleration structure is an opt-in of the native convention (the strict "
                                              "builds and the work counters follow the reference's linear scan)");
    if (cfg->shard_count > 1u && cfg->shard_index >= cfg->shard_count)
        return fail(nullptr, TRT_ERR_INVALID, "trt_create: shard_index >= shard_count");
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0) {
        g_create_err = std::string("trt_create: no CUDA device (") + cudaGetErrorString(e) + "); there is no CPU fallback";
        cudaGetLastError();
        return TRT_ERR_NO_DEVICE;
    }
    if (cfg->device < 0 || cfg->device >= n_dev) return fail(nullptr, TRT_ERR_INVALID, "trt_create: bad device ordinal");
    trt_ctx *ctx = new (std::nothrow) trt_ctx();
    if (!ctx) return fail(nullptr, TRT_ERR_NOMEM, "trt_create: out of host memory");
    ctx->cfg = *cfg;
    if (ctx->cfg.shard_count == 0) ctx->cfg.shard_count = 1;
    if (ctx->cfg.stripe_rows == 0) ctx->cfg.stripe_rows = 1;
    ctx->device = cfg->device;
    int rc = [&]() -> int {
        CK(cudaSetDevice(ctx->device));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, ctx->device));
        ctx->sm_count = prop.multiProcessorCount;
        CK(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        for (auto &ev : ctx->ev) CK(cudaEventCreate(&ev));
        ctx->max_pixels = (size_t)cfg->max_width * cfg->max_height;
        // `pixels` is a copy of `accumulation` after every frame (shader.wgsl:621 `pixels[index] = accumulation[index]`; the
        // reference keeps two buffers only because it copies one of them to the staging buffer, lib.rs:461-463): ONE device
        // buffer serves both names -- 16 B per pixel less to write, 132.7 MB less HBM at C3
        CK(cudaMalloc(&ctx->d_accum, ctx->max_pixels * sizeof(float4)));
        CK(cudaMalloc(&ctx->d_variance, ctx->max_pixels * sizeof(trt_variance)));
        CK(cudaMemsetAsync(ctx->d_accum, 0, ctx->max_pixels * sizeof(float4), ctx->stream));
        ctx->d_pixels = ctx->d_accum;
        CK(cudaMemsetAsync(ctx->d_variance, 0, ctx->max_pixels * sizeof(trt_variance), ctx->stream));
        CK(cudaMalloc(&ctx->d_counters, K_COUNT * sizeof(unsigned long long)));
        CK(cudaMemsetAsync(ctx->d_counters, 0, K_COUNT * sizeof(unsigned long long), ctx->stream));
        CK(cudaMalloc(&ctx->d_work, 64 + 2 * TRT_ORDER_BUCKETS * sizeof(unsigned int)));
        CK(cudaMalloc(&ctx->d_stats, TRT_STATS_BYTES));
        CK(cudaMemsetAsync(ctx->d_stats, 0, TRT_STATS_BYTES, ctx->stream));
        CK(cudaMalloc(&ctx->d_order, ctx->max_pixels * sizeof(uint32_t)));
        CK(cudaMalloc(&ctx->d_cost, ctx->max_pixels * sizeof(uint32_t)));
        CK(cudaMalloc(&ctx->d_row_map, (size_t)cfg->max_height * sizeof(uint32_t)));
        CK(cudaMalloc(&ctx->d_row_src, (size_t)cfg->max_height * sizeof(uint64_t)));
        CK(cudaMalloc(&ctx->d_row_src_g, (size_t)cfg->max_height * sizeof(uint64_t)));
        CK(cudaMalloc(&ctx->d_row_len, (size_t)cfg->max_height * sizeof(uint32_t)));
        CK(cudaMallocHost(&ctx->h_row_len, (size_t)cfg->max_height * sizeof(uint32_t)));
        CK(cudaEventCreateWithFlags(&ctx->ev_row_len, cudaEventDisableTiming));
        CK(cudaMalloc(&ctx->d_packed, ctx->max_pixels * sizeof(uint32_t)));
        const size_t nb = (ctx->max_pixels + TRT_Q_BLOCK - 1) / TRT_Q_BLOCK;
        CK(cudaMalloc(&ctx->d_block_len, nb * sizeof(uint32_t)));
        CK(cudaMalloc(&ctx->d_block_off, (nb + 1) * sizeof(uint64_t)));
        CK(cudaMalloc(&ctx->d_tile_state, ((ctx->max_pixels + TRT_QF_TILE - 1) / TRT_QF_TILE + 2) * sizeof(unsigned long long)));
        CK(cudaMalloc(&ctx->d_row_start, ((size_t)cfg->max_height + 1) * sizeof(uint64_t)));
        ctx->stream_cap = ctx->max_pixels * 26 + (size_t)cfg->max_height * 2 + 64;
        CK(cudaMalloc(&ctx->d_stream, ctx->stream_cap));
        CK(cudaMallocHost(&ctx->h_len, 64));
        memset(ctx->h_len, 0, 64);
        CK(cudaStreamSynchronize(ctx->stream));
        return TRT_OK;
    }();
    if (rc != TRT_OK) {
        g_create_err = ctx->err;
        trt_destroy(ctx);
        return rc;
    }
    *out = ctx;
    return TRT_OK;
}

// ---- scene ------------------------------------------------------------------------------------------
extern "C" int trt_upload_scene(trt_ctx *ctx, const trt_primitive *prims, uint32_t count) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!prims && count) return fail(ctx, TRT_ERR_INVALID, "trt_upload_scene: null primitives");
    CK(cudaSetDevice(ctx->device));
    // host-side index structures: type runs and the light list, in scan order (lib.rs:122-154 push order)
    std::vector<uint8_t> types(count);
    std::vector<int4> runs;
    std::vector<int> light_prim;
    for (uint32_t i = 0; i < count; i++) {
        const uint32_t t = prims[i].primitive_type;
        types[i] = t <= 2u ? (uint8_t)t : (uint8_t)255;
        if (t == 1u) {
            // axis-aligned plane: the normal is exactly +-1 on one axis and +-0 on the others and the point
            // is below 1e30 in magnitude (p - o cannot overflow); then dot(n, d) == +-d_axis and dot(p - o, n) == +-(p_axis - o_axis) bit for bit
            // for finite rays (scan_lean, trt_trace.cuh).  Device-side run types 3, 4, 5 = x, y, z.
            const float n[3] = {prims[i].plane_normal.x, prims[i].plane_normal.y, prims[i].plane_normal.z};
            const float p[3] = {prims[i].plane_point.x, prims[i].plane_point.y, prims[i].plane_point.z};
            int ones = 0, zeros = 0, axis = 0;
            for (int c = 0; c < 3; c++) {
                if (n[c] == 1.0f || n[c] == -1.0f) { ones++; axis = c; }
                else if (n[c] == 0.0f) zeros++;
            }
            if (ones == 1 && zeros == 2 && std::fabs(p[0]) < 1e30f && std::fabs(p[1]) < 1e30f && std::fabs(p[2]) < 1e30f)
                types[i] = (uint8_t)(3 + axis);
        }
        // shader.wgsl:346-347 light test, :352/:370 supported light types
        const float power = prims[i].emission.x + prims[i].emission.y + prims[i].emission.z;
        if (power > 0.001f && (t == 0u || t == 2u)) light_prim.push_back((int)i);
    }
    // slab pairs: two consecutive axis-aligned planes of the same axis at different positions (device run
    // types 6, 7, 8); gc of the first plane holds (p_lo, p_hi, first-plane-is-p_hi, 2^20 * gap) for scan_lean
    struct PairInfo { uint32_t index; float4 v; };
    std::vector<PairInfo> pairs;
    for (uint32_t i = 0; i + 1 < count; i++) {
        if (types[i] < 3 || types[i] > 5 || types[i + 1] != types[i]) continue;
        const int axis = types[i] - 3;
        const float p0 = (&prims[i].plane_point.x)[axis], p1 = (&prims[i + 1].plane_point.x)[axis];
        if (p0 == p1) continue;
        const float lo = p0 < p1 ? p0 : p1, hi = p0 < p1 ? p1 : p0;
        pairs.push_back({i, make_float4(lo, hi, p0 < p1 ? 0.0f : 1.0f, 1048576.0f * (hi - lo))});
        types[i] = types[i + 1] = (uint8_t)(6 + axis);
        i++;
    }
    for (uint32_t i = 0; i < count; i++) {
        // a pair run must start on a pair boundary: never merge a pair run into a preceding one of odd length
        if (runs.empty() || runs.back().x != (int)types[i]) runs.push_back(make_int4((int)types[i], (int)i, (int)i + 1, 0));
        else runs.back().z = (int)i + 1;
    }
    // canonical small scene (conv 2 unrolled scan): spheres, planes, triangles in this order, few of each
    {
        SmallScene sm{};
        uint32_t ns = 0, np = 0, nt = 0;
        bool canon = true;
        for (uint32_t i = 0; i < count && canon; i++) {
            const uint32_t t = prims[i].primitive_type;
            if (t == 0u) { canon = (np == 0 && nt == 0); ns++; }
            else if (t == 1u) { canon = (nt == 0); np++; }
            else if (t == 2u) nt++;
            else canon = false;
        }
        if (canon && ns <= TRT_SMALL_S && np <= TRT_SMALL_P && nt <= TRT_SMALL_T) {
            sm.jit_ok = 1;
            sm.ok = ns <= TRT_STATIC_S && np <= TRT_STATIC_P && nt <= TRT_STATIC_T;
            sm.n_sph = (int)ns; sm.n_pl = (int)np; sm.n_tri = (int)nt;
            for (uint32_t j = 0; j < ns; j++) {
                const trt_primitive &p = prims[j];
                sm.sph[j] = make_float4(p.sphere_center.x, p.sphere_center.y, p.sphere_center.z, p.sphere_radius);
                sm.sph_r2[j] = make_float4(p.sphere_radius * p.sphere_radius, 0.0f, 0.0f, 0.0f);
            }
            for (uint32_t k = 0; 2 * k + 1 < ns; k++) {
                const trt_primitive &a = prims[2 * k], &b = prims[2 * k + 1];
                sm.sph_pair[3 * k] = make_float4(a.sphere_center.x, b.sphere_center.x, a.sphere_center.y, b.sphere_center.y);
                sm.sph_pair[3 * k + 1] = make_float4(a.sphere_center.z, b.sphere_center.z, a.sphere_radius, b.sphere_radius);
                sm.sph_pair[3 * k + 2] = make_float4(a.sphere_radius * a.sphere_radius, b.sphere_radius * b.sphere_radius, 0.0f, 0.0f);
            }
            for (uint32_t j = 0; j < np; j++) {
                const trt_primitive &p = prims[ns + j];
                const uint8_t ty = types[ns + j];
                const int kind = ty >= 3 ? (ty - 3) % 3 : 3;
                sm.pl_p[j] = make_float4(p.plane_point.x, p.plane_point.y, p.plane_point.z, 0.0f);
                memcpy(&sm.pl_p[j].w, &kind, 4);
                sm.pl_n[j] = make_float4(p.plane_normal.x, p.plane_normal.y, p.plane_normal.z, 0.0f);
            }
            for (uint32_t k = 0; 2 * k + 1 < np; k++) {          // two walls on one axis: their coordinates side by side
                int ka, kb;
                memcpy(&ka, &sm.pl_p[2 * k].w, 4); memcpy(&kb, &sm.pl_p[2 * k + 1].w, 4);
                const float4 a = sm.pl_p[2 * k], b = sm.pl_p[2 * k + 1];
                sm.pl_pair[k] = ka == kb && ka < 3 ? make_float2(ka == 0 ? a.x : ka == 1 ? a.y : a.z, ka == 0 ? b.x : ka == 1 ? b.y : b.z)
                                                   : make_float2(0.0f, 0.0f);
            }
            for (uint32_t j = 0; j < nt; j++) {
                const trt_primitive &p = prims[ns + np + j];
                sm.tri_v0[j] = make_float4(p.triangle_v0.x, p.triangle_v0.y, p.triangle_v0.z, 0.0f);
                sm.tri_e1[j] = make_float4(p.triangle_v1.x - p.triangle_v0.x, p.triangle_v1.y - p.triangle_v0.y, p.triangle_v1.z - p.triangle_v0.z, 0.0f);
                sm.tri_e2[j] = make_float4(p.triangle_v2.x - p.triangle_v0.x, p.triangle_v2.y - p.triangle_v0.y, p.triangle_v2.z - p.triangle_v0.z, 0.0f);
            }
        }
        for (int j = 0; j < sm.n_tri && j < 5 && sm.jit_ok; j++) {
            const float e[6] = {sm.tri_e1[j].x, sm.tri_e1[j].y, sm.tri_e1[j].z, sm.tri_e2[j].x, sm.tri_e2[j].y, sm.tri_e2[j].z};
            for (int b = 0; b < 6; b++)
                if (e[b] == 0.0f) sm.tri_zero |= 1u << (6 * j + b);
        }
        sm.n_lights = (int)light_prim.size();
        // triangle lights 2k and 2k + 1 side by side for the packed reach test of the bounce (trt_native.cuh): unit normal,
        // plane offset dot(n, v0) and its magnitude, with the operation sequence of repack_kernel<0>
        for (size_t k = 0; 2 * k + 1 < light_prim.size() && k < TRT_SMALL_LPAIRS; k++) {
            float nx[2] = {0, 0}, ny[2] = {0, 0}, nz[2] = {0, 0}, off[2] = {0, 0};
            for (int h = 0; h < 2; h++) {
                const trt_primitive &p = prims[light_prim[2 * k + h]];
                if (p.primitive_type != 2u) continue;
                const float e1[3] = {p.triangle_v1.x - p.triangle_v0.x, p.triangle_v1.y - p.triangle_v0.y, p.triangle_v1.z - p.triangle_v0.z};
                const float e2[3] = {p.triangle_v2.x - p.triangle_v0.x, p.triangle_v2.y - p.triangle_v0.y, p.triangle_v2.z - p.triangle_v0.z};
                const float cx = e1[1] * e2[2] - e1[2] * e2[1], cy = e1[2] * e2[0] - e1[0] * e2[2], cz = e1[0] * e2[1] - e1[1] * e2[0];
                const float len = sqrtf(cx * cx + cy * cy + cz * cz);
                nx[h] = cx / len; ny[h] = cy / len; nz[h] = cz / len;
                off[h] = nx[h] * p.triangle_v0.x + ny[h] * p.triangle_v0.y + nz[h] * p.triangle_v0.z;
            }
            sm.lt_pair[3 * k] = make_float4(nx[0], nx[1], ny[0], ny[1]);
            sm.lt_pair[3 * k + 1] = make_float4(nz[0], nz[1], off[0], off[1]);
            sm.lt_pair[3 * k + 2] = make_float4(fabsf(off[0]), fabsf(off[1]), 0.0f, 0.0f);
        }
        for (size_t k = 0; k < light_prim.size() && k < 32; k++)
            if (prims[light_prim[k]].primitive_type == 0u) { sm.light_spheres |= 1 << k; sm.n_light_spheres++; }
        ctx->small = sm;
    }
    // tiles of the CTA-cooperative scan: at most TRT_TILE consecutive primitives of one run, in scan order
    std::vector<int4> tiles;
    for (const int4 &r : runs)
        for (int t0 = r.y; t0 < r.z; t0 += TRT_TILE) tiles.push_back(make_int4(r.x, t0, std::min(TRT_TILE, r.z - t0), 0));
    {   // triangle and sphere tiles: where their pair-packed cull entries start in gd (two entries per pair of primitives)
        int next = 0;
        for (int4 &t : tiles)
            if (t.x == 2 || t.x == 0) { t.w = next; next += 2 * ((t.z + 1) / 2); }
    }
    const uint32_t n_lights = (uint32_t)light_prim.size(), n_runs = (uint32_t)runs.size();
    CK(cudaStreamSynchronize(ctx->stream));
    if (count > ctx->cap_prims) {
        if (ctx->d_aos) { cudaFree(ctx->d_aos); cudaFree(ctx->d_geom); cudaFree(ctx->d_type); }
        ctx->d_aos = nullptr; ctx->d_geom = nullptr; ctx->d_type = nullptr; ctx->cap_prims = 0;
        const uint32_t cap = count < 64u ? 64u : count;
        CK(cudaMalloc(&ctx->d_aos, (size_t)cap * sizeof(trt_primitive)));
        CK(cudaMalloc(&ctx->d_geom, (size_t)cap * 8 * sizeof(float4)));     // ga gb gc gn mc me + gd (<= count + tiles entries)
        CK(cudaMalloc(&ctx->d_type, cap));
        ctx->cap_prims = cap;
    }
    if (n_lights > ctx->cap_lights || !ctx->d_lgeom) {
        if (ctx->d_lgeom) { cudaFree(ctx->d_lgeom); cudaFree(ctx->d_light_prim); cudaFree(ctx->d_lights); }
        ctx->d_lgeom = nullptr; ctx->d_light_prim = nullptr; ctx->d_lights = nullptr; ctx->cap_lights = 0;
        const uint32_t cap = n_lights < 16u ? 16u : n_lights;
        CK(cudaMalloc(&ctx->d_lgeom, (size_t)cap * 5 * sizeof(float4)));
        CK(cudaMalloc(&ctx->d_light_prim, (size_t)cap * sizeof(int)));
        CK(cudaMalloc(&ctx->d_lights, (size_t)cap * sizeof(int4)));
        ctx->cap_lights = cap;
    }
    if (n_runs > ctx->cap_runs || !ctx->d_runs) {
        if (ctx->d_runs) cudaFree(ctx->d_runs);
        ctx->d_runs = nullptr; ctx->cap_runs = 0;
        const uint32_t cap = n_runs < 16u ? 16u : n_runs;
        CK(cudaMalloc(&ctx->d_runs, (size_t)cap * sizeof(int4)));
        ctx->cap_runs = cap;
    }
    if (tiles.size() > ctx->cap_tiles || !ctx->d_tiles) {
        if (ctx->d_tiles) cudaFree(ctx->d_tiles);
        ctx->d_tiles = nullptr; ctx->cap_tiles = 0;
        const size_t cap = tiles.size() < 16 ? 16 : tiles.size();
        CK(cudaMalloc(&ctx->d_tiles, cap * sizeof(int4)));
        ctx->cap_tiles = cap;
    }
    if (!tiles.empty())
        CK(cudaMemcpyAsync(ctx->d_tiles, tiles.data(), tiles.size() * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream));
    if (count) {
        CK(cudaMemcpyAsync(ctx->d_aos, prims, (size_t)count * sizeof(trt_primitive), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->d_type, types.data(), count, cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaMemcpyAsync(ctx->d_runs, runs.data(), n_runs * sizeof(int4), cudaMemcpyHostToDevice, ctx->stream));
    }
    if (n_lights)
        CK(cudaMemcpyAsync(ctx->d_light_prim, light_prim.data(), n_lights * sizeof(int), cudaMemcpyHostToDevice, ctx->stream));
    const size_t cp = ctx->cap_prims, cl = ctx->cap_lights;
    RepackParams rp{};
    rp.aos = ctx->d_aos; rp.n_prims = (int)count;
    rp.ga = ctx->d_geom; rp.gb = ctx->d_geom + cp; rp.gc = ctx->d_geom + 2 * cp; rp.gn = ctx->d_geom + 3 * cp;
    rp.mc = ctx->d_geom + 4 * cp; rp.me = ctx->d_geom + 5 * cp;
    rp.light_prim = ctx->d_light_prim; rp.n_lights = (int)n_lights;
    rp.la = ctx->d_lgeom; rp.lb = ctx->d_lgeom + cl; rp.lc = ctx->d_lgeom + 2 * cl; rp.ln = ctx->d_lgeom + 3 * cl;
    rp.le = ctx->d_lgeom + 4 * cl;
    const uint32_t work = count > n_lights ? count : n_lights;
    if (work) {
        // conv 2 keeps the as-written derived values (unit normals, edges, r^2): only the per-ray math is native
        if (ctx->cfg.conv == 1u) repack_kernel<1><<<(work + 127) / 128, 128, 0, ctx->stream>>>(rp);
        else repack_kernel<0><<<(work + 127) / 128, 128, 0, ctx->stream>>>(rp);
        CK(cudaGetLastError());
    }
    if (!tiles.empty() && count) {
        pack_cull_kernel<<<(unsigned)tiles.size(), 64, 0, ctx->stream>>>(rp.ga, ctx->d_tiles, (int)tiles.size(), ctx->d_geom + 6 * cp);
        CK(cudaGetLastError());
    }
    for (const PairInfo &pi : pairs)
        CK(cudaMemcpyAsync(rp.gc + pi.index, &pi.v, sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    ctx->sv.ga = rp.ga; ctx->sv.gb = rp.gb; ctx->sv.gc = rp.gc; ctx->sv.gn = rp.gn; ctx->sv.mc = rp.mc; ctx->sv.me = rp.me;
    ctx->sv.gd = ctx->d_geom + 6 * cp;
    ctx->sv.type = ctx->d_type; ctx->sv.runs = ctx->d_runs; ctx->sv.lights = ctx->d_lights;
    ctx->sv.n_prims = (int)count; ctx->sv.n_runs = (int)n_runs; ctx->sv.n_lights = (int)n_lights;
    ctx->sv.tiles = ctx->d_tiles; ctx->sv.n_tiles = (int)tiles.size();
    // ---- opt-in uniform grid: the reference's build_uniform_grid (lib.rs:181-244) over spheres and triangles -------
    ctx->sv.grid_off = ctx->sv.grid_idx = nullptr; ctx->sv.grid_planes = nullptr; ctx->sv.n_grid_planes = 0;
    if (ctx->cfg.accel == TRT_ACCEL_GRID && count) {
        auto aabb = [&](const trt_primitive &p, float mn[3], float mx[3]) -> bool {      // lib.rs:158-179
            if (p.primitive_type == 0u) {
                const float c[3] = {p.sphere_center.x, p.sphere_center.y, p.sphere_center.z}, r = fabsf(p.sphere_radius);
                for (int a = 0; a < 3; a++) { mn[a] = c[a] - r; mx[a] = c[a] + r; }
                return true;
            }
            if (p.primitive_type == 2u) {
                const float v[3][3] = {{p.triangle_v0.x, p.triangle_v0.y, p.triangle_v0.z}, {p.triangle_v1.x, p.triangle_v1.y, p.triangle_v1.z},
                                       {p.triangle_v2.x, p.triangle_v2.y, p.triangle_v2.z}};
                for (int a = 0; a < 3; a++) { mn[a] = fminf(v[0][a], fminf(v[1][a], v[2][a])); mx[a] = fmaxf(v[0][a], fmaxf(v[1][a], v[2][a])); }
                return true;
            }
            return false;
        };
        float bmin[3] = {INFINITY, INFINITY, INFINITY}, bmax[3] = {-INFINITY, -INFINITY, -INFINITY};
        std::vector<int> plane_list;
        uint32_t bounded = 0;
        for (uint32_t i = 0; i < count; i++) {
            float mn[3], mx[3];
            if (aabb(prims[i], mn, mx)) { bounded++; for (int a = 0; a < 3; a++) { bmin[a] = fminf(bmin[a], mn[a]); bmax[a] = fmaxf(bmax[a], mx[a]); } }
            else if (prims[i].primitive_type == 1u) plane_list.push_back((int)i);
        }
        if (bounded) {
            for (int a = 0; a < 3; a++) { bmin[a] -= 1e-3f; bmax[a] += 1e-3f; }                         // lib.rs:191-193
            const float ext[3] = {bmax[0] - bmin[0], bmax[1] - bmin[1], bmax[2] - bmin[2]};
            const float emax = fmaxf(ext[0], fmaxf(ext[1], ext[2]));
            const float sres = cbrtf((float)bounded) * 1.5f;                                            // lib.rs:196-198
            int dims[3];
            for (int a = 0; a < 3; a++) { dims[a] = (int)lroundf(fmaxf(sres * (ext[a] / emax), 1.0f)); if (dims[a] < 1) dims[a] = 1; if (dims[a] > 256) dims[a] = 256; }
            const float cell[3] = {ext[0] / dims[0], ext[1] / dims[1], ext[2] / dims[2]};
            const size_t n_cells = (size_t)dims[0] * dims[1] * dims[2];
            std::vector<uint32_t> off(n_cells + 1, 0);
            auto range = [&](const float mn[3], const float mx[3], int lo[3], int hi[3]) {
                for (int a = 0; a < 3; a++) {
                    lo[a] = (int)fminf(fmaxf(floorf((mn[a] - bmin[a]) / cell[a]), 0.0f), (float)(dims[a] - 1));
                    hi[a] = (int)fminf(fmaxf(floorf((mx[a] - bmin[a]) / cell[a]), 0.0f), (float)(dims[a] - 1));
                }
            };
            for (int pass = 0; pass < 2; pass++) {                       // count, then fill (primitive order = scan order)
                std::vector<uint32_t> cur;
                std::vector<uint32_t> idx;
                if (pass == 1) {
                    uint32_t run = 0;
                    for (size_t c = 0; c < n_cells; c++) { const uint32_t n = off[c]; off[c] = run; run += n; }
                    off[n_cells] = run;
                    cur.assign(off.begin(), off.end() - 1);
                    idx.resize(run);
                }
                for (uint32_t i = 0; i < count; i++) {
                    float mn[3], mx[3];
                    if (!aabb(prims[i], mn, mx)) continue;
                    int lo[3], hi[3];
                    range(mn, mx, lo, hi);
                    for (int z = lo[2]; z <= hi[2]; z++)
                        for (int y = lo[1]; y <= hi[1]; y++)
                            for (int x = lo[0]; x <= hi[0]; x++) {
                                const size_t c = (size_t)x + (size_t)dims[0] * ((size_t)y + (size_t)dims[1] * z);
                                if (pass == 0) off[c]++; else idx[cur[c]++] = i;
                            }
                }
                if (pass == 1) {
                    if (ctx->d_grid_off) cudaFree(ctx->d_grid_off);
                    if (ctx->d_grid_idx) cudaFree(ctx->d_grid_idx);
                    if (ctx->d_grid_planes) cudaFree(ctx->d_grid_planes);
                    ctx->d_grid_off = ctx->d_grid_idx = nullptr; ctx->d_grid_planes = nullptr;
                    CK(cudaMalloc(&ctx->d_grid_off, off.size() * sizeof(uint32_t)));
                    CK(cudaMalloc(&ctx->d_grid_idx, (idx.size() + 1) * sizeof(uint32_t)));
                    CK(cudaMalloc(&ctx->d_grid_planes, (plane_list.size() + 1) * sizeof(int)));
                    CK(cudaMemcpy(ctx->d_grid_off, off.data(), off.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
                    if (!idx.empty()) CK(cudaMemcpy(ctx->d_grid_idx, idx.data(), idx.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
                    if (!plane_list.empty()) CK(cudaMemcpy(ctx->d_grid_planes, plane_list.data(), plane_list.size() * sizeof(int), cudaMemcpyHostToDevice));
                }
            }
            ctx->sv.grid_off = ctx->d_grid_off; ctx->sv.grid_idx = ctx->d_grid_idx;
            ctx->sv.grid_planes = ctx->d_grid_planes; ctx->sv.n_grid_planes = (int)plane_list.size();
            ctx->sv.gx = dims[0]; ctx->sv.gy = dims[1]; ctx->sv.gz = dims[2];
            ctx->sv.grid_min = make_float3(bmin[0], bmin[1], bmin[2]);
            ctx->sv.grid_cell = make_float3(cell[0], cell[1], cell[2]);
            ctx->sv.grid_inv_cell = make_float3(1.0f / cell[0], 1.0f / cell[1], 1.0f / cell[2]);
        }
    }
    {   // default Cornell layout (ShapeCornell) with exactly the run list RunsCornell: the strict specialised kernel
        bool c = ctx->small.ok && ctx->small.n_sph == ShapeCornell::NS && ctx->small.n_pl == ShapeCornell::NP &&
                 ctx->small.n_tri == ShapeCornell::NT && n_lights == (uint32_t)ShapeCornell::NL && ctx->small.n_light_spheres == 0 &&
                 (int)runs.size() == RunsCornell::kCount;
        for (int j = 0; c && j < ShapeCornell::NP; j++) {
            int kind;
            memcpy(&kind, &ctx->small.pl_p[j].w, 4);
            c = kind == ShapeCornell::kind(j);
        }
        for (int r = 0; c && r < RunsCornell::kCount; r++) {
            const int4 e = RunsCornell::run(r);
            c = runs[r].x == e.x && runs[r].y == e.y && runs[r].z == e.z;
        }
        ctx->cornell_runs = c;
    }
    ctx->lv.la = rp.la; ctx->lv.lb = rp.lb; ctx->lv.lc = rp.lc; ctx->lv.ln = rp.ln; ctx->lv.le = rp.le;
    CK(cudaStreamSynchronize(ctx->stream));
    ctx->have_scene = true;
    static std::atomic<unsigned long long> g_scene_serial{0};     // process-wide: a recycled ctx address never aliases
    ctx->scene_serial = ++g_scene_serial;
    return TRT_OK;
}

// ---- frame -------------------------------------------------------------------------------------------
// rows owned by this shard at frame height h: stripes of stripe_rows rows dealt round-robin
static void shard_rows(const trt_config &c, uint32_t h, std::vector<uint32_t> &rows) {
    rows.clear();
    for (uint32_t y = 0; y < h; y++) {
        const uint32_t stripe = y / c.stripe_rows;
        if (c.shard_count <= 1u || stripe % c.shard_count == c.shard_index) rows.push_back(y);
    }
}

extern "C" uint32_t trt_shard_rows_of(uint32_t shard_index, uint32_t shard_count, uint32_t stripe_rows, uint32_t height,
                                      uint32_t *rows, uint32_t cap) {
    trt_config c{};
    c.shard_index = shard_index;
    c.shard_count = shard_count ? shard_count : 1u;
    c.stripe_rows = stripe_rows ? stripe_rows : 1u;
    std::vector<uint32_t> r;
    shard_rows(c, height, r);
    for (uint32_t i = 0; i < r.size() && i < cap && rows; i++) rows[i] = r[i];
    return (uint32_t)r.size();
}

extern "C" uint32_t trt_shard_rows(const trt_ctx *ctx, uint32_t height, uint32_t *rows, uint32_t cap) {
    if (!ctx) return 0;
    std::vector<uint32_t> r;
    shard_rows(ctx->cfg, height, r);
    for (uint32_t i = 0; i < r.size() && i < cap && rows; i++) rows[i] = r[i];
    return (uint32_t)r.size();
}

static int ensure_rows(trt_ctx *ctx, uint32_t w, uint32_t h) {
    if (ctx->rows_height == h && ctx->rows_width == w) return TRT_OK;
    shard_rows(ctx->cfg, h, ctx->rows);
    std::vector<uint64_t> src(ctx->rows.size());
    for (size_t i = 0; i < ctx->rows.size(); i++) src[i] = (uint64_t)ctx->rows[i] * w;
    CK(cudaStreamSynchronize(ctx->stream));
    if (!ctx->rows.empty()) {
        CK(cudaMemcpy(ctx->d_row_map, ctx->rows.data(), ctx->rows.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
        CK(cudaMemcpy(ctx->d_row_src, src.data(), src.size() * sizeof(uint64_t), cudaMemcpyHostToDevice));
    }
    ctx->rows_height = h;
    ctx->rows_width = w;
    return TRT_OK;
}

static int check_uniforms(trt_ctx *ctx, const trt_uniforms *u) {
    if (!u) return fail(ctx, TRT_ERR_INVALID, "null uniforms");
    if (!ctx->have_scene) return fail(ctx, TRT_ERR_NO_SCENE, "render before trt_upload_scene");
    if (u->width > 65535u || u->height > 65535u) return fail(ctx, TRT_ERR_INVALID, "uniforms: frame larger than 65535 in a dimension");
    if (u->width == 0 || u->height == 0 || u->width > ctx->cfg.max_width || u->height > ctx->cfg.max_height ||
        (size_t)u->width * u->height > ctx->max_pixels)
        return fail(ctx, TRT_ERR_INVALID, "uniforms: frame size is zero or exceeds trt_config.max_width/max_height");
    if (u->primitive_count > (uint32_t)ctx->sv.n_prims)
        return fail(ctx, TRT_ERR_INVALID, "uniforms: primitive_count exceeds the uploaded scene");
    return TRT_OK;
}

// One launch of the trace kernel for the frame `u` describes; with n_batch > 0 for the n_batch frames whose
// (seed, frame_number) pairs are in ctx->d_batch (means into ctx->d_batch_cur, blended by the caller).
// Resident CTAs per SM for this frame.  A pixel is one serial chain (its random-number state runs through all its
// samples, shader.wgsl:517/:565): no frame can be shorter than its longest pixel, and a chain advances at the SM's
// issue rate divided by the resident warps.  When the frame is long against the longest possible chain (spp x
// max_depth bounces) every CTA that fits is worth its throughput (C3 on one GPU: 8 CTAs 760 ms, 7: 769 ms); when a
// shard of the frame gets short, fewer resident warps make the critical chains faster for a smaller loss (a quarter
// of C3: 8 CTAs 219 ms, 7: 199 ms; an eighth: 108 / 101 ms; a sixteenth: 61 / 59 / 55 ms with 6).  R = longest
// possible chain / bounces per resident lane, from the previous frame's bounce count; no history: all that fit.
static uint32_t frame_ctas_hint(trt_ctx *ctx, const FrameParams &fp) {
    if (!ctx->stats_valid || ctx->stats_w != fp.width || ctx->stats_h != fp.height || ctx->stats_pixels != fp.n_local_pixels ||
        ctx->stats_spp != fp.spp || ctx->stats_depth != fp.max_depth) return 0u;
    // the latest frame whose statistics have arrived (frames may be queued faster than they complete)
    if (cudaEventQuery(ctx->ev_stats) == cudaSuccess) ctx->stats_bounces = ctx->h_len[5];
    else cudaGetLastError();
    const double bounces = (double)ctx->stats_bounces;
    if (!(bounces > 0.0)) return 0u;
    const double lanes = (double)ctx->sm_count * 8.0 * 128.0;
    const double r = (double)fp.spp * (double)(fp.max_depth ? fp.max_depth : 1u) * lanes / bounces;
    return r < 0.5 ? 0u : r < 2.5 ? 7u : 6u;
}

// State shared by every context of the process: the per-device __constant__ scene mirrors with their owners
// (trt_persistent.cuh, trt_native.cu) and the cache of run-time compiled kernels (trt_jit.cuh).  Binding a scene and
// launching its frame is one critical section, so contexts owned by different host threads can render side by side
// (a context itself stays single-owner, include/trt_b200.h); the kernels run asynchronously outside the lock.
static std::mutex g_device_state_mutex;

static int trace_impl(trt_ctx *ctx, const trt_uniforms *u, uint32_t n_batch) {
    int rc = check_uniforms(ctx, u);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    if (u->primitive_count != (uint32_t)ctx->sv.n_prims)
        return fail(ctx, TRT_ERR_INVALID, "uniforms: primitive_count must equal the uploaded primitive count");
    rc = ensure_rows(ctx, u->width, u->height);
    if (rc) return rc;
    FrameParams fp{};
    fp.width = u->width; fp.height = u->height; fp.spp = u->samples_per_pixel; fp.max_depth = u->max_depth;
    fp.seed = u->seed; fp.frame_number = u->frame_number;
    fp.base_samples = u->samples_per_pixel / 4u < 4u ? 4u : u->samples_per_pixel / 4u;     // shader.wgsl:513
    fp.n_local_pixels = (uint32_t)(ctx->rows.size() * (size_t)u->width);
    fp.all_ones = 0xffffffffu;
    fp.char_aspect = u->char_aspect_ratio;
    fp.half_h = tanf(u->fov_rad / 2.0f);                 // shader.wgsl:522, frame-uniform: evaluated once (glibc)
    fp.half_w = u->aspect_ratio * fp.half_h;             // shader.wgsl:523
    fp.wm1 = (float)(u->width - 1u);
    fp.hm1 = (float)(u->height - 1u);
    fp.cam_kx = 2.0f * fp.half_w / fp.wm1;
    fp.cam_ky = 2.0f * fp.half_h / (fp.hm1 * fp.char_aspect);
    fp.cam_cy = fp.half_h / fp.char_aspect;
    fp.spp_f = (float)u->samples_per_pixel;
    fp.base_f = (float)fp.base_samples;
    fp.cam_pos = make_float3(u->camera_pos.x, u->camera_pos.y, u->camera_pos.z);
    fp.cam_fwd = make_float3(u->camera_forward.x, u->camera_forward.y, u->camera_forward.z);
    fp.cam_right = make_float3(u->camera_right.x, u->camera_right.y, u->camera_right.z);
    fp.cam_up = make_float3(u->camera_up.x, u->camera_up.y, u->camera_up.z);
    fp.row_map = ctx->cfg.shard_count > 1u ? ctx->d_row_map : nullptr;
    fp.pixels = ctx->d_pixels; fp.accumulation = ctx->d_accum; fp.variance = ctx->d_variance;
    fp.counters = ctx->d_counters; fp.work_counter = ctx->d_work; fp.stats = ctx->d_stats;
    if (n_batch) {
        if (ctx->cfg.count_work) return fail(ctx, TRT_ERR_INVALID, "frame batch on an instrumented (count_work) context");
        fp.batch = ctx->d_batch; fp.batch_cur = ctx->d_batch_cur; fp.n_batch = n_batch;
    } else {
        CK(cudaMemsetAsync(ctx->d_stats, 0, TRT_STATS_BYTES, ctx->stream));
        ctx->launches = 0;
        CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    }
    // longest pixels first, predicted by the variance buffer of the previous frame of the same size
    static const bool no_order = [] { const char *k = getenv("TRT_B200_NO_ORDER"); return k && *k == '1'; }();
    // (not for frames far smaller than the machine: with fewer pixels than two lanes per SM-resident warp slot every pixel
    // starts at once whatever the order, and the two sorting launches are 5 % of a 0.3 ms frame -- C5: p50 0.298 -> 0.281 ms;
    // C1's 80 000 pixels still gain 1.7 % from warps of like pixels)
    const bool worth_ordering = fp.n_local_pixels > 2u * (uint32_t)ctx->sm_count * 128u;
    if (!no_order && worth_ordering && !n_batch && !ctx->cfg.count_work && fp.n_local_pixels && ctx->order_w == u->width && ctx->order_h == u->height) {
        CK(cudaMemsetAsync(ctx->d_work + 16, 0, 2 * TRT_ORDER_BUCKETS * sizeof(unsigned int), ctx->stream));
        const uint32_t ob = (fp.n_local_pixels + 255u) / 256u;
        const uint32_t *cost = ctx->cost_valid ? ctx->d_cost : nullptr;
        const float max_cost = (float)fp.spp * (float)(fp.max_depth ? fp.max_depth : 1u);
        order_count_kernel<<<ob, 256, 0, ctx->stream>>>(ctx->d_variance, fp.row_map, fp.width, fp.n_local_pixels, fp.base_samples,
                                                       fp.spp, ctx->d_work + 16, cost, max_cost);
        order_scatter_kernel<<<ob, 256, 0, ctx->stream>>>(ctx->d_variance, fp.row_map, fp.width, fp.n_local_pixels, fp.base_samples,
                                                         fp.spp, ctx->d_work + 16, ctx->d_order, cost, max_cost);
        ctx->launches += 1;
        ctx->launches += 1;
        fp.order = ctx->d_order;
    }
    if (!n_batch) {
        fp.ctas_hint = frame_ctas_hint(ctx, fp);
        static const bool say = [] { const char *k = getenv("TRT_B200_VERBOSE"); return k && *k == '1'; }();
        if (say) fprintf(stderr, "[trt_b200] frame %ux%u, %u local pixels: resident CTAs hint %u (bounces of the last frame %llu)\n", fp.width,
                         fp.height, fp.n_local_pixels, fp.ctas_hint, (unsigned long long)ctx->stats_bounces);
    }
    if (fp.n_local_pixels) {
        std::lock_guard<std::mutex> lock(g_device_state_mutex);
        if (ctx->cfg.count_work) {
            const uint32_t blocks = (fp.n_local_pixels + 127u) / 128u;
            // the work counters are those of the as-written shader: conv 2 contexts count with conv 0 arithmetic
            if (ctx->cfg.conv == 1u) trace_pixel_kernel<1, true><<<blocks, 128, 0, ctx->stream>>>(ctx->sv, ctx->lv, fp);
            else trace_pixel_kernel<0, true><<<blocks, 128, 0, ctx->stream>>>(ctx->sv, ctx->lv, fp);
            ctx->kernel_name = "trace_pixel_kernel<CONV, count>";
            ctx->launches += 1;
        } else if (ctx->cfg.conv == 2u) {
            if (!n_batch) fp.cost = ctx->d_cost;
            cudaError_t e = native_trace_launch(ctx->device, ctx, ctx->scene_serial, ctx->sv, ctx->lv, &ctx->small, fp, ctx->sm_count,
                                                ctx->stream, &ctx->launches, &ctx->kernel_name);
            if (e != cudaSuccess) { ctx->err = std::string("native trace launch: ") + cudaGetErrorString(e); return TRT_ERR_CUDA; }
            ctx->cost_kernel = strstr(ctx->kernel_name, "native") || strstr(ctx->kernel_name, "jit") || strstr(ctx->kernel_name, "pool");   // the wave fallback records no cost
        } else {
            cudaError_t e = persistent_bind_scene(ctx->device, ctx, ctx->scene_serial, ctx->sv, ctx->lv, ctx->stream);
            // TRT_B200_KERNEL=v1 selects the previous production kernel (A/B measurements only)
            static const bool use_v1 = [] { const char *k = getenv("TRT_B200_KERNEL"); return k && !strcmp(k, "v1"); }();
            // Strict kernels: the deferred-NEE machine where its scan has a compile-time run list (the default Cornell
            // layout: 2 289 against 1 813 Msamples/s), else the round-1 wave kernel (faster than the deferred machine with
            // the run-time run loop: 1 813 against 1 664).  TRT_B200_STRICT=wave / deferred / generic force one of them
            // (deferred: the machine on every scene, the configuration the bit-exactness tests of the machine use).
            static const char *strict_env = getenv("TRT_B200_STRICT");
            const bool use_wave = strict_env && !strcmp(strict_env, "wave");
            const bool force_deferred = strict_env && (!strcmp(strict_env, "deferred") || !strcmp(strict_env, "generic"));
            const bool deferred = !use_v1 && !use_wave && ctx->sv.n_lights <= TRT_NV_MAX_LIGHTS && (ctx->cornell_runs || force_deferred);
            if (e == cudaSuccess) {
                if (deferred) {
                    if (!n_batch) fp.cost = ctx->d_cost;
                    static const bool generic = [] { const char *k = getenv("TRT_B200_STRICT"); return k && !strcmp(k, "generic"); }();
                    e = strict_native_launch(ctx->device, ctx->sv, ctx->lv, fp, (int)ctx->cfg.conv, ctx->sm_count, scene_fits_constant(ctx->sv),
                                             ctx->cornell_runs && !generic, ctx->stream, &ctx->launches, &ctx->kernel_name);
                } else if (use_v1 && !n_batch) e = persistent_trace_launch(ctx->sv, ctx->lv, fp, (int)ctx->cfg.conv, ctx->sm_count, ctx->stream, &ctx->launches);
                else e = wave_trace_launch(ctx->sv, ctx->lv, fp, (int)ctx->cfg.conv, ctx->sm_count, scene_fits_constant(ctx->sv),
                                           ctx->stream, &ctx->launches);
            }
            if (e != cudaSuccess) { ctx->err = std::string("trace launch: ") + cudaGetErrorString(e); return TRT_ERR_CUDA; }
            ctx->cost_kernel = deferred;
            if (!deferred)
                ctx->kernel_name = use_v1 && !n_batch ? "trace_persistent_kernel" : scene_fits_constant(ctx->sv) ? "trace_wave_kernel<CONV, SrcConst>"
                                                                                                              : "trace_wave_kernel<CONV, SrcGlobal>";
        }
        CK(cudaGetLastError());
    }
    if (!n_batch) {
        CK(cudaEventRecord(ctx->ev[1], ctx->stream));
        if (!ctx->ev_stats) CK(cudaEventCreateWithFlags(&ctx->ev_stats, cudaEventDisableTiming));
        CK(cudaMemcpyAsync(ctx->h_len + 4, ctx->d_stats, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
        CK(cudaEventRecord(ctx->ev_stats, ctx->stream));
        if (!(ctx->stats_valid && ctx->stats_w == u->width && ctx->stats_h == u->height && ctx->stats_pixels == fp.n_local_pixels &&
              ctx->stats_spp == fp.spp && ctx->stats_depth == fp.max_depth))
            ctx->stats_bounces = 0;                       // a different frame description: no history yet
        ctx->stats_valid = true;
        ctx->stats_w = u->width; ctx->stats_h = u->height; ctx->stats_pixels = fp.n_local_pixels;
        ctx->stats_spp = fp.spp; ctx->stats_depth = fp.max_depth;
    }
    ctx->last_u = *u;
    ctx->have_frame = true;
    ctx->order_w = u->width;
    ctx->order_h = u->height;
    ctx->cost_valid = fp.cost != nullptr && fp.n_local_pixels != 0u && ctx->cost_kernel;
    return TRT_OK;
}

extern "C" int trt_trace_async(trt_ctx *ctx, const trt_uniforms *u) {
    if (!ctx) return TRT_ERR_INVALID;
    return trace_impl(ctx, u, 0u);
}

// d_out / out_cap / h_len default to the context's stream buffer and pinned length word
static int launch_quantise(trt_ctx *ctx, const float4 *src, const uint64_t *row_src, uint32_t w, uint32_t rows, int mode,
                           uint8_t *d_out = nullptr, size_t out_cap = 0, uint64_t *h_len = nullptr) {
    if (mode != TRT_MODE_ASCII && mode != TRT_MODE_TRUECOLOR) return fail(ctx, TRT_ERR_INVALID, "quantise: bad mode");
    const uint64_t n = (uint64_t)w * rows;
    if (n > ctx->max_pixels) return fail(ctx, TRT_ERR_INVALID, "quantise: frame exceeds the context's capacity");
    QuantParams q{};
    q.src = src; q.row_src = row_src; q.width = w; q.height = rows; q.mode = mode;
    q.packed = ctx->d_packed; q.block_len = ctx->d_block_len; q.block_off = ctx->d_block_off;
    q.n_blocks = (uint32_t)((n + TRT_Q_BLOCK - 1) / TRT_Q_BLOCK);
    q.out = d_out ? d_out : ctx->d_stream; q.out_cap = d_out ? out_cap : ctx->stream_cap;
    // one pass (decoupled look-back) unless TRT_B200_QUANT=3pass asks for the round-1 three-launch form (A/B)
    static const bool three_pass = [] { const char *k = getenv("TRT_B200_QUANT"); return k && !strcmp(k, "3pass"); }();
    ctx->fused_rows = false;
    if (q.n_blocks && !three_pass) {
        q.n_tiles = (uint32_t)((n + TRT_QF_TILE - 1) / TRT_QF_TILE);
        q.ticket = (unsigned int *)ctx->d_tile_state;              // word 0: the ticket; states from word 1
        q.tile_state = ctx->d_tile_state + 1;
        q.row_start = d_out ? nullptr : ctx->d_row_start;
        q.total = (uint64_t *)(ctx->d_block_off + q.n_blocks);
        CK(cudaMemsetAsync(ctx->d_tile_state, 0, ((size_t)q.n_tiles + 1) * sizeof(unsigned long long), ctx->stream));
        quant_fused_kernel<<<q.n_tiles, TRT_Q_BLOCK, 0, ctx->stream>>>(q);
        ctx->launches += 1;
        CK(cudaGetLastError());
        ctx->fused_rows = d_out == nullptr;
    } else if (q.n_blocks) {
        quant_pack_kernel<<<q.n_blocks, TRT_Q_BLOCK, 0, ctx->stream>>>(q);
        quant_scan_kernel<<<1, 1024, 0, ctx->stream>>>(q);
        quant_emit_kernel<<<q.n_blocks, TRT_Q_BLOCK, 0, ctx->stream>>>(q);
        ctx->launches += 3;
        CK(cudaGetLastError());
    } else {
        CK(cudaMemsetAsync(ctx->d_block_off, 0, sizeof(uint64_t), ctx->stream));
    }
    ctx->last_q_blocks = q.n_blocks;
    ctx->last_mode = mode;
    ctx->have_row_len = false;
    CK(cudaMemcpyAsync(h_len ? h_len : ctx->h_len, ctx->d_block_off + q.n_blocks, sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    ctx->have_stream = d_out == nullptr;
    return TRT_OK;
}

extern "C" int trt_quantise_async(trt_ctx *ctx, int mode, size_t *len_needed) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!ctx->have_frame) return fail(ctx, TRT_ERR_INVALID, "quantise before any frame was traced");
    CK(cudaSetDevice(ctx->device));
    const bool sharded = ctx->cfg.shard_count > 1u;
    int rc = launch_quantise(ctx, ctx->d_pixels, sharded ? ctx->d_row_src : nullptr, ctx->last_u.width,
                             (uint32_t)ctx->rows.size(), mode);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    if (len_needed) {
        CK(cudaStreamSynchronize(ctx->stream));
        *len_needed = (size_t)*ctx->h_len;
    }
    return TRT_OK;
}

extern "C" int trt_quantise_gathered(trt_ctx *ctx, const void *d_gathered, size_t block_stride_bytes, uint32_t width,
                                     uint32_t height, uint32_t shard_count, uint32_t stripe_rows, int mode,
                                     size_t *len_needed) {
    if (!ctx || !d_gathered || !width || !height || !shard_count || !stripe_rows) return fail(ctx, TRT_ERR_INVALID, "quantise_gathered: bad argument");
    if (height > ctx->cfg.max_height || (size_t)width * height > ctx->max_pixels || block_stride_bytes % sizeof(float4))
        return fail(ctx, TRT_ERR_INVALID, "quantise_gathered: frame exceeds the context's capacity");
    CK(cudaSetDevice(ctx->device));
    // global row y lives in block (stripe % shard_count) at local row index = (rows of that shard below y).  The table
    // lives in its own buffer (this shard's d_row_src stays valid) and is uploaded once per layout.
    if (ctx->g_w != width || ctx->g_h != height || ctx->g_shards != shard_count || ctx->g_stripe != stripe_rows ||
        ctx->g_stride != block_stride_bytes) {
        std::vector<uint64_t> src(height);
        std::vector<uint32_t> local(shard_count, 0);
        for (uint32_t y = 0; y < height; y++) {
            const uint32_t owner = (y / stripe_rows) % shard_count;
            src[y] = (uint64_t)owner * (block_stride_bytes / sizeof(float4)) + (uint64_t)local[owner] * width;
            local[owner]++;
        }
        CK(cudaStreamSynchronize(ctx->stream));
        CK(cudaMemcpy(ctx->d_row_src_g, src.data(), height * sizeof(uint64_t), cudaMemcpyHostToDevice));
        ctx->g_w = width; ctx->g_h = height; ctx->g_shards = shard_count; ctx->g_stripe = stripe_rows; ctx->g_stride = block_stride_bytes;
    }
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    int rc = launch_quantise(ctx, (const float4 *)d_gathered, ctx->d_row_src_g, width, height, mode);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    if (len_needed) {
        CK(cudaStreamSynchronize(ctx->stream));
        *len_needed = (size_t)*ctx->h_len;
    }
    return TRT_OK;
}

// ---- row-addressed streams ---------------------------------------------------------------------------
// A frame's byte stream is its rows back to back (lib.rs:503-532), and a row's bytes depend on that row's pixels
// only.  So a sharded frame needs no pixel exchange at all: every shard quantises ITS rows on its GPU, publishes the
// byte length of each (4 B per row), and once the lengths of all rows are known -- one tiny all-gather between
// processes, or a host loop inside one -- every shard copies its rows device -> host over its own PCIe link, straight
// to their final place in one host buffer.
extern "C" int trt_quantise_rows_async(trt_ctx *ctx, int mode) {
    if (!ctx) return TRT_ERR_INVALID;
    int rc = trt_quantise_async(ctx, mode, nullptr);
    if (rc) return rc;
    const uint32_t rows = (uint32_t)ctx->rows.size();
    if (rows) {
        if (ctx->fused_rows) quant_row_len_from_starts_kernel<<<(rows + 255u) / 256u, 256, 0, ctx->stream>>>(ctx->d_row_start, rows, ctx->d_row_len);
        else quant_row_len_kernel<<<(rows + 7u) / 8u, 256, 0, ctx->stream>>>(ctx->d_packed, ctx->last_u.width, rows, mode, ctx->d_row_len);
        ctx->launches += 1;
        CK(cudaGetLastError());
        CK(cudaMemcpyAsync(ctx->h_row_len, ctx->d_row_len, rows * sizeof(uint32_t), cudaMemcpyDeviceToHost, ctx->stream));
    }
    CK(cudaEventRecord(ctx->ev_row_len, ctx->stream));
    CK(cudaEventRecord(ctx->ev[2], ctx->stream));
    ctx->row_len_rows = rows;
    ctx->have_row_len = true;
    return TRT_OK;
}

extern "C" int trt_row_lengths(trt_ctx *ctx, uint32_t *out, uint32_t cap, uint32_t *n_rows) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!ctx->have_row_len) return fail(ctx, TRT_ERR_INVALID, "row_lengths before trt_quantise_rows_async");
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventSynchronize(ctx->ev_row_len));
    if (n_rows) *n_rows = ctx->row_len_rows;
    if (ctx->row_len_rows > cap || (!out && ctx->row_len_rows)) return fail(ctx, TRT_ERR_TOO_SMALL, "row_lengths: destination too small");
    if (ctx->row_len_rows) memcpy(out, ctx->h_row_len, ctx->row_len_rows * sizeof(uint32_t));
    return TRT_OK;
}

// Issues the copies of this shard's rows into host_base + row_off[y] (row_off: height + 1 offsets of the whole frame's
// rows; runs of consecutive rows -- stripes -- are one copy each).  wait != 0: returns when the bytes have landed.
extern "C" int trt_read_stream_rows(trt_ctx *ctx, uint8_t *host_base, size_t cap, const uint64_t *row_off, uint32_t height, int wait) {
    if (!ctx || !row_off) return TRT_ERR_INVALID;
    if (!ctx->have_row_len || !ctx->have_stream) return fail(ctx, TRT_ERR_INVALID, "read_stream_rows before trt_quantise_rows_async");
    if (height != ctx->last_u.height) return fail(ctx, TRT_ERR_INVALID, "read_stream_rows: height differs from the last frame");
    if (row_off[height] > cap || (!host_base && row_off[height])) return fail(ctx, TRT_ERR_TOO_SMALL, "read_stream_rows: output buffer too small for the frame stream");
    CK(cudaSetDevice(ctx->device));
    CK(cudaEventSynchronize(ctx->ev_row_len));
    uint64_t local = 0;
    const size_t n = ctx->rows.size();
    for (size_t i = 0; i < n;) {
        size_t j = i;
        uint64_t bytes = 0;
        while (j < n && ctx->rows[j] == ctx->rows[i] + (uint32_t)(j - i)) { bytes += ctx->h_row_len[j]; j++; }
        const uint32_t y0 = ctx->rows[i], y1 = ctx->rows[j - 1] + 1u;
        if (row_off[y1] - row_off[y0] != bytes) return fail(ctx, TRT_ERR_INVALID, "read_stream_rows: row offsets disagree with this shard's row lengths");
        if (bytes) CK(cudaMemcpyAsync(host_base + row_off[y0], ctx->d_stream + local, bytes, cudaMemcpyDeviceToHost, ctx->stream));
        local += bytes;
        i = j;
    }
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    if (wait) CK(cudaStreamSynchronize(ctx->stream));
    return TRT_OK;
}

extern "C" int trt_host_register(void *p, size_t bytes) {
    if (!p || !bytes) return TRT_ERR_INVALID;
    cudaError_t e = cudaHostRegister(p, bytes, cudaHostRegisterPortable);
    if (e != cudaSuccess) { g_create_err = std::string("trt_host_register: ") + cudaGetErrorString(e); cudaGetLastError(); return TRT_ERR_CUDA; }
    return TRT_OK;
}
extern "C" int trt_host_unregister(void *p) {
    if (!p) return TRT_ERR_INVALID;
    cudaError_t e = cudaHostUnregister(p);
    if (e != cudaSuccess) { cudaGetLastError(); return TRT_ERR_CUDA; }
    return TRT_OK;
}

// ---- pipelined frames -----------------------------------------------------------------------------------
// The reference's loop is strictly serial: dispatch, map + poll(Wait), format, print (lib.rs:442-477, 503-545).  Here
// up to two frames are in flight: trt_frame_submit queues trace + quantise of frame f (into one of two device stream
// buffers) and returns; trt_frame_collect hands out the OLDEST submitted frame -- its device -> host copy runs on a
// second stream behind an event, so the copy, and whatever the caller does with the bytes (stdout), overlap the trace
// of the frame submitted after it.  The frames, their order and their bytes are those of the serial loop.
extern "C" int trt_frame_submit(trt_ctx *ctx, const trt_uniforms *u, int mode) {
    if (!ctx) return TRT_ERR_INVALID;
    if (mode != TRT_MODE_ASCII && mode != TRT_MODE_TRUECOLOR) return fail(ctx, TRT_ERR_INVALID, "frame_submit: bad mode");
    if (ctx->cfg.shard_count > 1u) return fail(ctx, TRT_ERR_INVALID, "frame_submit: not on a sharded context");
    if (ctx->pipe_submitted - ctx->pipe_collected >= 2u) return fail(ctx, TRT_ERR_INVALID, "frame_submit: two frames already in flight (collect one)");
    CK(cudaSetDevice(ctx->device));
    if (!ctx->copy_stream) {
        CK(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (auto &e : ctx->ev_q) CK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
        CK(cudaMalloc(&ctx->d_stream2, ctx->stream_cap));
    }
    const uint32_t slot = ctx->pipe_submitted & 1u;
    // the slot's previous copy (frame f - 2) was waited for by its collect, so its buffer is free
    int rc = trace_impl(ctx, u, 0u);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    rc = launch_quantise(ctx, ctx->d_pixels, nullptr, u->width, u->height, mode, slot ? ctx->d_stream2 : ctx->d_stream, ctx->stream_cap,
                         ctx->h_len + 2 + slot);
    if (rc) return rc;
    ctx->have_stream = false;          // the plain read_stream path does not see pipelined frames
    CK(cudaEventRecord(ctx->ev_q[slot], ctx->stream));
    ctx->pipe_mode[slot] = mode;
    ctx->pipe_submitted++;
    return TRT_OK;
}

extern "C" int trt_frame_collect(trt_ctx *ctx, uint8_t *out, size_t cap, size_t *len) {
    if (!ctx || !len) return TRT_ERR_INVALID;
    if (ctx->pipe_submitted == ctx->pipe_collected) return fail(ctx, TRT_ERR_INVALID, "frame_collect: no frame in flight");
    CK(cudaSetDevice(ctx->device));
    const uint32_t slot = ctx->pipe_collected & 1u;
    CK(cudaEventSynchronize(ctx->ev_q[slot]));
    const size_t n = (size_t)ctx->h_len[2 + slot];
    *len = n;
    ctx->pipe_collected++;
    if (n > ctx->stream_cap) return fail(ctx, TRT_ERR_INVALID, "internal: stream longer than its buffer");
    if (n > cap || (!out && n)) return fail(ctx, TRT_ERR_TOO_SMALL, "output buffer too small for the frame stream");
    if (n) CK(cudaMemcpyAsync(out, slot ? ctx->d_stream2 : ctx->d_stream, n, cudaMemcpyDeviceToHost, ctx->copy_stream));
    CK(cudaStreamSynchronize(ctx->copy_stream));
    return TRT_OK;
}

extern "C" int trt_sync(trt_ctx *ctx) {
    if (!ctx) return TRT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    return TRT_OK;
}

extern "C" int trt_read_stream(trt_ctx *ctx, uint8_t *out, size_t cap, size_t *len) {
    if (!ctx || !len) return TRT_ERR_INVALID;
    if (!ctx->have_stream) return fail(ctx, TRT_ERR_INVALID, "read_stream before quantise");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    const size_t n = (size_t)*ctx->h_len;
    *len = n;
    if (n > ctx->stream_cap) return fail(ctx, TRT_ERR_INVALID, "internal: stream longer than its buffer");
    if (n > cap || (!out && n)) return fail(ctx, TRT_ERR_TOO_SMALL, "output buffer too small for the frame stream");
    if (n) CK(cudaMemcpyAsync(out, ctx->d_stream, n, cudaMemcpyDeviceToHost, ctx->stream));
    CK(cudaEventRecord(ctx->ev[3], ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    return TRT_OK;
}

extern "C" int trt_render_frame(trt_ctx *ctx, const trt_uniforms *u, int mode, uint8_t *out, size_t cap, size_t *len) {
    if (!ctx) return TRT_ERR_INVALID;
    if (mode != TRT_MODE_ASCII && mode != TRT_MODE_TRUECOLOR && mode != TRT_MODE_NONE)
        return fail(ctx, TRT_ERR_INVALID, "render_frame: bad mode");
    int rc = trt_trace_async(ctx, u);
    if (rc) return rc;
    if (mode == TRT_MODE_NONE) {
        if (len) *len = 0;
        CK(cudaEventRecord(ctx->ev[2], ctx->stream));
        CK(cudaEventRecord(ctx->ev[3], ctx->stream));
        return trt_sync(ctx);
    }
    rc = trt_quantise_async(ctx, mode, nullptr);
    if (rc) return rc;
    size_t n = 0;
    rc = trt_read_stream(ctx, out, cap, &n);
    if (len) *len = n;
    return rc;
}

// ---- frame batch ---------------------------------------------------------------------------------------
// The reference's loop (lib.rs:414-560) renders the frames of a still camera one at a time, each a full GPU
// round trip; at terminal sizes a frame is far too small to fill a B200 (80 000 serial pixel chains against
// 151 552 resident lanes).  A batch traces n such frames -- same scene, size and camera, their own
// (seed, frame_number) -- as ONE launch over n x pixels work items, then blends them in frame order.
extern "C" int trt_render_frames(trt_ctx *ctx, const trt_uniforms *u, uint32_t n, int mode, int emit_all, uint8_t *out,
                                 size_t cap, size_t *len, size_t *frame_len) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!u || n == 0u) return fail(ctx, TRT_ERR_INVALID, "render_frames: no frames");
    if (mode != TRT_MODE_ASCII && mode != TRT_MODE_TRUECOLOR && mode != TRT_MODE_NONE)
        return fail(ctx, TRT_ERR_INVALID, "render_frames: bad mode");
    if (ctx->cfg.shard_count > 1u && mode != TRT_MODE_NONE)
        return fail(ctx, TRT_ERR_INVALID, "render_frames: a sharded context has no whole frame to emit (use TRT_MODE_NONE and gather)");
    for (uint32_t i = 1; i < n; i++) {
        trt_uniforms a = u[0], b = u[i];
        a.seed = b.seed = 0u; a.frame_number = b.frame_number = 0u;
        if (memcmp(&a, &b, sizeof(trt_uniforms)) != 0)
            return fail(ctx, TRT_ERR_INVALID, "render_frames: the frames of a batch may differ only in seed and frame_number");
    }
    int rc = check_uniforms(ctx, &u[0]);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    rc = ensure_rows(ctx, u[0].width, u[0].height);
    if (rc) return rc;
    const size_t np = ctx->rows.size() * (size_t)u[0].width;
    if (len) *len = 0;
    if (np == 0) return TRT_OK;
    const bool emit = emit_all && mode != TRT_MODE_NONE;
    const size_t stride = (np * 26 + (size_t)u[0].height * 2 + 64 + 255) & ~(size_t)255;
    // frames per launch: staging (16 B/pixel/frame, plus the stream when every frame is emitted) within 6 GB,
    // and the work-item index within 31 bits
    const size_t per_frame = np * sizeof(float4) + (emit ? stride : 0);
    size_t chunk = (size_t(6) << 30) / per_frame;
    if (chunk > ((size_t(1) << 31) - 1) / np) chunk = ((size_t(1) << 31) - 1) / np;
    if (chunk < 1) chunk = 1;
    if (chunk > n) chunk = n;
    if (ctx->batch_cur_cap < chunk * np) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->d_batch_cur) cudaFree(ctx->d_batch_cur);
        ctx->d_batch_cur = nullptr; ctx->batch_cur_cap = 0;
        CK(cudaMalloc(&ctx->d_batch_cur, chunk * np * sizeof(float4)));
        ctx->batch_cur_cap = chunk * np;
    }
    if (ctx->batch_cap < chunk) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->d_batch) cudaFree(ctx->d_batch);
        ctx->d_batch = nullptr; ctx->batch_cap = 0;
        CK(cudaMalloc(&ctx->d_batch, chunk * sizeof(uint2)));
        ctx->batch_cap = (uint32_t)chunk;
    }
    if (emit) {
        if (ctx->batch_stream_cap < chunk * stride) {
            CK(cudaStreamSynchronize(ctx->stream));
            if (ctx->d_batch_stream) cudaFree(ctx->d_batch_stream);
            ctx->d_batch_stream = nullptr; ctx->batch_stream_cap = 0;
            CK(cudaMalloc(&ctx->d_batch_stream, chunk * stride));
            ctx->batch_stream_cap = chunk * stride;
        }
        if (ctx->batch_len_cap < chunk) {
            if (ctx->h_batch_len) cudaFreeHost(ctx->h_batch_len);
            ctx->h_batch_len = nullptr; ctx->batch_len_cap = 0;
            CK(cudaMallocHost(&ctx->h_batch_len, chunk * sizeof(uint64_t)));
            ctx->batch_len_cap = chunk;
        }
    }
    CK(cudaMemsetAsync(ctx->d_stats, 0, TRT_STATS_BYTES, ctx->stream));
    ctx->launches = 0;
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    size_t off = 0;
    std::vector<uint2> hb(chunk);
    for (uint32_t c0 = 0; c0 < n; c0 += (uint32_t)chunk) {
        const uint32_t m = (uint32_t)std::min<size_t>(chunk, n - c0);
        for (uint32_t f = 0; f < m; f++) hb[f] = make_uint2(u[c0 + f].seed, u[c0 + f].frame_number);
        CK(cudaMemcpyAsync(ctx->d_batch, hb.data(), m * sizeof(uint2), cudaMemcpyHostToDevice, ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));          // hb is reused by the next chunk
        rc = trace_impl(ctx, &u[c0], m);
        if (rc) return rc;
        const uint32_t *row_map = ctx->cfg.shard_count > 1u ? ctx->d_row_map : nullptr;
        if (ctx->cfg.conv == 2u) {
            cudaError_t e = native_blend_launch(ctx->d_batch_cur, ctx->d_batch, m, (uint32_t)np, row_map, u[0].width, ctx->d_accum,
                                                ctx->d_pixels, ctx->stream);
            if (e != cudaSuccess) { ctx->err = std::string("blend launch: ") + cudaGetErrorString(e); return TRT_ERR_CUDA; }
        } else {
            blend_frames_kernel<0><<<((uint32_t)np + 255u) / 256u, 256, 0, ctx->stream>>>(ctx->d_batch_cur, ctx->d_batch, m, (uint32_t)np,
                                                                                         row_map, u[0].width, ctx->d_accum, ctx->d_pixels);
            CK(cudaGetLastError());
        }
        ctx->launches += 1;
        if (c0 + m == n) CK(cudaEventRecord(ctx->ev[1], ctx->stream));
        if (emit) {
            for (uint32_t f = 0; f < m; f++) {
                rc = launch_quantise(ctx, ctx->d_batch_cur + (size_t)f * np, nullptr, u[0].width, u[0].height, mode,
                                     ctx->d_batch_stream + (size_t)f * stride, stride, ctx->h_batch_len + f);
                if (rc) return rc;
            }
            if (c0 + m == n) CK(cudaEventRecord(ctx->ev[2], ctx->stream));
            CK(cudaStreamSynchronize(ctx->stream));
            for (uint32_t f = 0; f < m; f++) {
                const size_t nf = (size_t)ctx->h_batch_len[f];
                if (frame_len) frame_len[c0 + f] = nf;
                if (nf > stride) return fail(ctx, TRT_ERR_INVALID, "internal: stream longer than its buffer");
                if (off + nf > cap || (!out && nf)) return fail(ctx, TRT_ERR_TOO_SMALL, "output buffer too small for the batch's streams");
                if (nf) CK(cudaMemcpyAsync(out + off, ctx->d_batch_stream + (size_t)f * stride, nf, cudaMemcpyDeviceToHost, ctx->stream));
                off += nf;
            }
        }
    }
    ctx->last_u = u[n - 1];
    if (emit) {
        CK(cudaEventRecord(ctx->ev[3], ctx->stream));
        CK(cudaStreamSynchronize(ctx->stream));
        if (len) *len = off;
        return TRT_OK;
    }
    if (mode == TRT_MODE_NONE) {
        CK(cudaEventRecord(ctx->ev[2], ctx->stream));
        CK(cudaEventRecord(ctx->ev[3], ctx->stream));
        return trt_sync(ctx);
    }
    rc = trt_quantise_async(ctx, mode, nullptr);
    if (rc) return rc;
    size_t nlast = 0;
    rc = trt_read_stream(ctx, out, cap, &nlast);
    if (len) *len = nlast;
    if (frame_len) { for (uint32_t i = 0; i + 1 < n; i++) frame_len[i] = 0; frame_len[n - 1] = nlast; }
    return rc;
}

// ---- frames over GPUs: the two halves of a frame batch as separate calls -----------------------------------
// trt_trace_frames_async traces n frames of a still camera (one launch, work item = (frame, pixel)) and leaves every
// frame's per-pixel mean in the context's staging buffer (trt_frame_means), WITHOUT the temporal blend;
// trt_blend_frames_async applies the blend (shader.wgsl:610-619) in frame order to n means that lie in this device's
// memory, wherever they were traced.  trt_multi_render_frames deals the frames of a batch over the GPUs (frame f on
// GPU f mod G), gathers the means on the first GPU and blends them there: bit-identical to one GPU, G times the lanes.
static int same_but_seed(trt_ctx *ctx, const trt_uniforms *u, uint32_t n, const char *who) {
    for (uint32_t i = 1; i < n; i++) {
        trt_uniforms a = u[0], b = u[i];
        a.seed = b.seed = 0u; a.frame_number = b.frame_number = 0u;
        if (memcmp(&a, &b, sizeof(trt_uniforms)) != 0) {
            ctx->err = std::string(who) + ": the frames of a batch may differ only in seed and frame_number";
            return TRT_ERR_INVALID;
        }
    }
    return TRT_OK;
}

static int ensure_batch_buffers(trt_ctx *ctx, size_t frames, size_t np, bool means) {
    if (means && ctx->batch_cur_cap < frames * np) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->d_batch_cur) cudaFree(ctx->d_batch_cur);
        ctx->d_batch_cur = nullptr; ctx->batch_cur_cap = 0;
        CK(cudaMalloc(&ctx->d_batch_cur, frames * np * sizeof(float4)));
        ctx->batch_cur_cap = frames * np;
    }
    if (ctx->batch_cap < frames) {
        CK(cudaStreamSynchronize(ctx->stream));
        if (ctx->d_batch) cudaFree(ctx->d_batch);
        ctx->d_batch = nullptr; ctx->batch_cap = 0;
        CK(cudaMalloc(&ctx->d_batch, frames * sizeof(uint2)));
        ctx->batch_cap = (uint32_t)frames;
    }
    return TRT_OK;
}

extern "C" int trt_trace_frames_async(trt_ctx *ctx, const trt_uniforms *u, uint32_t n) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!u || n == 0u) return fail(ctx, TRT_ERR_INVALID, "trace_frames: no frames");
    int rc = same_but_seed(ctx, u, n, "trace_frames");
    if (rc) return rc;
    rc = check_uniforms(ctx, &u[0]);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    rc = ensure_rows(ctx, u[0].width, u[0].height);
    if (rc) return rc;
    const size_t np = ctx->rows.size() * (size_t)u[0].width;
    if (np == 0) return TRT_OK;
    if ((size_t)n * np * sizeof(float4) > (size_t(6) << 30) || (size_t)n * np > ((size_t(1) << 31) - 1))
        return fail(ctx, TRT_ERR_INVALID, "trace_frames: the batch does not fit one launch (use trt_render_frames, which chunks)");
    rc = ensure_batch_buffers(ctx, n, np, true);
    if (rc) return rc;
    std::vector<uint2> hb(n);
    for (uint32_t f = 0; f < n; f++) hb[f] = make_uint2(u[f].seed, u[f].frame_number);
    CK(cudaMemcpyAsync(ctx->d_batch, hb.data(), n * sizeof(uint2), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));              // hb leaves scope
    CK(cudaMemsetAsync(ctx->d_stats, 0, TRT_STATS_BYTES, ctx->stream));
    ctx->launches = 0;
    CK(cudaEventRecord(ctx->ev[0], ctx->stream));
    rc = trace_impl(ctx, u, n);
    if (rc) return rc;
    CK(cudaEventRecord(ctx->ev[1], ctx->stream));
    return TRT_OK;
}

extern "C" int trt_frame_means(trt_ctx *ctx, void **dptr, size_t *frame_stride_bytes) {
    if (!ctx || !dptr) return TRT_ERR_INVALID;
    if (!ctx->d_batch_cur || !ctx->have_frame) return fail(ctx, TRT_ERR_INVALID, "frame_means before trt_trace_frames_async");
    *dptr = ctx->d_batch_cur;
    if (frame_stride_bytes) *frame_stride_bytes = ctx->rows.size() * (size_t)ctx->last_u.width * sizeof(float4);
    return TRT_OK;
}

extern "C" int trt_blend_frames_async(trt_ctx *ctx, const void *d_means, const trt_uniforms *u, uint32_t n) {
    if (!ctx) return TRT_ERR_INVALID;
    if (!d_means || !u || n == 0u) return fail(ctx, TRT_ERR_INVALID, "blend_frames: no frames");
    int rc = same_but_seed(ctx, u, n, "blend_frames");
    if (rc) return rc;
    rc = check_uniforms(ctx, &u[0]);
    if (rc) return rc;
    CK(cudaSetDevice(ctx->device));
    rc = ensure_rows(ctx, u[0].width, u[0].height);
    if (rc) return rc;
    const size_t np = ctx->rows.size() * (size_t)u[0].width;
    if (np == 0) return TRT_OK;
    rc = ensure_batch_buffers(ctx, n, np, false);
    if (rc) return rc;
    std::vector<uint2> hb(n);
    for (uint32_t f = 0; f < n; f++) hb[f] = make_uint2(u[f].seed, u[f].frame_number);
    CK(cudaMemcpyAsync(ctx->d_batch, hb.data(), n * sizeof(uint2), cudaMemcpyHostToDevice, ctx->stream));
    CK(cudaStreamSynchronize(ctx->stream));
    const uint32_t *row_map = ctx->cfg.shard_count > 1u ? ctx->d_row_map : nullptr;
    if (ctx->cfg.conv == 2u) {
        cudaError_t e = native_blend_launch((float4 *)d_means, ctx->d_batch, n, (uint32_t)np, row_map, u[0].width, ctx->d_accum, ctx->d_pixels,
                                            ctx->stream);
        if (e != cudaSuccess) { ctx->err = std::string("blend launch: ") + cudaGetErrorString(e); return TRT_ERR_CUDA; }
    } else if (ctx->cfg.conv == 1u) {
        blend_frames_kernel<1><<<((uint32_t)np + 255u) / 256u, 256, 0, ctx->stream>>>((float4 *)d_means, ctx->d_batch, n, (uint32_t)np, row_map,
                                                                                     u[0].width, ctx->d_accum, ctx->d_pixels);
        CK(cudaGetLastError());
    } else {
        blend_frames_kernel<0><<<((uint32_t)np + 255u) / 256u, 256, 0, ctx->stream>>>((float4 *)d_means, ctx->d_batch, n, (uint32_t)np, row_map,
                                                                                     u[0].width, ctx->d_accum, ctx->d_pixels);
        CK(cudaGetLastError());
    }
    ctx->launches += 1;
    ctx->last_u = u[n - 1];
    ctx->have_frame = true;
    return TRT_OK;
}

extern "C" const char *trt_trace_kernel_name(const trt_ctx *ctx) { return ctx ? ctx->kernel_name : ""; }

extern "C" int trt_get_timings(trt_ctx *ctx, trt_timings *out) {
    if (!ctx || !out) return TRT_ERR_INVALID;
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    trt_timings t{};
    float ms = 0.0f;
    if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[1]) == cudaSuccess) t.trace_ms = ms;
    if (cudaEventElapsedTime(&ms, ctx->ev[1], ctx->ev[2]) == cudaSuccess) t.quantise_ms = ms;
    if (cudaEventElapsedTime(&ms, ctx->ev[2], ctx->ev[3]) == cudaSuccess) t.d2h_ms = ms;
    if (cudaEventElapsedTime(&ms, ctx->ev[0], ctx->ev[3]) == cudaSuccess) t.total_ms = ms;
    cudaGetLastError();
    t.kernel_launches = ctx->launches;
    *out = t;
    return TRT_OK;
}

// ---- parity / checkpoint access ---------------------------------------------------------------------
// Copies the rows this shard owns from a full-frame device buffer into the caller's full-frame array.
static int read_rows(trt_ctx *ctx, const void *d_src, void *dst, size_t elem, size_t n_pixels) {
    if (!ctx->have_frame) return fail(ctx, TRT_ERR_INVALID, "read before any frame was traced");
    const size_t w = ctx->last_u.width, h = ctx->last_u.height;
    if (!dst || n_pixels < w * h) return fail(ctx, TRT_ERR_TOO_SMALL, "read: destination smaller than width*height");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    if (ctx->cfg.shard_count <= 1u) {
        CK(cudaMemcpy(dst, d_src, w * h * elem, cudaMemcpyDeviceToHost));
        return TRT_OK;
    }
    std::vector<uint32_t> rows;
    shard_rows(ctx->cfg, (uint32_t)h, rows);
    for (uint32_t y : rows)
        CK(cudaMemcpy((char *)dst + (size_t)y * w * elem, (const char *)d_src + (size_t)y * w * elem, w * elem,
                      cudaMemcpyDeviceToHost));
    return TRT_OK;
}

extern "C" int trt_read_pixels(trt_ctx *ctx, float *dst4, size_t n_pixels) {
    if (!ctx) return TRT_ERR_INVALID;
    return read_rows(ctx, ctx->d_pixels, dst4, sizeof(float4), n_pixels);
}
extern "C" int trt_read_accumulation(trt_ctx *ctx, float *dst4, size_t n_pixels) {
    if (!ctx) return TRT_ERR_INVALID;
    return read_rows(ctx, ctx->d_accum, dst4, sizeof(float4), n_pixels);
}
extern "C" int trt_read_variance(trt_ctx *ctx, trt_variance *dst, size_t n_pixels) {
    if (!ctx) return TRT_ERR_INVALID;
    return read_rows(ctx, ctx->d_variance, dst, sizeof(trt_variance), n_pixels);
}
extern "C" int trt_write_accumulation(trt_ctx *ctx, const float *src4, size_t n_pixels) {
    if (!ctx || !src4) return TRT_ERR_INVALID;
    if (n_pixels > ctx->max_pixels) return fail(ctx, TRT_ERR_INVALID, "write_accumulation: more pixels than the context holds");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(ctx->d_accum, src4, n_pixels * sizeof(float4), cudaMemcpyHostToDevice));
    return TRT_OK;
}

extern "C" int trt_get_counters(trt_ctx *ctx, trt_counters *out) {
    if (!ctx || !out) return TRT_ERR_INVALID;
    static_assert(sizeof(trt_counters) == K_COUNT * sizeof(unsigned long long), "counter layout");
    if (!ctx->cfg.count_work) return fail(ctx, TRT_ERR_INVALID, "counters need trt_config.count_work = 1");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    CK(cudaMemcpy(out, ctx->d_counters, sizeof(trt_counters), cudaMemcpyDeviceToHost));
    CK(cudaMemset(ctx->d_counters, 0, sizeof(trt_counters)));      // counters are per read interval
    return TRT_OK;
}

// ---- multi-GPU plumbing -------------------------------------------------------------------------------
extern "C" int trt_device_buffer(trt_ctx *ctx, int which, void **dptr, size_t *bytes) {
    if (!ctx || !dptr) return TRT_ERR_INVALID;
    size_t n = 0;
    switch (which) {
        case 0: *dptr = ctx->d_pixels; n = ctx->max_pixels * sizeof(float4); break;
        case 1: *dptr = ctx->d_accum; n = ctx->max_pixels * sizeof(float4); break;
        case 2: *dptr = ctx->d_variance; n = ctx->max_pixels * sizeof(trt_variance); break;
        case 3: *dptr = ctx->d_stream; n = ctx->stream_cap; break;
        case 4: *dptr = ctx->d_stats; n = TRT_STATS_BYTES; break;
        case 5: *dptr = ctx->d_row_len; n = (size_t)ctx->cfg.max_height * sizeof(uint32_t); break;   // trt_quantise_rows_async      // frame statistics + the diagnostic words of a measurement build
        default: return fail(ctx, TRT_ERR_INVALID, "device_buffer: unknown buffer id");
    }
    if (bytes) *bytes = n;
    return TRT_OK;
}

extern "C" int trt_cuda_stream(trt_ctx *ctx, void **stream) {
    if (!ctx || !stream) return TRT_ERR_INVALID;
    *stream = (void *)ctx->stream;
    return TRT_OK;
}

extern "C" int trt_pack_shard_pixels(trt_ctx *ctx, void *d_dst, size_t dst_bytes) {
    if (!ctx || !d_dst) return TRT_ERR_INVALID;
    if (!ctx->have_frame) return fail(ctx, TRT_ERR_INVALID, "pack before any frame was traced");
    CK(cudaSetDevice(ctx->device));
    int rc = ensure_rows(ctx, ctx->last_u.width, ctx->last_u.height);
    if (rc) return rc;
    const uint32_t n = (uint32_t)(ctx->rows.size() * (size_t)ctx->last_u.width);
    if ((size_t)n * sizeof(float4) > dst_bytes) return fail(ctx, TRT_ERR_TOO_SMALL, "pack: destination too small");
    if (n) {
        pack_rows_kernel<<<(n + 255u) / 256u, 256, 0, ctx->stream>>>(ctx->d_pixels, (float4 *)d_dst,
                                                                     ctx->cfg.shard_count > 1u ? ctx->d_row_map : nullptr,
                                                                     ctx->last_u.width, n);
        ctx->launches += 1;
        CK(cudaGetLastError());
    }
    return TRT_OK;
}

extern "C" int trt_alloc_host(size_t bytes, void **out) {
    if (!out) return TRT_ERR_INVALID;
    *out = nullptr;
    cudaError_t e = cudaMallocHost(out, bytes ? bytes : 1);
    if (e != cudaSuccess) { g_create_err = std::string("trt_alloc_host: ") + cudaGetErrorString(e); cudaGetLastError(); return TRT_ERR_CUDA; }
    return TRT_OK;
}
extern "C" void trt_free_host(void *p) { if (p) cudaFreeHost(p); }

extern "C" int trt_get_frame_stats(trt_ctx *ctx, trt_frame_stats *out) {
    if (!ctx || !out) return TRT_ERR_INVALID;
    if (!ctx->have_frame) return fail(ctx, TRT_ERR_INVALID, "frame_stats before any frame was traced");
    CK(cudaSetDevice(ctx->device));
    CK(cudaStreamSynchronize(ctx->stream));
    unsigned long long h[2] = {0, 0};
    CK(cudaMemcpy(h, ctx->d_stats, sizeof h, cudaMemcpyDeviceToHost));
    out->samples_traced = h[0];
    out->bounces = h[1];
    out->pixels = (uint64_t)ctx->rows.size() * ctx->last_u.width;
    out->reserved = 0;
    return TRT_OK;
}

// ---- FP32 peak micro-benchmark -------------------------------------------------------------------------
// 8 independent FFMA chains per thread, 1024 threads per SM-resident CTA pair: issue-bound on the FMA pipe.
__global__ void __launch_bounds__(512) ffma_peak_kernel(float *out, int iters, float a, float b) {
    float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int k = 0; k < 16; k++) {
            x0 = __fmaf_rn(x0, a, b); x1 = __fmaf_rn(x1, a, b); x2 = __fmaf_rn(x2, a, b); x3 = __fmaf_rn(x3, a, b);
            x4 = __fmaf_rn(x4, a, b); x5 = __fmaf_rn(x5, a, b); x6 = __fmaf_rn(x6, a, b); x7 = __fmaf_rn(x7, a, b);
        }
    }
    float s = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
    if (s == 12345.678f) out[0] = s;
}

extern "C" int trt_measure_fp32_peak(int device, float *tflops, int *sm_count, int *clock_khz) {
    if (!tflops) return TRT_ERR_INVALID;
    trt_ctx *ctx = nullptr;     // CK() target for error text
    trt_ctx tmp;
    ctx = &tmp;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); g_create_err = "no CUDA device"; return TRT_ERR_NO_DEVICE; }
    int rc = [&]() -> int {
        CK(cudaSetDevice(device));
        cudaDeviceProp prop;
        CK(cudaGetDeviceProperties(&prop, device));
        if (sm_count) *sm_count = prop.multiProcessorCount;
        if (clock_khz) *clock_khz = prop.clockRate;
        float *d = nullptr;
        CK(cudaMalloc(&d, 256));
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0));
        CK(cudaEventCreate(&e1));
        const int blocks = prop.multiProcessorCount * 4, iters = 4096;
        float best = 0.0f;
        for (int rep = 0; rep < 6; rep++) {
            CK(cudaEventRecord(e0));
            ffma_peak_kernel<<<blocks, 512>>>(d, iters, 1.0000001f, 1e-9f);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms = 0.0f;
            CK(cudaEventElapsedTime(&ms, e0, e1));
            const double flops = 2.0 * 8 * 16 * (double)iters * 512.0 * blocks;
            const float tf = (float)(flops / (ms * 1e-3) / 1e12);
            if (rep >= 2 && tf > best) best = tf;
        }
        *tflops = best;
        cudaEventDestroy(e0);
        cudaEventDestroy(e1);
        cudaFree(d);
        return TRT_OK;
    }();
    if (rc) g_create_err = tmp.err;
    return rc;
}

// ---- self-test of div3_rn against the IEEE `/` operator (tests/test_gpu_division.py) ---------------------
__device__ __forceinline__ uint32_t mix32(uint32_t v) { v ^= v >> 16; v *= 0x7feb352du; v ^= v >> 15; v *= 0x846ca68bu; v ^= v >> 16; return v; }
__global__ void div3_selftest_kernel(unsigned long long n, uint32_t seed, int dist, unsigned long long *mismatch, float *first_bad) {
    unsigned long long bad = 0;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        uint32_t h = mix32((uint32_t)i * 4u + seed), g = mix32((uint32_t)(i >> 32) + 0x9e3779b9u * ((uint32_t)i * 4u + 1u) + seed);
        uint32_t b[4] = {h, mix32(h + g), mix32(g ^ 0x85ebca6bu), mix32(h * 3u + g)};
        float v[4];
        for (int c = 0; c < 4; c++) {
            uint32_t bits = b[c];
            if (dist == 1) {            // scene-like magnitudes: exponent in [2^-24, 2^8), some exact zeros
                bits = (bits & 0x807fffffu) | ((103u + (b[(c + 1) & 3] >> 27)) << 23);
                if ((b[(c + 2) & 3] & 15u) == 0u && c < 3) bits &= 0x80000000u;
            } else if (dist == 2) {     // mantissa corner cases
                const uint32_t m[8] = {0u, 1u, 0x7fffffu, 0x7ffffeu, 0x400000u, 0x3fffffu, 0x555555u, 0x2aaaaau};
                bits = (bits & 0x80000000u) | ((120u + ((bits >> 8) & 15u)) << 23) | m[bits & 7u];
            }                           // dist 0: raw random bits (NaN, inf, denormals included)
            v[c] = __uint_as_float(bits);
        }
        const float3 got = div3_rn(make_float3(v[0], v[1], v[2]), v[3]);
        const float ex[3] = {v[0] / v[3], v[1] / v[3], v[2] / v[3]}, gt[3] = {got.x, got.y, got.z};
        for (int c = 0; c < 3; c++) {
            const bool same = (__float_as_uint(ex[c]) == __float_as_uint(gt[c])) || (ex[c] != ex[c] && gt[c] != gt[c]);
            if (!same) {
                if (bad == 0 && atomicAdd(mismatch + 1, 1ull) == 0ull) { first_bad[0] = v[c]; first_bad[1] = v[3]; first_bad[2] = ex[c]; first_bad[3] = gt[c]; }
                bad++;
            }
        }
    }
    if (bad) atomicAdd(mismatch, bad);
}

extern "C" int trt_selftest_div3(int device, uint64_t n_cases, uint32_t seed, int distribution, uint64_t *mismatches, float *first_bad4) {
    if (!mismatches) return TRT_ERR_INVALID;
    trt_ctx tmp;
    trt_ctx *ctx = &tmp;
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); g_create_err = "no CUDA device"; return TRT_ERR_NO_DEVICE; }
    int rc = [&]() -> int {
        CK(cudaSetDevice(device));
        unsigned long long *d = nullptr;
        float *fb = nullptr;
        CK(cudaMalloc(&d, 16));
        CK(cudaMalloc(&fb, 16));
        CK(cudaMemset(d, 0, 16));
        CK(cudaMemset(fb, 0, 16));
        div3_selftest_kernel<<<148 * 8, 256>>>(n_cases, seed, distribution, d, fb);
        CK(cudaGetLastError());
        CK(cudaDeviceSynchronize());
        unsigned long long h[2];
        CK(cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost));
        if (first_bad4) CK(cudaMemcpy(first_bad4, fb, 16, cudaMemcpyDeviceToHost));
        *mismatches = h[0];
        cudaFree(d);
        cudaFree(fb);
        return TRT_OK;
    }();
    if (rc) g_create_err = tmp.err;
    return rc;
}
