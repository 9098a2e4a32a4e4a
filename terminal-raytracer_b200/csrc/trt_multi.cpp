// trt_multi.cpp -- one frame over several GPUs of one host process, behind the C ABI (include/trt_b200.h).
//
// The reference has one entry and one device (lib.rs:100, :248-250).  A host that calls the C ABI -- the Rust run()
// of INTEGRATION.md, or trt_run below it -- reaches N GPUs through ONE object: trt_multi_create makes one row-sharded
// context per device (stripes of rows dealt round-robin, trt_config.shard_*), trt_multi_render_frame traces the
// shards concurrently (every call below is asynchronous on its context's stream), lets every GPU quantise its own
// rows, adds up the row lengths on the host and has every GPU copy its rows over its own PCIe link straight to
// their place in the caller's buffer (row-addressed streams, trt_api.cu).  No pixel leaves its GPU, so there is
// nothing for a collective to do inside one process; the multi-PROCESS driver (multigpu.py, one rank per GPU, as the
// benchmark contract wants) exchanges the same row lengths with one NCCL all-gather.  Pixel indices and RNG seeds
// are global, so the stream is byte-identical to a single-GPU frame.
#include <cstring>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/trt_b200.h"

struct trt_multi {
    std::vector<trt_ctx *> ctx;
    // frames over GPUs (trt_multi_render_frames): one whole-frame context per device, created on first use
    std::vector<trt_ctx *> full;
    std::vector<int32_t> devices;
    trt_config cfg{};
    std::vector<trt_primitive> prims;             // the scene, for contexts created later
    void *d_gather = nullptr;                     // on devices[0]: the means of all frames of a batch, frame-major
    size_t gather_cap = 0;
    std::vector<std::vector<uint32_t>> rows;      // rows of every shard at rows_height
    uint32_t rows_height = 0;
    std::vector<uint32_t> len;                    // scratch: one shard's row lengths
    std::vector<uint64_t> row_off;                // height + 1
    std::string err;
};

static thread_local std::string g_multi_err;

static int mfail(trt_multi *m, int code, const std::string &msg) {
    if (m) m->err = msg; else g_multi_err = msg;
    return code;
}
static int cfail(trt_multi *m, int code, trt_ctx *c) { return mfail(m, code, trt_last_error(c)); }

extern "C" const char *trt_multi_last_error(const trt_multi *m) { return m ? m->err.c_str() : g_multi_err.c_str(); }

extern "C" void trt_multi_destroy(trt_multi *m) {
    if (!m) return;
    for (trt_ctx *c : m->ctx) trt_destroy(c);
    for (trt_ctx *c : m->full) trt_destroy(c);
    if (m->d_gather) { cudaSetDevice(m->devices[0]); cudaFree(m->d_gather); }
    delete m;
}

extern "C" int trt_multi_create(const int32_t *devices, uint32_t n_devices, const trt_config *cfg, trt_multi **out) {
    if (!devices || !n_devices || !cfg || !out) return mfail(nullptr, TRT_ERR_INVALID, "trt_multi_create: null argument");
    *out = nullptr;
    trt_multi *m = new (std::nothrow) trt_multi();
    if (!m) return mfail(nullptr, TRT_ERR_NOMEM, "trt_multi_create: out of host memory");
    for (uint32_t i = 0; i < n_devices; i++) {
        trt_config c = *cfg;
        c.device = devices[i];
        c.shard_index = i;
        c.shard_count = n_devices;
        if (c.stripe_rows == 0) c.stripe_rows = 1;      // finest deal = best balance (tools/shard_sweep.py: 8 shards of C3 within 0.07 % of each other)
        trt_ctx *x = nullptr;
        const int rc = trt_create(&c, &x);
        if (rc != TRT_OK) {
            g_multi_err = std::string("trt_multi_create: device ") + std::to_string(devices[i]) + ": " + trt_last_error(nullptr);
            trt_multi_destroy(m);
            return rc;
        }
        m->ctx.push_back(x);
    }
    m->devices.assign(devices, devices + n_devices);
    m->cfg = *cfg;
    *out = m;
    return TRT_OK;
}

extern "C" uint32_t trt_multi_size(const trt_multi *m) { return m ? (uint32_t)m->ctx.size() : 0u; }
extern "C" trt_ctx *trt_multi_context(trt_multi *m, uint32_t i) { return m && i < m->ctx.size() ? m->ctx[i] : nullptr; }

extern "C" int trt_multi_upload_scene(trt_multi *m, const trt_primitive *prims, uint32_t count) {
    if (!m) return TRT_ERR_INVALID;
    for (trt_ctx *c : m->ctx) {
        const int rc = trt_upload_scene(c, prims, count);
        if (rc != TRT_OK) return cfail(m, rc, c);
    }
    for (trt_ctx *c : m->full) {
        const int rc = trt_upload_scene(c, prims, count);
        if (rc != TRT_OK) return cfail(m, rc, c);
    }
    m->prims.assign(prims, prims + count);
    return TRT_OK;
}

extern "C" int trt_multi_render_frame(trt_multi *m, const trt_uniforms *u, int mode, uint8_t *out, size_t cap, size_t *len) {
    if (!m || !u) return TRT_ERR_INVALID;
    if (len) *len = 0;
    for (trt_ctx *c : m->ctx) {                               // all shards in flight before anything waits
        const int rc = trt_trace_async(c, u);
        if (rc != TRT_OK) return cfail(m, rc, c);
    }
    if (mode == TRT_MODE_NONE) {
        for (trt_ctx *c : m->ctx) { const int rc = trt_sync(c); if (rc != TRT_OK) return cfail(m, rc, c); }
        return TRT_OK;
    }
    for (trt_ctx *c : m->ctx) {
        const int rc = trt_quantise_rows_async(c, mode);
        if (rc != TRT_OK) return cfail(m, rc, c);
    }
    const uint32_t h = u->height;
    if (m->rows_height != h) {
        m->rows.assign(m->ctx.size(), {});
        for (size_t i = 0; i < m->ctx.size(); i++) {
            m->rows[i].resize(h);
            m->rows[i].resize(trt_shard_rows(m->ctx[i], h, m->rows[i].data(), h));
        }
        m->rows_height = h;
    }
    m->row_off.assign((size_t)h + 1, 0);
    m->len.resize(h);
    for (size_t i = 0; i < m->ctx.size(); i++) {               // waits for shard i's quantiser (its trace is the frame time)
        uint32_t n = 0;
        const int rc = trt_row_lengths(m->ctx[i], m->len.data(), h, &n);
        if (rc != TRT_OK) return cfail(m, rc, m->ctx[i]);
        if (n != m->rows[i].size()) return mfail(m, TRT_ERR_INVALID, "trt_multi_render_frame: shard row count changed under the frame");
        for (uint32_t k = 0; k < n; k++) m->row_off[(size_t)m->rows[i][k] + 1] = m->len[k];
    }
    for (uint32_t y = 0; y < h; y++) m->row_off[(size_t)y + 1] += m->row_off[y];
    const size_t total = (size_t)m->row_off[h];
    if (len) *len = total;
    if (total > cap || (!out && total)) return mfail(m, TRT_ERR_TOO_SMALL, "output buffer too small for the frame stream");
    for (trt_ctx *c : m->ctx) {                               // every GPU copies over its own link, concurrently
        const int rc = trt_read_stream_rows(c, out, cap, m->row_off.data(), h, 0);
        if (rc != TRT_OK) return cfail(m, rc, c);
    }
    for (trt_ctx *c : m->ctx) { const int rc = trt_sync(c); if (rc != TRT_OK) return cfail(m, rc, c); }
    return TRT_OK;
}

// float frame / variance of the whole image: every shard fills in its rows
extern "C" int trt_multi_read_pixels(trt_multi *m, float *dst4, size_t n_pixels) {
    if (!m) return TRT_ERR_INVALID;
    for (trt_ctx *c : m->ctx) { const int rc = trt_read_pixels(c, dst4, n_pixels); if (rc != TRT_OK) return cfail(m, rc, c); }
    return TRT_OK;
}
extern "C" int trt_multi_read_variance(trt_multi *m, trt_variance *dst, size_t n_pixels) {
    if (!m) return TRT_ERR_INVALID;
    for (trt_ctx *c : m->ctx) { const int rc = trt_read_variance(c, dst, n_pixels); if (rc != TRT_OK) return cfail(m, rc, c); }
    return TRT_OK;
}

extern "C" int trt_multi_get_frame_stats(trt_multi *m, trt_frame_stats *out) {
    if (!m || !out) return TRT_ERR_INVALID;
    memset(out, 0, sizeof *out);
    for (trt_ctx *c : m->ctx) {
        trt_frame_stats s{};
        const int rc = trt_get_frame_stats(c, &s);
        if (rc != TRT_OK) return cfail(m, rc, c);
        out->samples_traced += s.samples_traced; out->bounces += s.bounces; out->pixels += s.pixels;
    }
    return TRT_OK;
}

// slowest shard's device times of the last frame (the frame is as long as its longest shard)
extern "C" int trt_multi_get_timings(trt_multi *m, trt_timings *out) {
    if (!m || !out) return TRT_ERR_INVALID;
    memset(out, 0, sizeof *out);
    for (trt_ctx *c : m->ctx) {
        trt_timings t{};
        const int rc = trt_get_timings(c, &t);
        if (rc != TRT_OK) return cfail(m, rc, c);
        if (t.trace_ms > out->trace_ms) out->trace_ms = t.trace_ms;
        if (t.quantise_ms > out->quantise_ms) out->quantise_ms = t.quantise_ms;
        if (t.d2h_ms > out->d2h_ms) out->d2h_ms = t.d2h_ms;
        if (t.total_ms > out->total_ms) out->total_ms = t.total_ms;
        out->kernel_launches += t.kernel_launches;
    }
    return TRT_OK;
}

// ---- the second axis: the frames of a still camera over the GPUs -------------------------------------------
// A pixel's samples cannot be split over GPUs (one serial random-number chain, shader.wgsl:507,517,565), but frames are
// independent given (seed, frame_number) and meet only in the temporal blend (:610-619), which is per pixel and
// sequential in the frame index.  GPU g traces frames f = g, g + G, ... as ONE launch each (frame batch), the per-frame
// means are copied to the first GPU (peer copies over NVLink) and blended there IN FRAME ORDER by the same function a
// single GPU uses: accumulation, pixels and the byte stream are bit-identical to n trt_render_frame calls on one GPU.
// This is what fills several GPUs at terminal sizes (C2: 100 frames of 400x200), where row shards are latency-bound.
extern "C" int trt_multi_render_frames(trt_multi *m, const trt_uniforms *u, uint32_t n, int mode, uint8_t *out, size_t cap, size_t *len) {
    if (!m || !u || n == 0u) return TRT_ERR_INVALID;
    if (len) *len = 0;
    const uint32_t G = (uint32_t)m->devices.size();
    if (m->full.empty()) {
        for (uint32_t g = 0; g < G; g++) {
            trt_config c = m->cfg;
            c.device = m->devices[g];
            c.shard_index = 0; c.shard_count = 1; c.stripe_rows = 1;
            trt_ctx *x = nullptr;
            int rc = trt_create(&c, &x);
            if (rc != TRT_OK) return mfail(m, rc, std::string("trt_multi_render_frames: device ") + std::to_string(c.device) + ": " + trt_last_error(nullptr));
            m->full.push_back(x);
            rc = trt_upload_scene(x, m->prims.data(), (uint32_t)m->prims.size());
            if (rc != TRT_OK) return cfail(m, rc, x);
        }
    }
    const size_t np = (size_t)u[0].width * u[0].height;
    // every GPU's share, queued before anything waits
    std::vector<std::vector<trt_uniforms>> share(G);
    for (uint32_t f = 0; f < n; f++) share[f % G].push_back(u[f]);
    for (uint32_t g = 0; g < G; g++) {
        if (share[g].empty()) continue;
        const int rc = trt_trace_frames_async(m->full[g], share[g].data(), (uint32_t)share[g].size());
        if (rc != TRT_OK) return cfail(m, rc, m->full[g]);
    }
    const int root_dev = m->devices[0];
    if (m->gather_cap < (size_t)n * np * 16) {
        cudaSetDevice(root_dev);
        if (m->d_gather) cudaFree(m->d_gather);
        m->d_gather = nullptr; m->gather_cap = 0;
        if (cudaMalloc(&m->d_gather, (size_t)n * np * 16) != cudaSuccess) { cudaGetLastError(); return mfail(m, TRT_ERR_NOMEM, "trt_multi_render_frames: gather buffer"); }
        m->gather_cap = (size_t)n * np * 16;
    }
    for (uint32_t g = 0; g < G; g++) {                          // means -> first GPU, into frame order
        if (share[g].empty()) continue;
        int rc = trt_sync(m->full[g]);
        if (rc != TRT_OK) return cfail(m, rc, m->full[g]);
        void *means = nullptr;
        size_t stride = 0;
        rc = trt_frame_means(m->full[g], &means, &stride);
        if (rc != TRT_OK) return cfail(m, rc, m->full[g]);
        for (size_t k = 0; k < share[g].size(); k++) {
            const size_t f = g + k * G;
            if (cudaMemcpyPeer((char *)m->d_gather + f * np * 16, root_dev, (const char *)means + k * stride, m->devices[g], np * 16) != cudaSuccess) {
                cudaGetLastError();
                return mfail(m, TRT_ERR_CUDA, "trt_multi_render_frames: peer copy of a frame's means failed");
            }
        }
    }
    {   // only the last frame of a sequence leaves its variance behind (shader.wgsl:608 overwrites every frame)
        const uint32_t gl = (n - 1u) % G;
        if (gl != 0u) {
            void *src = nullptr, *dst = nullptr;
            size_t bytes = 0;
            if (trt_device_buffer(m->full[gl], 2, &src, &bytes) != TRT_OK || trt_device_buffer(m->full[0], 2, &dst, &bytes) != TRT_OK)
                return mfail(m, TRT_ERR_INVALID, "trt_multi_render_frames: variance buffers");
            if (cudaMemcpyPeer(dst, root_dev, src, m->devices[gl], np * sizeof(trt_variance)) != cudaSuccess) { cudaGetLastError(); return mfail(m, TRT_ERR_CUDA, "variance peer copy failed"); }
        }
    }
    int rc = trt_blend_frames_async(m->full[0], m->d_gather, u, n);
    if (rc != TRT_OK) return cfail(m, rc, m->full[0]);
    if (mode == TRT_MODE_NONE) { rc = trt_sync(m->full[0]); return rc == TRT_OK ? TRT_OK : cfail(m, rc, m->full[0]); }
    rc = trt_quantise_async(m->full[0], mode, nullptr);
    if (rc != TRT_OK) return cfail(m, rc, m->full[0]);
    size_t nlast = 0;
    rc = trt_read_stream(m->full[0], out, cap, &nlast);
    if (len) *len = nlast;
    return rc == TRT_OK ? TRT_OK : cfail(m, rc, m->full[0]);
}

// the whole-frame context the blended batch lives in (pixels / accumulation / variance of trt_multi_render_frames)
extern "C" trt_ctx *trt_multi_frames_context(trt_multi *m) { return m && !m->full.empty() ? m->full[0] : nullptr; }
