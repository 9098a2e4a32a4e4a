/*
 * trt_b200.h -- C ABI of the B200-native path tracer that drops in for Terminal-Raytracer's
 * per-pixel path-tracing hot path.
 *
 * What it replaces (all citations relative to /root/reference/):
 *   the wgpu device / buffer / pipeline / staging layer inside `run()`  src/lib.rs:248-352,371-383
 *   the per-frame upload, dispatch, copy and blocking map               src/lib.rs:442-477,538-539
 *   the WGSL compute shader                                             src/shader.wgsl:1-621
 *   the rayon row formatter (quantiser)                                 src/lib.rs:503-532
 * plus the host logic around them that defines the inputs: scene JSON schema (lib.rs:52-154),
 * uniforms (lib.rs:417-439), camera (camera.rs:4-26), key semantics (lib.rs:392-412), CLI flags
 * (main.rs:6-39) and the `run(full_color, verbose, scene_path)` entry (lib.rs:100).
 *
 * Conventions: plain C types only; every function returns trt_status (0 = ok, < 0 = error) unless
 * stated; the caller owns every host pointer and the library copies before returning; a context
 * is single-owner (one host thread at a time); no exception or abort crosses this boundary;
 * there is NO CPU fallback -- without a CUDA device trt_create fails with TRT_ERR_NO_DEVICE.
 * The reference-side binding (Rust `extern "C"` block) is in INTEGRATION.md.
 */
#ifndef TRT_B200_H
#define TRT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TRT_ABI_VERSION 1

typedef enum {
    TRT_OK = 0,
    TRT_ERR_INVALID = -1,      /* bad argument                                             */
    TRT_ERR_NO_DEVICE = -2,    /* no CUDA device / driver (there is no CPU fallback)       */
    TRT_ERR_CUDA = -3,         /* a CUDA call failed; text in trt_last_error               */
    TRT_ERR_NO_SCENE = -4,     /* render before trt_upload_scene                           */
    TRT_ERR_TOO_SMALL = -5,    /* output buffer too small; *len holds the size needed      */
    TRT_ERR_PARSE = -6,        /* scene JSON: syntax error, missing mandatory key          */
    TRT_ERR_IO = -7,           /* scene file unreadable                                    */
    TRT_ERR_NOMEM = -8
} trt_status;

/* ---- wire formats: byte-identical to the reference's #[repr(C)] Pod structs ------------------ */

/* src/vec3.rs:5-12 -- 16 bytes */
typedef struct trt_vec3 { float x, y, z, _padding; } trt_vec3;

/* src/primitive.rs:4-34, src/shader.wgsl:8-32 -- 180 bytes, AoS tagged union */
typedef struct trt_primitive {
    uint32_t primitive_type;                 /* 0 sphere, 1 plane, 2 triangle */
    uint32_t _padding1, _padding2, _padding3;
    trt_vec3 color;
    trt_vec3 emission;
    float reflectivity, _padding4, _padding5, _padding6;
    trt_vec3 sphere_center;
    float sphere_radius;
    float _sphere_padding[3];
    trt_vec3 plane_point;
    trt_vec3 plane_normal;
    trt_vec3 triangle_v0, triangle_v1, triangle_v2;
    float _triangle_padding;
} trt_primitive;

/* src/lib.rs:23-50 -- 160 bytes on the host; the shader reads the first 112 (shader.wgsl:49-69).
 * The four grid fields are dead in the reference (never read by the shader) and ignored here. */
typedef struct trt_uniforms {
    uint32_t width, height, samples_per_pixel, max_depth;
    uint32_t seed, frame_number, primitive_count, _padding1;
    float aspect_ratio, char_aspect_ratio, fov_rad, _padding2;
    trt_vec3 camera_pos, camera_forward, camera_right, camera_up;
    trt_vec3 grid_min, inv_cell_size;
    uint32_t grid_dims[3];
    uint32_t _padding3;
} trt_uniforms;

/* src/shader.wgsl:71-74 -- 8 bytes */
typedef struct trt_variance { float variance; uint32_t sample_count; } trt_variance;

/* ---- context --------------------------------------------------------------------------------- */
typedef struct trt_ctx trt_ctx;

/* arithmetic convention of the trace kernel (DESIGN.md "Arithmetic conventions") */
#define TRT_CONV_AS_WRITTEN 0     /* one IEEE rounding per WGSL operator, no contraction */
#define TRT_CONV_FUSED      1     /* fused multiply-add inside the vector helpers        */
#define TRT_CONV_NATIVE     2     /* what a GPU shader compiler gives the same WGSL: FFMA contraction and the SM's
                                     reciprocal / rsqrt / sqrt / sin / cos units (1-2 ulp, inside WGSL's accuracy
                                     bounds); tracks conv 0 to rounding noise, not bit for bit */

/* output mode of the quantiser: lib.rs:509-513 (truecolor) / lib.rs:515-524 (ASCII) */
#define TRT_MODE_ASCII      0
#define TRT_MODE_TRUECOLOR  1
#define TRT_MODE_NONE       2     /* trace + accumulate only, no byte stream */

#define TRT_ACCEL_NONE      0
#define TRT_ACCEL_GRID      1

typedef struct trt_config {
    uint32_t abi_version;         /* TRT_ABI_VERSION */
    int32_t  device;              /* CUDA device ordinal (one context = one GPU) */
    uint32_t max_width, max_height;   /* buffers are sized once (lib.rs:269-319 does the same) */
    uint32_t conv;                /* TRT_CONV_* */
    uint32_t count_work;          /* 1: run the instrumented kernel that fills trt_counters */
    /* Row sharding of one frame over `shard_count` contexts (one per GPU).  Rows are dealt in
     * stripes of `stripe_rows` rows, stripe s belongs to shard (s % shard_count).  Pixel indices
     * and RNG seeds always use the global (x, y, width): the union of the shards' rows is
     * bit-identical to a single-context frame.  shard_count 0 or 1 = whole frame. */
    uint32_t shard_index, shard_count, stripe_rows;
    /* TRT_ACCEL_NONE: the reference's linear scan (the parity contract).  TRT_ACCEL_GRID (opt-in, conv 2 only, outside the
     * parity contract, reported separately): hit_scene through the uniform grid the reference builds and never uses
     * (lib.rs:181-244): same (t, index) per ray, O(cells) instead of O(primitives) tests. */
    uint32_t accel;
    uint32_t reserved[6];
} trt_config;

/* work counters of the instrumented build (same fields as the oracle's) */
typedef struct trt_counters {
    uint64_t pixels, samples_base, samples_extra, closest_scans, hits, misses, shadow_scans,
             shadow_unoccluded, shadow_lit, shadow_needed, mirror, diffuse, roulette_tests,
             roulette_kills, prim_tests, prim_tests_min, flops_written, flops_min, rng_draws;
} trt_counters;

/* device-side timings of the last frame, CUDA events on the context's stream, milliseconds */
typedef struct trt_timings {
    float trace_ms, quantise_ms, d2h_ms, total_ms;
    uint32_t kernel_launches;     /* kernels launched by the last call */
    uint32_t _pad;
} trt_timings;

/* always-on statistics of the last traced frame (this shard's rows) */
typedef struct trt_frame_stats {
    uint64_t samples_traced;      /* ray_color calls: sum over pixels of base + adaptive extra samples */
    uint64_t bounces;             /* closest-hit scans (shader.wgsl:445) */
    uint64_t pixels;
    uint64_t reserved;
} trt_frame_stats;

/* replaces Instance/adapter/device + buffer creation, lib.rs:248-319 */
int trt_create(const trt_config *cfg, trt_ctx **out);
void trt_destroy(trt_ctx *ctx);
/* message for the last error on this context (ctx may be NULL: last trt_create failure) */
const char *trt_last_error(const trt_ctx *ctx);

/* replaces create_buffer_init(primitive_buffer), lib.rs:301-305: takes the reference's 180-byte AoS
 * verbatim, in scan order, and re-lays it as structure-of-arrays device buffers. */
int trt_upload_scene(trt_ctx *ctx, const trt_primitive *prims, uint32_t count);

/* replaces write_buffer + dispatch + copy + map/poll + row formatting, lib.rs:442-477,503-532.
 * Traces one frame (shader.wgsl main), updates accumulation/variance, quantises on the device and
 * copies ONLY the byte stream to `out`.  Blocks until the bytes are in `out` (like poll(Wait)).
 * mode TRT_MODE_NONE skips quantise and copy (out may be NULL).  *len = bytes written (or needed). */
int trt_render_frame(trt_ctx *ctx, const trt_uniforms *u, int mode, uint8_t *out, size_t cap, size_t *len);

/* n iterations of the reference's frame loop (lib.rs:414-560) in one call: the frames of a still camera --
 * u[0..n) equal in everything but seed and frame_number -- are traced by ONE launch over n x pixels work items
 * and blended in frame order (shader.wgsl:610-619), so accumulation / pixels / variance end up bit-identical to
 * n trt_render_frame calls.  emit_all != 0: the byte streams of all n frames, back to back, go to `out`
 * (frame_len[i] = bytes of frame i, may be NULL); emit_all == 0: only the last frame's stream (what a
 * terminal would still be showing).  mode TRT_MODE_NONE: no stream.  Frames are processed in chunks that
 * keep the staging buffers (16 B per pixel and frame, plus the streams) within 6 GB.  A sharded context
 * accepts TRT_MODE_NONE only. */
int trt_render_frames(trt_ctx *ctx, const trt_uniforms *u, uint32_t n, int mode, int emit_all, uint8_t *out, size_t cap,
                      size_t *len, size_t *frame_len_or_null);

/* the two halves of trt_render_frame, asynchronous on the context's stream (device-resident use:
 * benchmarks, multi-GPU drivers).  trt_sync waits for the stream. */
int trt_trace_async(trt_ctx *ctx, const trt_uniforms *u);
int trt_quantise_async(trt_ctx *ctx, int mode, size_t *len_needed_or_null);
int trt_sync(trt_ctx *ctx);
int trt_read_stream(trt_ctx *ctx, uint8_t *out, size_t cap, size_t *len);

/* parity / checkpoint access (the reference keeps these device-only: lib.rs:269-284).
 * Layout: float4 per pixel / trt_variance per pixel, rows of the whole frame; rows owned by other
 * shards are left untouched in `dst`. */
int trt_read_pixels(trt_ctx *ctx, float *dst4, size_t n_pixels);
int trt_read_accumulation(trt_ctx *ctx, float *dst4, size_t n_pixels);
int trt_read_variance(trt_ctx *ctx, trt_variance *dst, size_t n_pixels);
int trt_write_accumulation(trt_ctx *ctx, const float *src4, size_t n_pixels);
int trt_get_counters(trt_ctx *ctx, trt_counters *out);
int trt_get_timings(trt_ctx *ctx, trt_timings *out);
/* name of the trace kernel the last frame ran (static builds, or a scene-specialised kernel compiled at upload time
 * by NVRTC -- the counterpart of the reference's create_shader_module + create_compute_pipeline, lib.rs:321-352) */
const char *trt_trace_kernel_name(const trt_ctx *ctx);
int trt_get_frame_stats(trt_ctx *ctx, trt_frame_stats *out);
/* register-resident FFMA loop on every SM: the measured FP32 FMA peak of this GPU in TFLOP/s
 * (roofline denominator cross-check; nominal = SMs x 128 lanes x 2 x clock) */
int trt_measure_fp32_peak(int device, float *tflops, int *sm_count, int *clock_khz);
/* device self-test: the kernels' shared-reciprocal vec3 division against the IEEE `/` operator on
 * n_cases pseudo-random (a.x, a.y, a.z, s) tuples; distribution 0 raw bits, 1 scene-like, 2 mantissa
 * corner cases.  *mismatches = components whose bits differ (NaN == NaN). */
int trt_selftest_div3(int device, uint64_t n_cases, uint32_t seed, int distribution, uint64_t *mismatches, float *first_bad4);

/* multi-GPU plumbing: raw device pointers of this context's buffers so that the host driver can
 * hand them to NCCL (gather of the shards' pixel rows), and a quantise that reads a caller-owned
 * device frame.  which: 0 pixels (float4), 1 accumulation, 2 variance, 3 byte stream, 4 frame statistics,
 * 5 row lengths of the last trt_quantise_rows_async (u32 per local row). */
int trt_device_buffer(trt_ctx *ctx, int which, void **dptr, size_t *bytes);
int trt_cuda_stream(trt_ctx *ctx, void **stream);
/* rows this shard owns, ascending; returns the count (writes at most cap entries) */
uint32_t trt_shard_rows(const trt_ctx *ctx, uint32_t height, uint32_t *rows, uint32_t cap);
/* the same rule without a context (host-side planning, tests) */
uint32_t trt_shard_rows_of(uint32_t shard_index, uint32_t shard_count, uint32_t stripe_rows, uint32_t height,
                           uint32_t *rows, uint32_t cap);
/* pack this shard's pixel rows contiguously (local row order) into d_dst (device, float4) */
int trt_pack_shard_pixels(trt_ctx *ctx, void *d_dst, size_t dst_bytes);
/* quantise a frame assembled from per-shard packed blocks: block r holds shard r's rows, blocks
 * are `block_stride_bytes` apart in d_gathered (what an NCCL gather produces on the root) */
int trt_quantise_gathered(trt_ctx *ctx, const void *d_gathered, size_t block_stride_bytes,
                          uint32_t width, uint32_t height, uint32_t shard_count, uint32_t stripe_rows,
                          int mode, size_t *len_needed_or_null);

/* Quantiser that ALLOCATES NOTHING.  No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same kernel and same bytes as trt_quantise_async / trt_render_frame (lib.rs:503-532 on the device); the kernel's
 * scratch (tile ticket, look-back states, stream length) is a static device array of TRT_NOALLOC_SCRATCH_BYTES that
 * every GPU holds from the moment it loads the library until exit.  d_pixels: width x height float4 pixels, row-major,
 * device memory, 16-byte aligned; d_out: device buffer of out_cap bytes for the stream (width * height * 26 + 2 *
 * height bytes always suffice; a stream longer than out_cap is not written, *d_len still tells its size); d_len: device
 * uint64_t, receives the stream's length.  Asynchronous on `cuda_stream` (a cudaStream_t, NULL = the default stream);
 * no context, no cudaMalloc / cudaFree, no host synchronisation.  The first call per device looks the array's address
 * up (make that call outside a graph capture).  Frames of more than TRT_NOALLOC_MAX_PIXELS pixels are rejected with
 * TRT_ERR_INVALID and a text in `err` (NULL allowed).  All calls on a device share the array: calls that overlap on
 * the GPU (different streams without a dependency between them) race -- serialise them. */
#define TRT_NOALLOC_MAX_PIXELS    8388608u   /* 8192 tiles of 1024 pixels; a 3840x2160 frame (8 294 400) fits */
#define TRT_NOALLOC_SCRATCH_BYTES 65792u
int trt_quantise_noalloc(int device, const void *d_pixels, uint32_t width, uint32_t height, int mode, void *d_out,
                         size_t out_cap, void *d_len, void *cuda_stream, char *err, size_t err_cap);

/* pinned host memory for `out` buffers (optional; any host pointer works) */
int trt_alloc_host(size_t bytes, void **out);
void trt_free_host(void *p);
/* page-lock memory the caller owns (e.g. a POSIX shared-memory segment several ranks copy their rows into) */
int trt_host_register(void *p, size_t bytes);
int trt_host_unregister(void *p);

/* ---- row-addressed streams (sharded frames without moving a pixel) -------------------------------
 * The frame's stream is its rows back to back (lib.rs:503-532) and a row's bytes depend on that row only: every
 * shard quantises ITS rows (trt_quantise_rows_async: stream of the rows it owns, in ascending row order, plus the
 * byte length of each -- device buffer 5, u32 per local row), the lengths are exchanged (4 B per row instead of
 * 16 B per pixel), and every shard copies its rows device -> host straight to their place in ONE host buffer
 * (trt_read_stream_rows; row_off = height + 1 byte offsets of the whole frame's rows, row_off[height] = total).
 * wait == 0: the copies are queued on the context's stream (trt_sync to wait). */
int trt_quantise_rows_async(trt_ctx *ctx, int mode);
int trt_row_lengths(trt_ctx *ctx, uint32_t *out, uint32_t cap, uint32_t *n_rows);
int trt_read_stream_rows(trt_ctx *ctx, uint8_t *host_base, size_t cap, const uint64_t *row_off, uint32_t height, int wait);

/* ---- pipelined frames: replaces the strictly serial dispatch / map+poll / format / print of lib.rs:442-477,
 * 503-545.  Up to two frames in flight: submit queues trace + quantise and returns; collect hands out the OLDEST
 * submitted frame, its device -> host copy running on a second stream while the next frame is being traced.  Same
 * frames, same order, same bytes as trt_render_frame called once per frame. */
int trt_frame_submit(trt_ctx *ctx, const trt_uniforms *u, int mode);
int trt_frame_collect(trt_ctx *ctx, uint8_t *out, size_t cap, size_t *len);

/* ---- one frame over several GPUs of one process (the reference has one device: lib.rs:248-250) --------
 * One row-sharded context per device (cfg->device / shard_* are filled in; stripe_rows 0 = 1).  render_frame has
 * trt_render_frame's contract: blocks until the assembled stream is in `out` (pinned memory lets the GPUs copy
 * concurrently); the stream is byte-identical to a single-GPU frame. */
typedef struct trt_multi trt_multi;
int trt_multi_create(const int32_t *devices, uint32_t n_devices, const trt_config *cfg, trt_multi **out);
void trt_multi_destroy(trt_multi *m);
const char *trt_multi_last_error(const trt_multi *m);
uint32_t trt_multi_size(const trt_multi *m);
trt_ctx *trt_multi_context(trt_multi *m, uint32_t i);
int trt_multi_upload_scene(trt_multi *m, const trt_primitive *prims, uint32_t count);
int trt_multi_render_frame(trt_multi *m, const trt_uniforms *u, int mode, uint8_t *out, size_t cap, size_t *len);
int trt_multi_read_pixels(trt_multi *m, float *dst4, size_t n_pixels);
int trt_multi_read_variance(trt_multi *m, trt_variance *dst, size_t n_pixels);
int trt_multi_get_frame_stats(trt_multi *m, trt_frame_stats *out);
int trt_multi_get_timings(trt_multi *m, trt_timings *out);
/* The second axis (north_star: "by image tiles or by sample batches"): n frames of a still camera -- u[0..n) equal but for
 * seed and frame_number, as in trt_render_frames -- dealt over the GPUs (frame f on GPU f mod G, one launch per GPU),
 * their per-frame means gathered on the first GPU and blended there in frame order (shader.wgsl:610-619): accumulation,
 * pixels, variance and the last frame's stream are bit-identical to n trt_render_frame calls on one GPU.  The result
 * lives in trt_multi_frames_context(m) (a whole-frame context on the first device). */
int trt_multi_render_frames(trt_multi *m, const trt_uniforms *u, uint32_t n, int mode, uint8_t *out, size_t cap, size_t *len);
trt_ctx *trt_multi_frames_context(trt_multi *m);
/* the two halves of a frame batch on one context: trace n frames into the staging buffer (no blend; means of frame i at
 * trt_frame_means + i * stride), and blend n means lying in this device's memory into the accumulation, in order */
int trt_trace_frames_async(trt_ctx *ctx, const trt_uniforms *u, uint32_t n);
int trt_frame_means(trt_ctx *ctx, void **dptr, size_t *frame_stride_bytes);
int trt_blend_frames_async(trt_ctx *ctx, const void *d_means, const trt_uniforms *u, uint32_t n);

/* ---- host logic of the reference, restated natively (no GPU needed) --------------------------- */

/* src/camera.rs:4-26 */
typedef struct trt_camera { trt_vec3 position; float yaw, pitch; } trt_camera;
void trt_camera_default(trt_camera *cam);                               /* lib.rs:118 */
void trt_camera_vectors(const trt_camera *cam, trt_vec3 *forward, trt_vec3 *right, trt_vec3 *up);
/* lib.rs:392-405; key: 'w' 'a' 's' 'd' or TRT_KEY_*; returns 1 if the camera moved */
#define TRT_KEY_UP 1001
#define TRT_KEY_DOWN 1002
#define TRT_KEY_LEFT 1003
#define TRT_KEY_RIGHT 1004
int trt_camera_key(trt_camera *cam, int key);

/* src/lib.rs:52-98: the scene file */
typedef struct trt_scene trt_scene;
typedef struct trt_scene_info {
    uint32_t width, height, samples_per_pixel, max_depth, frames_to_accumulate;
    float fov_degrees, char_aspect_ratio;
    uint32_t n_spheres, n_planes, n_triangles;
} trt_scene_info;
int trt_scene_parse(const char *json_text, size_t json_len, trt_scene **out, char *err, size_t err_cap);
int trt_scene_load(const char *path_or_null, trt_scene **out, char *err, size_t err_cap); /* NULL: embedded Cornell box, lib.rs:107 */
void trt_scene_free(trt_scene *s);
int trt_scene_get_info(const trt_scene *s, trt_scene_info *info);
/* lib.rs:114-115 clamp to the terminal (cols, rows-2); only applied by trt_run on a TTY */
void trt_scene_clamp(trt_scene *s, uint32_t term_cols, uint32_t term_rows);
void trt_scene_set_size(trt_scene *s, uint32_t width, uint32_t height);
/* lib.rs:120-154: spheres, planes, triangles in file order -> 180-byte primitives */
const trt_primitive *trt_scene_primitives(const trt_scene *s, uint32_t *count);
/* lib.rs:417-439 */
void trt_fill_uniforms(const trt_scene *s, const trt_camera *cam, uint32_t seed, uint32_t frame_number,
                       trt_uniforms *out);

/* the reference's public entry, `pub async fn run(full_color, verbose, scene_path)` lib.rs:100.
 * `opts` (may be NULL) carries the non-reference extensions needed without a TTY. */
typedef struct trt_run_options {
    uint32_t width, height;       /* 0: from the scene (clamped to the TTY when there is one) */
    uint32_t frames;              /* 0: frames_to_accumulate */
    uint32_t seed;                /* with has_seed: base seed, frame f uses seed + 0x9E3779B9*f */
    uint32_t has_seed;
    const char *keys;             /* comma-separated scripted keys, one per frame, or NULL */
    const char *out_path;         /* write the frame stream here instead of stdout, or NULL */
    int32_t device;
    uint32_t conv;
    uint32_t quiet;               /* suppress the status line */
    uint32_t batch_frames;        /* > 1: while the camera is still, render this many frames per GPU round trip
                                   * (trt_render_frames) and show the last one; 0 / 1: one frame per round trip as
                                   * the reference does */
    uint32_t gpus;                /* > 1: every frame row-sharded over devices device .. device + gpus - 1 (trt_multi) */
    uint32_t accel;               /* TRT_ACCEL_*: opt-in acceleration structure (conv 2) */
    uint32_t pipeline;            /* 1: frame f is copied back and written while frame f + 1 is traced
                                   * (trt_frame_submit / trt_frame_collect); same frames, same bytes, same order */
} trt_run_options;
int trt_run(int full_color, int verbose, const char *scene_path_or_null, const trt_run_options *opts);

const char *trt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TRT_B200_H */
