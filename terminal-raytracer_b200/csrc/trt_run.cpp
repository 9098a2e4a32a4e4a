// trt_run.cpp -- the reference's public entry `run(full_color, verbose, scene_path)`
// (/root/reference/src/lib.rs:100-571) over the C ABI: scene load + terminal clamp (lib.rs:102-115),
// camera (lib.rs:118), the frame loop with key handling and accumulation reset (lib.rs:386-565),
// frame framing ESC[1;1H + rows + ESC[h+1;1H + status line (lib.rs:497,544,551-558).
// The wgpu calls inside that loop are replaced by trt_upload_scene / trt_render_frame.
// Extensions that the reference does not have (needed without a TTY) live in trt_run_options.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <random>
#include <string>
#include <thread>
#include <vector>

#include <poll.h>
#include <sys/ioctl.h>
#include <termios.h>
#include <unistd.h>

#include "../../include/trt_b200.h"

namespace {

using clk = std::chrono::steady_clock;
double ms_since(clk::time_point t0) { return std::chrono::duration<double, std::milli>(clk::now() - t0).count(); }

int parse_key(const std::string &k) {
    if (k == "w" || k == "a" || k == "s" || k == "d") return k[0];
    if (k == "Up") return TRT_KEY_UP;
    if (k == "Down") return TRT_KEY_DOWN;
    if (k == "Left") return TRT_KEY_LEFT;
    if (k == "Right") return TRT_KEY_RIGHT;
    if (k == "Esc") return 27;
    return 0;
}

struct RawTty {                       // crossterm enable_raw_mode / cursor::Hide (lib.rs:355-356,568-569)
    termios saved{};
    bool active = false;
    void enter() {
        if (tcgetattr(STDIN_FILENO, &saved) != 0) return;
        termios raw = saved;
        cfmakeraw(&raw);
        if (tcsetattr(STDIN_FILENO, TCSANOW, &raw) == 0) active = true;
    }
    void leave() { if (active) tcsetattr(STDIN_FILENO, TCSANOW, &saved); active = false; }
    ~RawTty() { leave(); }
};

// one key event within 1 ms (lib.rs:390-391); returns 0 if none
int poll_key() {
    pollfd pf{STDIN_FILENO, POLLIN, 0};
    if (poll(&pf, 1, 1) <= 0) return 0;
    unsigned char b[8];
    ssize_t n = read(STDIN_FILENO, b, sizeof b);
    if (n <= 0) return 0;
    if (b[0] == 27) {
        if (n >= 3 && b[1] == '[') {
            switch (b[2]) { case 'A': return TRT_KEY_UP; case 'B': return TRT_KEY_DOWN; case 'C': return TRT_KEY_RIGHT; case 'D': return TRT_KEY_LEFT; }
            return -1;
        }
        return 27;
    }
    return b[0];
}

}  // namespace

extern "C" int trt_run(int full_color, int verbose, const char *scene_path, const trt_run_options *opts_in) {
    trt_run_options opts{};
    if (opts_in) opts = *opts_in;
    char err[512] = {0};
    trt_scene *scene = nullptr;
    int rc = trt_scene_load(scene_path, &scene, err, sizeof err);
    if (rc != TRT_OK) { fprintf(stderr, "%s\n", err); return rc; }

    FILE *sink = stdout;
    if (opts.out_path) {
        sink = fopen(opts.out_path, "wb");
        if (!sink) { fprintf(stderr, "cannot open '%s'\n", opts.out_path); trt_scene_free(scene); return TRT_ERR_IO; }
    }
    const bool tty = !opts.out_path && isatty(STDOUT_FILENO) && isatty(STDIN_FILENO);
    if (tty) {                                                   // lib.rs:102,114-115
        winsize ws{};
        if (ioctl(STDOUT_FILENO, TIOCGWINSZ, &ws) == 0 && ws.ws_col && ws.ws_row) trt_scene_clamp(scene, ws.ws_col, ws.ws_row);
    }
    trt_scene_info info{};
    trt_scene_get_info(scene, &info);
    if (opts.width && opts.height) { trt_scene_set_size(scene, opts.width, opts.height); trt_scene_get_info(scene, &info); }
    if (info.width == 0 || info.height == 0) { fprintf(stderr, "scene has an empty frame\n"); trt_scene_free(scene); return TRT_ERR_INVALID; }

    trt_camera cam;
    trt_camera_default(&cam);                                    // lib.rs:118
    uint32_t n_prims = 0;
    const trt_primitive *prims = trt_scene_primitives(scene, &n_prims);

    trt_config cfg{};
    cfg.abi_version = TRT_ABI_VERSION;
    cfg.device = opts.device;
    cfg.max_width = info.width;
    cfg.max_height = info.height;
    cfg.conv = opts.conv;
    cfg.accel = opts.conv == TRT_CONV_NATIVE ? opts.accel : TRT_ACCEL_NONE;
    trt_ctx *ctx = nullptr;
    trt_multi *multi = nullptr;                                  // --gpus N: every frame row-sharded over N devices
    if (opts.gpus > 1u) {
        std::vector<int32_t> devs(opts.gpus);
        // TRT_B200_MULTI_SAME_DEVICE=1 (tests on a one-GPU box): all shards on opts.device
        const char *same = getenv("TRT_B200_MULTI_SAME_DEVICE");
        for (uint32_t i = 0; i < opts.gpus; i++) devs[i] = opts.device + (same && *same == '1' ? 0 : (int32_t)i);
        rc = trt_multi_create(devs.data(), opts.gpus, &cfg, &multi);
        if (rc != TRT_OK) { fprintf(stderr, "%s\n", trt_multi_last_error(nullptr)); trt_scene_free(scene); if (sink != stdout) fclose(sink); return rc; }
        rc = trt_multi_upload_scene(multi, prims, n_prims);
        if (rc != TRT_OK) { fprintf(stderr, "%s\n", trt_multi_last_error(multi)); trt_multi_destroy(multi); trt_scene_free(scene); if (sink != stdout) fclose(sink); return rc; }
    } else {
        rc = trt_create(&cfg, &ctx);
        if (rc != TRT_OK) { fprintf(stderr, "%s\n", trt_last_error(nullptr)); trt_scene_free(scene); if (sink != stdout) fclose(sink); return rc; }
        rc = trt_upload_scene(ctx, prims, n_prims);
        if (rc != TRT_OK) { fprintf(stderr, "%s\n", trt_last_error(ctx)); trt_destroy(ctx); trt_scene_free(scene); if (sink != stdout) fclose(sink); return rc; }
    }
    if (verbose) fprintf(stderr, "%s: %u primitives, %ux%u, %u spp, depth %u, %u GPU(s)\n", trt_version(), n_prims, info.width, info.height,
                         info.samples_per_pixel, info.max_depth, multi ? opts.gpus : 1u);

    const size_t cap = (size_t)info.width * info.height * 26 + (size_t)info.height * 2 + 64;
    uint8_t *bytes = nullptr, *bytes2 = nullptr;
    if (trt_alloc_host(cap, (void **)&bytes) != TRT_OK) { if (ctx) trt_destroy(ctx); trt_multi_destroy(multi); trt_scene_free(scene); if (sink != stdout) fclose(sink); return TRT_ERR_NOMEM; }
    // pipelined loop (one GPU, one frame per round trip): frame f is collected, written and flushed while frame f + 1
    // is being traced; the displayed stream is the serial loop's, one frame later
    const bool pipelined = opts.pipeline && !multi && opts.batch_frames <= 1u;
    if (pipelined && trt_alloc_host(cap, (void **)&bytes2) != TRT_OK) bytes2 = nullptr;
    struct Pending { bool live = false; uint32_t frame_count = 0; clk::time_point t_submit; } pending;

    std::vector<int> script;
    if (opts.keys) {
        std::string all(opts.keys), cur;
        for (char c : all + ",") { if (c == ',') { if (!cur.empty()) script.push_back(parse_key(cur)); cur.clear(); } else if (c != ' ') cur += c; }
    }
    RawTty raw;
    if (tty) { raw.enter(); fputs("\x1b[?25l\x1b[2J\x1b[1;1H", sink); fflush(sink); }   // lib.rs:355-368

    std::random_device rd;
    std::vector<trt_uniforms> batch_u;
    uint32_t frame_count = 0, rendered = 0;
    bool camera_moved = true;
    std::deque<double> frame_times;                              // lib.rs:362-364
    auto last_frame = clk::now();
    const uint32_t frame_budget = opts.frames ? opts.frames : (tty ? 0xFFFFFFFFu : info.frames_to_accumulate);
    const int mode = full_color ? TRT_MODE_TRUECOLOR : TRT_MODE_ASCII;
    size_t script_pos = 0;
    int status = TRT_OK;

    // the frame still in flight of a pipelined loop (at exit, or when the loop goes idle)
    auto flush_pending = [&]() {
        if (!pending.live || status != TRT_OK) return;
        pending.live = false;
        size_t len = 0;
        uint8_t *buf = (rendered & 1u) ? bytes2 : bytes;
        status = trt_frame_collect(ctx, buf, cap, &len);
        if (status != TRT_OK) { fprintf(stderr, "%s\n", trt_last_error(ctx)); return; }
        fputs("\x1b[1;1H", sink);
        fwrite(buf, 1, len, sink);
        if (!opts.quiet) fprintf(sink, "\x1b[%u;1HFrame: %u/%u | FPS: %.1f | GPU: %dms | CPU: %dms | Cleanup: %dms | Other: %dms | Total: %dms\r\n",
                                 info.height + 1u, pending.frame_count, info.frames_to_accumulate, 0.0, 0, 0, 0, 0, 0);
        fflush(sink);
    };

    for (;;) {
        auto t_total = clk::now();
        bool moved = false;
        int key = 0;
        if (!script.empty()) { if (script_pos < script.size()) key = script[script_pos++]; }
        else if (tty) key = poll_key();
        if (key == 27) break;                                    // lib.rs:402
        if (key > 0) moved = trt_camera_key(&cam, key) != 0;      // lib.rs:392-405
        if (moved) { camera_moved = true; frame_count = 0; }     // lib.rs:409-412
        if (rendered >= frame_budget) break;

        if (frame_count < info.frames_to_accumulate) {           // lib.rs:415
            // frames of this GPU round trip: 1 as in the reference, or a batch while the camera is still (extension)
            uint32_t k = opts.batch_frames > 1u ? opts.batch_frames : 1u;
            if (k > info.frames_to_accumulate - frame_count) k = info.frames_to_accumulate - frame_count;
            if (k > frame_budget - rendered) k = frame_budget - rendered;
            batch_u.resize(k);
            for (uint32_t j = 0; j < k; j++) {
                const uint32_t seed = opts.has_seed ? opts.seed + 0x9E3779B9u * (rendered + j) : (uint32_t)rd() + frame_count + j;   // lib.rs:423
                trt_fill_uniforms(scene, &cam, seed, camera_moved && j == 0 ? 0u : frame_count + j, &batch_u[j]);                  // lib.rs:424
            }
            size_t len = 0;
            auto t_gpu = clk::now();
            // presents one finished frame: cursor home + rows (lib.rs:497,544) + status line (lib.rs:551-558)
            auto present = [&](const uint8_t *data, size_t n, uint32_t shown_count, double gpu_ms) {
                auto now = clk::now();
                frame_times.push_back(std::chrono::duration<double>(now - last_frame).count());
                last_frame = now;
                if (frame_times.size() > 30) frame_times.pop_front();
                double sum = 0.0;
                for (double t : frame_times) sum += t;
                const double fps = sum > 0.0 ? (double)frame_times.size() / sum : 0.0;
                auto t_io = clk::now();
                fputs("\x1b[1;1H", sink);                        // lib.rs:497
                fwrite(data, 1, n, sink);                        // lib.rs:544
                fflush(sink);
                const double io_ms = ms_since(t_io);
                const double total_ms = ms_since(t_total);
                if (!opts.quiet) {                               // lib.rs:551-558
                    double other = total_ms - gpu_ms - io_ms;
                    if (other < 0.0) other = 0.0;
                    fprintf(sink, "\x1b[%u;1HFrame: %u/%u | FPS: %.1f | GPU: %dms | CPU: %dms | Cleanup: %dms | Other: %dms | Total: %dms\r\n",
                            info.height + 1u, shown_count, info.frames_to_accumulate, fps, (int)gpu_ms, 0, 0, (int)other, (int)total_ms);
                    fflush(sink);
                }
            };
            if (pipelined && bytes2) {
                status = trt_frame_submit(ctx, &batch_u[0], mode);           // frame f: queued, not waited for
                if (status != TRT_OK) { fprintf(stderr, "%s\n", trt_last_error(ctx)); break; }
                frame_count += 1;                                // lib.rs:457
                rendered += 1;
                if (pending.live) {                              // frame f - 1: copy + write overlap the trace of frame f
                    uint8_t *buf = (rendered & 1u) ? bytes : bytes2;
                    status = trt_frame_collect(ctx, buf, cap, &len);
                    if (status != TRT_OK) { fprintf(stderr, "%s\n", trt_last_error(ctx)); break; }
                    present(buf, len, pending.frame_count, std::chrono::duration<double, std::milli>(clk::now() - pending.t_submit).count());
                }
                pending.live = true; pending.frame_count = frame_count; pending.t_submit = t_gpu;
            } else {
                if (multi) status = trt_multi_render_frame(multi, &batch_u[0], mode, bytes, cap, &len);
                else status = k == 1u ? trt_render_frame(ctx, &batch_u[0], mode, bytes, cap, &len)
                                      : trt_render_frames(ctx, batch_u.data(), k, mode, 0, bytes, cap, &len, nullptr);
                if (status != TRT_OK) { fprintf(stderr, "%s\n", multi ? trt_multi_last_error(multi) : trt_last_error(ctx)); break; }
                const double gpu_ms = ms_since(t_gpu);
                frame_count += k;                                // lib.rs:457
                rendered += k;
                present(bytes, len, frame_count, gpu_ms);
            }
            camera_moved = false;                                // lib.rs:560
        } else {
            flush_pending();
            if (status != TRT_OK) break;
            if (!tty && script_pos >= script.size()) break;      // headless: nothing more can happen
            std::this_thread::sleep_for(std::chrono::milliseconds(10));   // lib.rs:563
        }
    }
    flush_pending();                                             // the last frame of a pipelined loop
    if (tty) { fputs("\x1b[?25h", sink); fflush(sink); }         // lib.rs:568
    raw.leave();
    if (tty || verbose) fprintf(stderr, "Exiting.\n");           // lib.rs:570
    trt_free_host(bytes);
    if (bytes2) trt_free_host(bytes2);
    if (ctx) trt_destroy(ctx);
    trt_multi_destroy(multi);
    trt_scene_free(scene);
    if (sink != stdout) fclose(sink);
    return status;
}
