// trt_main.cpp -- command line of the drop-in binary.  Same flags as the reference's
// /root/reference/src/main.rs:6-39 (--full-color, --verbose, --path <file>, --threads N; unknown
// flags are ignored, main.rs:9-20) plus clearly separate extensions for running without a TTY.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>

#include "../../include/trt_b200.h"

int main(int argc, char **argv) {
    bool full_color = false, verbose = false;
    const char *path = nullptr;
    long threads = -1;
    trt_run_options o{};
    o.conv = TRT_CONV_NATIVE;          // the production arithmetic; --strict / --fused select the bit-exact builds
    for (int i = 1; i < argc; i++) {
        const std::string a = argv[i];
        const char *next = i + 1 < argc ? argv[i + 1] : nullptr;
        if (a == "--full-color") full_color = true;
        else if (a == "--verbose") verbose = true;
        else if (a == "--path" && next) path = next;                       // main.rs:11-13
        else if (a == "--threads" && next) threads = strtol(next, nullptr, 10);   // main.rs:15-20
        // ---- extensions (not in the reference) ----
        else if (a == "--size" && next) { unsigned w = 0, h = 0; if (sscanf(next, "%ux%u", &w, &h) == 2) { o.width = w; o.height = h; } }
        else if (a == "--frames" && next) o.frames = (uint32_t)strtoul(next, nullptr, 10);
        else if (a == "--seed" && next) { o.seed = (uint32_t)strtoul(next, nullptr, 0); o.has_seed = 1; }
        else if (a == "--keys" && next) o.keys = next;
        else if (a == "--out" && next) o.out_path = next;
        else if (a == "--device" && next) o.device = (int32_t)strtol(next, nullptr, 10);
        else if (a == "--fused") o.conv = TRT_CONV_FUSED;
        else if (a == "--strict") o.conv = TRT_CONV_AS_WRITTEN;
        else if (a == "--native") o.conv = TRT_CONV_NATIVE;
        else if (a == "--batch" && next) o.batch_frames = (uint32_t)strtoul(next, nullptr, 10);
        else if (a == "--quiet") o.quiet = 1;
        else if (a == "--gpus" && next) o.gpus = (uint32_t)strtoul(next, nullptr, 10);     // rows of every frame over N devices
        else if (a == "--accel" && next) o.accel = !strcmp(next, "grid") ? TRT_ACCEL_GRID : TRT_ACCEL_NONE;   // opt-in, outside the parity contract
        else if (a == "--pipeline") o.pipeline = 1;                                        // overlap copy + stdout with the next frame's trace
    }
    if (threads <= 0) {                                                    // main.rs:18-19
        threads = (long)std::thread::hardware_concurrency();
        if (threads <= 0) threads = 4;
    }
    // main.rs:27-35 banners.  --threads sized the reference's rayon pool for the CPU row formatter
    // (lib.rs:503); that stage now runs on the GPU, so the value is accepted and has no effect.
    fprintf(stderr, full_color ? "outputting with \xE2\x96\x88 characters\n" : "outputting with ASCII characters\n");
    if (verbose) fprintf(stderr, "rayon threads: %ld (unused: rows are formatted on the GPU)\n", threads);
    return trt_run(full_color ? 1 : 0, verbose ? 1 : 0, path, &o) == TRT_OK ? 0 : 1;
}
