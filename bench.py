#!/usr/bin/env python
"""
bench.py -- Msamples/s (path samples per second) of the per-pixel path-tracing hot path on B200.

  python bench.py --gpus N --steps K --warmup W                 (our arm; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W (reference arm: CPU restatement)

A "step" is one frame of the workload through the hot path.  Default workload = BASELINE.json
configs[2] "c3": Cornell box 3840x2160, 1024 spp, max_depth 32, truecolor stream, rows sharded over
the N GPUs (the configuration the metric "Cornell Box at 1/2/4/8 B200" is quoted on; total work is
fixed as N grows => "scaling": "strong").  --workload c1|c2|c4 select the other configs.

`value`  : device-resident (scene + buffers in HBM; trace + accumulate + quantise of the own rows + row-length all-gather).
`e2e`    : the same through trt_render_frame with HOST buffers: uniforms in, byte stream out (pinned).
`roofline`: trace kernel, FP32: achieved = F_min (algorithmic flops) / CUDA-event time.
`roofline_quantise`: the on-device quantiser (one pass: quant_fused_kernel), algorithmic bytes / its CUDA-event time / HBM peak.
`cpu_baseline`: the reference's shader compiled for the host (oracle/_ref: shader.wgsl transliterated to C, kind
            "reference"; the hand-written port, kind "port", where _ref did not travel) on a bounded strided pixel
            sample of the same frame, all host threads.
`config.parity`: the named configuration itself checked against that CPU build on a strided pixel subset: the strict
            (conv 0) frame bit for bit, the benchmarked conv 2 frame by north_star's tolerances.
The product arm is fed by the product's own host code (Scene.parse / trt_fill_uniforms); only the parity,
cpu_baseline and --impl reference legs touch oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import __graft_entry__ as graft  # noqa: E402

NOMINAL_FP32_TFLOPS = 148 * 128 * 2 * 1.965e9 / 1e12      # 74.45: SMs x lanes x 2 x max clock
QUANT_KERNELS = "quant_fused_kernel (one pass, decoupled look-back) + quant_row_len_from_starts_kernel + the memset of the tile states"


def workload(name, O, args):
    """-> dict(scene, width, height, spp, depth, mode, frames_accumulate, seed_fn, label)"""
    if name in ("c1", "c2", "c3"):
        sc = O.CORNELL
        if name == "c1":
            w, h, spp, mode, acc = 400, 200, 16, 0, False
        elif name == "c2":
            w, h, spp, mode, acc = 400, 200, 16, 1, True
        else:
            w, h, spp, mode, acc = 3840, 2160, 1024, 1, False
        depth = 32
        seed_fn = (lambda f: O.c2_seed(f)) if acc else (lambda f: O.C1_SEED)
        label = {"c1": "C1 Cornell_Box 400x200 16spp depth32 ASCII single frame",
                 "c2": "C2 Cornell_Box 400x200 16spp depth32 truecolor, temporal accumulation (frame_number = step)",
                 "c3": "C3 Cornell_Box 3840x2160 1024spp depth32 truecolor, rows sharded over the GPUs"}[name]
    elif name == "c4":
        sc = O.make_c4_scene()
        w, h, spp, depth, mode, acc = 1920, 1080, 256, 32, 0, False
        seed_fn = lambda f: O.C4_SEED  # noqa: E731
        label = "C4 synthetic 4096 spheres + 16384 triangles + 2 emissive quads 1920x1080 256spp (linear scan)"
    else:
        raise SystemExit(f"unknown workload {name}")
    if args.width:
        w = args.width
    if args.height:
        h = args.height
    if args.spp:
        spp = args.spp
    if args.width or args.height or args.spp:
        label += f" [OVERRIDE {w}x{h} {spp}spp: not the named config]"
    return dict(scene=sc, w=w, h=h, spp=spp, depth=depth, mode=mode, acc=acc, seed_fn=seed_fn, label=label, name=name)


def sample_indices(w, h, n):
    """n pixel indices on a regular stride over the frame (deterministic)."""
    total = w * h
    n = max(1, min(n, total))
    return np.unique((np.arange(n, dtype=np.float64) * (total / n)).astype(np.uint32))


def host_threads():
    """Threads the CPU legs use: every core this process may run on.  Set explicitly -- torchrun exports
    OMP_NUM_THREADS=1, which omp_get_max_threads() would obey."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def cpu_kind(O):
    return "reference" if O.ref_available() else "port"


def cpu_build(O):
    if O.ref_available():
        return ("oracle/_ref/libref.so: /root/reference/src/shader.wgsl transliterated to C (oracle/wgsl2c.py), "
                "gcc -O2 -ffp-contract=off, OpenMP over pixels")
    return "oracle/liboracle_fast.so: hand-written port, gcc -O3 -ffp-contract=off, OpenMP over pixels"


def cpu_run(O, wl, prims, n_pixels, threads, frame=0):
    """One timed pass of the CPU build over a strided pixel sample: oracle/_ref (the reference's shader.wgsl
    transliterated to C) when it is there, else the hand-written port (-O3 build; same bits)."""
    idx = sample_indices(wl["w"], wl["h"], n_pixels)
    u = O.make_uniforms(wl["scene"], len(prims), wl["seed_fn"](frame), 0, width=wl["w"], height=wl["h"], spp=wl["spp"],
                        max_depth=wl["depth"])
    if O.ref_available():
        t0 = time.perf_counter()
        r = O.ref_render(prims, u, idx=idx, threads=threads)
        dt = time.perf_counter() - t0
        return O.samples_from_variance(r["variance"], wl["spp"]), dt, len(idx)
    t0 = time.perf_counter()
    r = O.render(prims, u, idx=idx, threads=threads, fast=True)
    dt = time.perf_counter() - t0
    return r["samples_traced"], dt, len(idx)


def cpu_sample_size(O, wl, prims, threads, budget_s):
    """pick a pixel count worth ~budget_s seconds of CPU work from a short probe"""
    O.render(prims, O.make_uniforms(wl["scene"], len(prims), 1, 0, width=32, height=16, spp=4, max_depth=4))   # spin up
    probe = max(64, min(4096, wl["w"] * wl["h"] // 64))
    s, dt, n = cpu_run(O, wl, prims, probe, threads)
    while dt < 0.3 and probe < wl["w"] * wl["h"]:
        probe = min(wl["w"] * wl["h"], probe * 4)
        s, dt, n = cpu_run(O, wl, prims, probe, threads)
    return int(max(64, min(wl["w"] * wl["h"], n * budget_s / max(dt, 1e-6))))


def reference_arm(args):
    """The reference's own implementation of the path is a WGSL shader behind wgpu (needs a graphics adapter:
    unavailable on this box).  The reference arm runs that shader's text compiled for the host (oracle/_ref, built
    by oracle/wgsl2c.py from /root/reference/src/shader.wgsl; kind "reference") on all host threads, on a bounded
    strided sample of the same frame; without _ref it falls back to the hand-written port (kind "port")."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    O = graft.load_oracle()
    wl = workload(args.workload, O, args)
    prims = O.flatten_scene(wl["scene"])
    threads = host_threads()
    budget = 60.0 / max(1, args.steps + args.warmup)
    n_px = cpu_sample_size(O, wl, prims, threads, min(20.0, max(2.0, budget)))
    for _ in range(args.warmup):
        cpu_run(O, wl, prims, n_px, threads)
    samples, secs = 0, 0.0
    for k in range(args.steps):
        s, dt, n = cpu_run(O, wl, prims, n_px, threads, frame=k if wl["acc"] else 0)
        samples += s
        secs += dt
    val = samples / secs / 1e6
    sample = f"{n} of {wl['w'] * wl['h']} pixels on a regular stride, every sample and bounce of those pixels"
    line = {"impl": "reference", "metric": "Msamples/s", "value": val, "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["label"], "l2": "n/a (CPU)"},
            "cpu_baseline": {"value": val, "unit": "Msamples/s", "cores": threads, "kind": cpu_kind(O), "sample": sample,
                             "build": cpu_build(O)},
            "e2e": {"value": val, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def parity_block(pkg, O, wl, prims_np, scene, cam, local, conv, n_pixels, native_renderer):
    """The named configuration itself against the CPU build of the reference's shader, on a strided pixel subset:
    strict (conv 0) frame bit for bit; the benchmarked convention by north_star's bars (pixels over 1e-4 relative,
    RMSE against the Monte-Carlo noise floor)."""
    w, h = wl["w"], wl["h"]
    idx = sample_indices(w, h, n_pixels)
    seed = wl["seed_fn"](0)
    u_np = O.make_uniforms(wl["scene"], len(prims_np), seed, 0, width=w, height=h, spp=wl["spp"], max_depth=wl["depth"])
    u = scene.uniforms(cam, seed, 0)
    same_uniforms = bytes(u)[:112] == u_np.tobytes()[:112]          # product host code == oracle restatement
    use_ref = O.ref_available()
    t0 = time.perf_counter()
    ref = O.ref_render(prims_np, u_np, idx=idx) if use_ref else O.render(prims_np, u_np, idx=idx)
    cpu_s = time.perf_counter() - t0
    want = ref["pixels"]
    out = {"checker": "oracle/_ref (shader.wgsl transliterated to C)" if use_ref else "oracle (hand-written port)",
           "subset_pixels": int(len(idx)), "uniforms_equal_oracle": bool(same_uniforms), "cpu_seconds": round(cpu_s, 2)}
    # strict build: bits
    rs = pkg.Renderer(w, h, device=local, conv=0)
    rs.upload_scene(scene)
    rs.render_frame(u, pkg.MODE_NONE)
    got = rs.read_pixels()[idx]
    var = rs.read_variance()[idx]
    rs.close()
    out["strict_bit_mismatches"] = int((got.view(np.uint32) != want.view(np.uint32)).any(axis=1).sum())
    out["strict_variance_bit_mismatches"] = int((var["variance"].view(np.uint32) != ref["variance"]["variance"].view(np.uint32)).sum())
    # the benchmarked convention: tolerance
    if conv != 0:
        native_renderer.render_frame(u, pkg.MODE_NONE)
        nat = native_renderer.read_pixels()[idx][:, :3].astype(np.float64)
        wt = want[:, :3].astype(np.float64)
        rel = np.abs(nat - wt) / np.maximum(np.abs(wt), 1e-2)
        out["native_over_1e-4"] = int((rel.max(axis=1) > 1e-4).sum())
        u2 = O.make_uniforms(wl["scene"], len(prims_np), (seed + 0x9E3779B9) & 0xFFFFFFFF, 0, width=w, height=h, spp=wl["spp"],
                             max_depth=wl["depth"])
        other = (O.ref_render(prims_np, u2, idx=idx) if use_ref else O.render(prims_np, u2, idx=idx))["pixels"][:, :3].astype(np.float64)
        noise = float(np.sqrt(((other - wt) ** 2).mean()))
        rmse = float(np.sqrt(((nat - wt) ** 2).mean()))
        out["native_rmse"] = rmse
        out["noise_rmse_seed_to_seed"] = noise
        out["native_rmse_vs_noise"] = rmse / noise if noise > 0 else None
        out["native_mean_rel_diff"] = float(abs(nat.mean() - wt.mean()) / max(wt.mean(), 1e-12))
        if use_ref:
            try:
                fma = O.ref_render(prims_np, u_np, idx=idx, fma=True)["pixels"][:, :3].astype(np.float64)
                relf = np.abs(fma - wt) / np.maximum(np.abs(wt), 1e-2)
                out["yardstick_ref_contracted_over_1e-4"] = int((relf.max(axis=1) > 1e-4).sum())
                out["yardstick_ref_contracted_rmse_vs_noise"] = float(np.sqrt(((fma - wt) ** 2).mean())) / noise if noise > 0 else None
            except Exception:
                pass
        out["note"] = ("strict = conv 0 kernels, compared bit for bit; native = the benchmarked conv 2 kernel by north_star's bars; "
                       "yardstick = the reference's own text compiled with FMA contraction allowed (another legal rounding of it) "
                       "against its strict compilation, same pixels")
    return out


class ClockSampler:
    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.p = None

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200",
                                       "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out, _ = self.p.communicate(timeout=5)
        except subprocess.TimeoutExpired:
            self.p.kill()
            out, _ = self.p.communicate()
        sm, mx, reasons, power = [], [], set(), []
        for ln in out.splitlines():
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": statistics.median(sm), "sm_max_mhz": max(mx), "reasons": sorted(reasons),
                "power_w_max": max(power), "samples": len(sm)}


def ours(args):
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
        args.gpus = world
    pkg = graft.load_package()
    pkg.api.lib()                                     # fails loudly if the CUDA library is missing
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from terminal_raytracer_b200 import multigpu

    OO = graft.load_oracle()                           # module import only: the workload DEFINITIONS (scene data, seeds)
    O = OO if rank == 0 else None                      # the CPU build itself: parity / cpu_baseline legs on rank 0
    wl = workload(args.workload, OO, args)
    w, h, mode = wl["w"], wl["h"], wl["mode"]
    # the product's own host code: JSON -> SceneConfig -> primitives (trt_scene_parse), camera (trt_camera_default),
    # per-frame uniforms (trt_fill_uniforms) -- lib.rs:52-154, :118, :417-439
    scene = pkg.Scene.parse(json.dumps(dict(wl["scene"], width=w, height=h, samples_per_pixel=wl["spp"], max_depth=wl["depth"])))
    cam = pkg.Camera()

    def uniforms(f):
        return scene.uniforms(cam, wl["seed_fn"](f), f if wl["acc"] else 0)

    sr = multigpu.ShardedRenderer(pkg, w, h, rank, world, stripe_rows=args.stripe_rows, device=local, conv=args.conv)
    sr.upload_scene(scene)
    stream = sr.stream
    frame_bytes = w * h * 16
    small = frame_bytes * 3 < 126 * 2**20
    flush = torch.empty(256 * 2**20, dtype=torch.uint8, device=f"cuda:{local}") if small else None

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident timing ------------------------------------------------------------------
    for k in range(args.warmup):
        sr.step(uniforms(k), mode)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    trace_ms, combine_ms, quant_ms, samples_local, launches = [], [], [], 0, 0
    barrier()
    for k in range(args.steps):
        if flush is not None:
            with torch.cuda.stream(stream):
                flush.fill_(k & 0xFF)
        evs[k][0].record(stream)
        sr.step(uniforms(args.warmup + k), mode)
        evs[k][1].record(stream)
        # per-step statistics are read back after the step's stop event (they sync the stream)
        t = sr.r.timings()
        st = sr.step_times()                          # this rank's trace kernel alone / row-length exchange / quantise of its rows
        trace_ms.append(st[0])
        combine_ms.append(st[1] + st[2])
        quant_ms.append(st[2])
        launches += t.kernel_launches                 # order + trace + 3 quantise kernels + row lengths
        samples_local += sr.r.frame_stats().samples_traced
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = sum(a.elapsed_time(b) for a, b in evs)
    tt = torch.tensor([dev_ms, float(samples_local), float(sum(trace_ms))], dtype=torch.float64, device=f"cuda:{local}")
    if world > 1:
        mx = tt.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        sm = tt.clone()
        dist.all_reduce(sm, op=dist.ReduceOp.SUM)
        dev_ms, samples_total, trace_sum_max = mx[0].item(), sm[1].item(), mx[2].item()
        per_rank = [torch.zeros(2, dtype=torch.float64, device=f"cuda:{local}") for _ in range(world)]
        dist.all_gather(per_rank, torch.tensor([sum(trace_ms) / args.steps, sum(combine_ms) / args.steps], dtype=torch.float64,
                                               device=f"cuda:{local}"))
        ranks = {"trace_ms": [round(v[0].item(), 2) for v in per_rank],
                 "combine_ms_rank0": round(per_rank[0][1].item(), 2),
                 "note": "per rank: its trace kernel alone; rank 0: quantise of its rows + all-gather of the row lengths (waits for the slowest rank)"}
    else:
        samples_total, trace_sum_max = float(samples_local), float(sum(trace_ms))
        ranks = None

    # ---- end to end through trt_render_frame with host buffers (rank-local frame; N = 1 only is exact) ---
    e2e = None
    if world == 1:
        r1 = sr.r
        cap = w * h * 26 + 2 * h + 64
        host = pkg.api._P()
        assert pkg.api.lib().trt_alloc_host(cap, __import__("ctypes").byref(host)) == 0
        for k in range(max(1, args.warmup // 2)):
            r1.render_frame_into(uniforms(k), mode, host, cap)
        torch.cuda.synchronize()
        t_e2e, s_e2e, out_bytes = 0.0, 0, 0
        for k in range(args.steps):
            if flush is not None:
                flush.fill_(k & 0xFF)
                torch.cuda.synchronize()
            t0 = time.perf_counter()
            n = r1.render_frame_into(uniforms(args.warmup + k), mode, host, cap)    # blocks until bytes are on the host
            t_e2e += time.perf_counter() - t0
            s_e2e += r1.frame_stats().samples_traced
            out_bytes += n
        pkg.api.lib().trt_free_host(host)
        e2e = {"value": s_e2e / t_e2e / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 160,
               "d2h_bytes_per_step": out_bytes // args.steps + 8, "ms_per_step": t_e2e / args.steps * 1e3}
    else:
        # N > 1: every rank quantises its rows and copies them over its own PCIe link to their place in the host buffer
        # the ranks share; the frame is on the host when the slowest rank's copy has landed (barrier)
        barrier()
        t0 = time.perf_counter()
        out_bytes = 0
        for k in range(args.steps):
            sr.step(uniforms(args.warmup + k), mode)
            out_bytes += sr.read_stream()                          # blocks until this rank's rows are in host memory
            barrier()
        t_e2e = time.perf_counter() - t0
        e2e = {"value": samples_total / t_e2e / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": 160 * world,
               "d2h_bytes_per_step": out_bytes // args.steps + 4 * sr.max_rows * world * world, "ms_per_step": t_e2e / args.steps * 1e3,
               "note": "d2h = the frame's byte stream (each rank its own rows, its own link) + the all-gathered row-length table on every rank"}

    if rank != 0:
        sr.close()
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the trace kernel: F_min per sample from the instrumented kernel on a row subset ----
    sub = 1 if w * h * wl["spp"] <= 2_000_000 else max(1, args.count_stride)
    rc = pkg.Renderer(w, h, device=local, conv=args.conv, count_work=True, shard_index=0, shard_count=sub, stripe_rows=1)
    rc.upload_scene(scene)
    rc.render_frame(uniforms(args.warmup), pkg.MODE_NONE)
    cnt = rc.counters()
    rc.close()
    n_cnt = cnt["samples_base"] + cnt["samples_extra"]
    fmin_per_sample = cnt["flops_min"] / max(1, n_cnt)
    fwritten_per_sample = cnt["flops_written"] / max(1, n_cnt)
    trace_s = trace_sum_max / 1e3
    achieved = fmin_per_sample * samples_total / world / trace_s / 1e12 if trace_s > 0 else 0.0   # per GPU
    ffma, sms, khz = pkg.api.measure_fp32_peak(local)
    # DRAM bytes per launch of the trace kernel from the committed ncu --set full capture of this workload
    traffic = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        ent = tj.get(f"{wl['name']}_conv{args.conv}")
        if ent and (ent["width"], ent["height"], ent["spp"]) == (w, h, wl["spp"]) and world == 1:
            traffic = ent["dram_bytes_per_launch"]
    except (OSError, ValueError, KeyError):
        pass
    roofline = {"bound": "fp32", "kernel": sr.r.kernel_name(),
                "achieved": achieved, "peak": NOMINAL_FP32_TFLOPS,
                "unit": "TFLOP/s", "frac": achieved / NOMINAL_FP32_TFLOPS, "traffic": traffic,
                "traffic_unit": "bytes of DRAM read + written per launch (ncu --set full, profiles/traffic.json)",
                "peak_source": "nominal 148 SM x 128 lanes x 2 x 1.965 GHz (MEASURED_PEAKS.json holds no FP32 figure)",
                "measured_ffma_tflops": ffma, "frac_of_measured_ffma": achieved / ffma if ffma else None,
                "flops_min_per_sample": fmin_per_sample, "flops_written_per_sample": fwritten_per_sample,
                "flops_basis": f"instrumented kernel on rows y % {sub} == 0 ({n_cnt} samples), DESIGN.md cost table",
                "trace_ms_per_step": trace_s * 1e3 / args.steps, "per_gpu": True}

    # ---- quantiser (north_star: "achieved HBM GB/s for the accumulate/quantise pass"): K4 over the assembled frame;
    # algorithmic bytes = 16 B/px read + the exact stream length written -------------------------------------
    roofline_q = None
    if world == 1 and mode != pkg.MODE_NONE and quant_ms:
        stream_len = sr.r.quantise_async(mode, want_len=True)
        sr.r.sync()
        hbm, src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        try:
            hbm = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
            src = "MEASURED_PEAKS.json hbm_gbs"
        except (OSError, ValueError, KeyError):
            pass
        q_ms = statistics.median(quant_ms)
        q_bytes = w * h * 16 + stream_len
        gbs = q_bytes / (q_ms / 1e3) / 1e9 if q_ms > 0 else None
        roofline_q = {"bound": "hbm", "kernel": QUANT_KERNELS, "achieved": gbs, "peak": hbm, "unit": "GB/s",
                      "frac": gbs / hbm if gbs else None, "peak_source": src,
                      "algorithmic_bytes": q_bytes, "ms": q_ms, "stream_bytes": stream_len,
                      "note": "CUDA events around the quantiser's launches on the context's stream, median of the timed steps"}

    # ---- CPU baseline (oracle port, bounded sample) -----------------------------------------------------
    cpu = None
    prims_np = O.flatten_scene(wl["scene"])
    if world == 1 and not args.no_cpu:
        threads = host_threads()
        n_px = cpu_sample_size(O, wl, prims_np, threads, args.cpu_seconds)
        s, dt, n = cpu_run(O, wl, prims_np, n_px, threads)
        cpu = {"value": s / dt / 1e6, "unit": "Msamples/s", "cores": threads, "kind": cpu_kind(O), "build": cpu_build(O),
               "sample": f"{n} of {w * h} pixels on a regular stride ({s} samples, {dt:.1f} s)"}

    # ---- parity of the named configuration itself (strided subset against the CPU build of the reference's shader) ----
    parity = None
    if world == 1 and not args.no_parity:
        # (a C4 pixel costs the CPU ~10^4 times a Cornell pixel: 20 484 tests per scan, five scans per bounce)
        n_par = min(args.parity_pixels, 160) if wl["name"] == "c4" else args.parity_pixels
        parity = parity_block(pkg, O, wl, prims_np, scene, cam, local, args.conv, n_par, sr.r)

    # ---- the same frame in the strict convention (conv 0: bit-exact against the oracle), for comparison ----
    strict = None
    if world == 1 and args.conv != 0 and not args.no_strict:
        rs = pkg.Renderer(w, h, device=local, conv=0)
        rs.upload_scene(scene)
        rs.render_frame(uniforms(0), pkg.MODE_NONE)
        ms, smp = 0.0, 0
        for k in range(2):
            rs.render_frame(uniforms(args.warmup + k), pkg.MODE_NONE)
            ms += rs.timings().trace_ms
            smp += rs.frame_stats().samples_traced
        rs_name = rs.kernel_name()
        rs.close()
        strict = {"conv": 0, "value": smp / ms / 1e3, "unit": "Msamples/s (trace kernel, device)", "trace_ms_per_step": ms / 2,
                  "roofline_frac": fmin_per_sample * smp / (ms / 1e3) / 1e12 / NOMINAL_FP32_TFLOPS,
                  "kernel": rs_name,
                  "note": "one IEEE rounding per WGSL operator; bit-exact against oracle/_ref (config.parity, tests/test_gpu_parity.py)"}

    value = samples_total / (dev_ms / 1e3) / 1e6
    line = {"metric": "Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": wl["label"], "samples_traced_per_step": samples_total / args.steps,
                       "samples_nominal_per_step": w * h * wl["spp"], "conv": args.conv,
                       "arithmetic": {0: "strict: one IEEE rounding per WGSL operator (bit-exact vs oracle)",
                                      1: "strict + fused vector helpers (bit-exact vs oracle conv 1)",
                                      2: "device-native: FFMA contraction + MUFU rcp/rsqrt/sqrt/sin/cos, within north_star tolerances (tests/test_gpu_native.py)"}[args.conv],
                       "stripe_rows": args.stripe_rows,
                       "l2": "256 MiB flush between timed steps" if small else "frame buffers (3 x %d MB) exceed the 126 MB L2" % (frame_bytes >> 20),
                       "parallelism": f"rows x{world}", "parity": parity, "strict": strict, "ranks": ranks},
            "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roofline, "roofline_quantise": roofline_q,
            "cpu_baseline": cpu}
    print(json.dumps(line), flush=True)
    sr.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=["c1", "c2", "c3", "c4"])
    ap.add_argument("--width", type=int, default=0)
    ap.add_argument("--height", type=int, default=0)
    ap.add_argument("--spp", type=int, default=0)
    ap.add_argument("--conv", type=int, default=2, choices=[0, 1, 2],
                    help="arithmetic convention: 2 device-native (default, the production build), 0 strict IEEE as written, 1 fused helpers")
    ap.add_argument("--no-strict", action="store_true", help="skip the strict (conv 0) comparison leg")
    ap.add_argument("--stripe-rows", type=int, default=1,
                    help="rows per stripe of the round-robin row deal (1: every 8th row per rank at N = 8; tools/shard_sweep.py)")
    ap.add_argument("--count-stride", type=int, default=24, help="instrumented kernel runs on every n-th row")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity check of the named configuration (config.parity)")
    ap.add_argument("--parity-pixels", type=int, default=20000)
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        reference_arm(args)
    else:
        ours(args)


if __name__ == "__main__":
    main()
