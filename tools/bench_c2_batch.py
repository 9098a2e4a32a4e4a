"""BASELINE.json configs[1] (C2) through the frame batch: Cornell 400x200, 16 spp, depth 32, truecolor, 100 accumulated
frames (frame_number = f, seed_f = C1_SEED + 0x9E3779B9 f) as ONE trt_render_frames call, against the same frames one
trt_render_frame call at a time.  Wall clock around the calls: uniforms in -> byte stream(s) in host memory.
usage: python tools/bench_c2_batch.py [frames] [conv]"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
conv = int(sys.argv[2]) if len(sys.argv) > 2 else 2
w, h, spp, depth = 400, 200, 16, 32
prims = O.flatten_scene(O.CORNELL)
us = [pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), O.c2_seed(f), f, width=w, height=h, spp=spp, max_depth=depth).tobytes())
      for f in range(n)]
res = {"config": f"C2 Cornell_Box {w}x{h} {spp}spp depth{depth} truecolor, {n} accumulated frames", "conv": conv}
r = pkg.Renderer(w, h, conv=conv)
r.upload_scene(prims)
for rep in range(3):                                   # sequential, one call per frame
    t0 = time.perf_counter(); samples = 0; nbytes = 0
    for u in us:
        nbytes += len(r.render_frame(u, pkg.MODE_TRUECOLOR)); samples += r.frame_stats().samples_traced
    dt = time.perf_counter() - t0
res["sequential"] = {"ms": dt * 1e3, "ms_per_frame": dt * 1e3 / n, "Msamples_per_s": samples / dt / 1e6, "stream_bytes": nbytes}
seq_px = r.read_pixels().copy()
for name, emit in (("batch_all_streams", True), ("batch_last_stream", False)):
    for rep in range(3):
        t0 = time.perf_counter()
        out = r.render_frames(us, pkg.MODE_TRUECOLOR, emit_all=emit)
        dt = time.perf_counter() - t0
    st, tm = r.frame_stats(), r.timings()
    res[name] = {"ms": dt * 1e3, "ms_per_frame": dt * 1e3 / n, "Msamples_per_s": st.samples_traced / dt / 1e6,
                 "stream_bytes": sum(len(o) for o in out), "trace_plus_blend_ms": tm.trace_ms, "kernel_launches": tm.kernel_launches}
import numpy as np
d = np.abs(r.read_pixels() - seq_px).max()
res["max_abs_diff_batch_vs_sequential"] = float(d)
print(json.dumps(res))
