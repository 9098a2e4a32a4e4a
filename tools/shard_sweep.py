"""Per-rank trace times of an N-way row-sharded C3 frame, measured on ONE GPU: shard k of N is rendered by a sharded
context exactly as rank k of an N-GPU run would (same rows, same kernel, same hand-out), so max over k / (whole-frame
time / N) is the strong-scaling efficiency the trace kernel allows, without an N-GPU box.
usage: python tools/shard_sweep.py N stripe [w h spp]      (TRT_B200_CTAS etc. from the environment)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
n, stripe = int(sys.argv[1]), int(sys.argv[2])
w, h, spp = (int(v) for v in sys.argv[3:6]) if len(sys.argv) > 5 else (3840, 2160, 1024)
prims = O.flatten_scene(O.CORNELL)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=32).tobytes())
times, samples = [], []
for k in range(n):
    r = pkg.Renderer(w, h, conv=2, shard_index=k, shard_count=n, stripe_rows=stripe)
    r.upload_scene(prims)
    best = 1e30
    for f in range(4):                       # frame 0 has no cost history; later frames hand out longest chains first
        r.render_frame(u, pkg.MODE_NONE)
        if f >= 2:
            best = min(best, r.timings().trace_ms)
    times.append(best); samples.append(int(r.frame_stats().samples_traced))
    r.close()
tot = sum(samples)
print(f"N={n} stripe={stripe} CTAS={os.environ.get('TRT_B200_CTAS', 'hint')}: per-shard trace ms {[round(t, 2) for t in times]}; max {max(times):.2f}, mean {sum(times) / n:.2f}; "
      f"samples per shard min/max {min(samples)}/{max(samples)} ({100.0 * (max(samples) * n / tot - 1):.2f} % over the mean)", flush=True)
