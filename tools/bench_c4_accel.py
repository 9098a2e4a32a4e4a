"""C4 scene (20 484 primitives) through the opt-in uniform grid (trt_config.accel = TRT_ACCEL_GRID) against the linear
scan, same frame: 1920x1080, reduced spp (the acceleration structure is outside the parity contract and reported
separately).  usage: python tools/bench_c4_accel.py [spp]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4
w, h = 1920, 1080
sc = O.make_c4_scene(width=w, height=h, spp=spp)
prims = O.flatten_scene(sc)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(sc, len(prims), O.C4_SEED, 0).tobytes())
res = {"config": f"C4 scene {w}x{h} {spp} spp depth 32, conv 2"}
frames = {}
for tag, accel in (("linear_scan", 0), ("uniform_grid", pkg.api.ACCEL_GRID)):
    r = pkg.Renderer(w, h, conv=2, accel=accel)
    r.upload_scene(prims)
    best = 1e30
    for k in range(3):
        r.render_frame(u, pkg.MODE_NONE)
        best = min(best, r.timings().trace_ms)
    st = r.frame_stats()
    frames[tag] = r.read_pixels()
    res[tag] = {"kernel": r.kernel_name(), "trace_ms": best, "Msamples_per_s": st.samples_traced / best / 1e3, "samples": int(st.samples_traced)}
    r.close()
a, b = frames["linear_scan"][:, :3].astype(np.float64), frames["uniform_grid"][:, :3].astype(np.float64)
res["speedup"] = res["linear_scan"]["trace_ms"] / res["uniform_grid"]["trace_ms"]
res["pixels_bit_equal"] = int((frames["linear_scan"].view(np.uint32) == frames["uniform_grid"].view(np.uint32)).all(axis=1).sum())
res["rmse_between_frames"] = float(np.sqrt(((a - b) ** 2).mean()))
print(json.dumps(res))
