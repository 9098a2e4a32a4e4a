"""A/B of the run-time specialised kernel: python tools/ab_jit.py [scene2|demo|cornell|cornell_swapped] (set TRT_B200_JIT=0 / force outside)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
which = sys.argv[1] if len(sys.argv) > 1 else "scene2"
S = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes")
sc = json.load(open(os.path.join(S, {"scene2": "scene2.json", "demo": "demo.json"}.get(which, "Cornell_Box.json"))))
if which == "cornell_swapped":
    sc["planes"] = sc["planes"][2:] + sc["planes"][:2]
w, h, spp, depth = 1920, 1080, int(sys.argv[2]) if len(sys.argv) > 2 else 64, 32
prims = O.flatten_scene(sc)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(sc, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=depth).tobytes())
r = pkg.Renderer(w, h, conv=2)
t0 = time.perf_counter(); r.upload_scene(prims); r.render_frame(u, pkg.MODE_NONE); first = time.perf_counter() - t0
best = 1e9
for k in range(5):
    r.render_frame(u, pkg.MODE_NONE); best = min(best, r.timings().trace_ms)
st = r.frame_stats()
px = r.read_pixels().reshape(-1, 4)[:, :3].astype('float64')
ref_path = os.environ.get("AB_REF")            # first run writes the frame, later runs report their RMSE against it
if ref_path:
    import numpy as np
    if os.path.exists(ref_path):
        ref = np.load(ref_path)
        print(f"  vs {ref_path}: rmse {float(np.sqrt(((px - ref) ** 2).mean())):.3e}, mean {px.mean():.6f} (ref {ref.mean():.6f}), nan {int(np.isnan(px).sum())}")
    else:
        np.save(ref_path, px)
print(f"{os.environ.get('TRT_B200_JIT_DEFS', '')} {which}: {r.kernel_name()}: trace {best:.2f} ms, {st.samples_traced / best / 1e3:.1f} Msamples/s, {st.bounces / st.samples_traced:.2f} bounces/sample; first frame incl. upload/compile {first * 1e3:.0f} ms", flush=True)
