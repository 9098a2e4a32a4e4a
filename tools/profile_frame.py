"""One frame of the hot path, for ncu: python tools/profile_frame.py [w h spp conv scene]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g

pkg, O = g.load_package(), g.load_oracle()
w, h, spp = (int(v) for v in (sys.argv[1:4] + ["960", "540", "64"][len(sys.argv[1:4]):]))
conv = int(sys.argv[4]) if len(sys.argv) > 4 else 0
scene = sys.argv[5] if len(sys.argv) > 5 else "cornell"
sc = {"cornell": O.CORNELL, "scene2": O.SCENE2}.get(scene) or O.make_c4_scene()
prims = O.flatten_scene(sc)
r = pkg.Renderer(w, h, conv=conv)
r.upload_scene(prims)
for f in range(2):
    u = O.make_uniforms(sc, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=32)
    r.render_frame(pkg.api.uniforms_from_bytes(u.tobytes()), pkg.MODE_TRUECOLOR)
t, st = r.timings(), r.frame_stats()
print(f"{scene} {w}x{h} spp {spp} conv {conv}: trace {t.trace_ms:.3f} ms quantise {t.quantise_ms:.3f} ms d2h {t.d2h_ms:.3f} ms; "
      f"{st.samples_traced} samples, {st.bounces} bounces -> {st.samples_traced / t.trace_ms / 1e3:.1f} Msamples/s; "
      f"pixels sha1 {__import__('hashlib').sha1(r.read_pixels().tobytes()).hexdigest()[:12]}")
r.close()
