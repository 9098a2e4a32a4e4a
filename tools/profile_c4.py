"""C4 timing probe: python tools/profile_c4.py w h spp"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
w, h, spp = (int(v) for v in sys.argv[1:4])
sc = O.make_c4_scene(width=w, height=h, spp=spp)
prims = O.flatten_scene(sc)
r = pkg.Renderer(w, h)
t0 = time.time(); r.upload_scene(prims); print("upload %.3f s, %d prims" % (time.time() - t0, len(prims)))
u = pkg.api.uniforms_from_bytes(O.make_uniforms(sc, len(prims), O.C4_SEED, 0).tobytes())
for _ in range(2):
    r.render_frame(u, pkg.MODE_ASCII)
    t, st = r.timings(), r.frame_stats()
    print(f"C4 {w}x{h} spp {spp}: trace {t.trace_ms:.1f} ms; {st.samples_traced} samples {st.bounces} bounces -> {st.samples_traced / t.trace_ms / 1e3:.2f} Msamples/s, {st.bounces / t.trace_ms / 1e3:.2f} Mbounces/s")
