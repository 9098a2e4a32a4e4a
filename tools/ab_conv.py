"""A/B timing of the trace kernel across arithmetic conventions on a reduced C3 frame (same scene,
depth and sampling law; fewer pixels).  usage: python tools/ab_conv.py [width height spp] [convs]"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g

pkg, O = g.load_package(), g.load_oracle()
w, h, spp = (int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])) if len(sys.argv) > 3 else (1920, 1080, 256)
convs = [int(c) for c in sys.argv[4].split(",")] if len(sys.argv) > 4 else [0, 2]
prims = O.flatten_scene(O.CORNELL)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=32).tobytes())
for conv in convs:
    r = pkg.Renderer(w, h, conv=conv)
    r.upload_scene(prims)
    best = 1e9
    for k in range(4):
        r.render_frame(u, pkg.MODE_NONE)
        t = r.timings().trace_ms
        if k:
            best = min(best, t)
    st = r.frame_stats()
    print(f"conv {conv}: {w}x{h} {spp}spp trace {best:.2f} ms, {st.samples_traced / best / 1e3:.1f} Msamples/s, "
          f"{st.bounces / st.samples_traced:.2f} bounces/sample", flush=True)
    r.close()
