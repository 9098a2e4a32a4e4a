"""Dump conv 0 / conv 2 frames for offline analysis: python tools/dump_native.py scene w h spp depth seed out.npz"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
name, w, h, spp, depth, seed = sys.argv[1], *(int(v) for v in sys.argv[2:7])
sc = json.load(open(os.path.join("tests/golden/scenes", name)))
prims = O.flatten_scene(sc)
u = O.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=depth)
uu = pkg.api.uniforms_from_bytes(u.tobytes())
out = {}
for conv in (0, 2):
    r = pkg.Renderer(w, h, conv=conv); r.upload_scene(prims); r.render_frame(uu, pkg.MODE_NONE)
    out[f"px{conv}"] = r.read_pixels(); out[f"var{conv}"] = r.read_variance()["variance"]; r.close()
np.savez_compressed(sys.argv[7], **out)
