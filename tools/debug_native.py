"""Where does conv 2 differ from the strict frame?  python tools/debug_native.py scene w h spp depth seed"""
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
    out[conv] = (r.read_pixels(), r.read_variance(), r.frame_stats().samples_traced); r.close()
a, b = out[0][0][:, :3].astype(np.float64), out[2][0][:, :3].astype(np.float64)
err = np.abs(a - b).max(axis=1) / np.maximum(np.abs(a).max(axis=1), 1e-3)
bad = np.nonzero(err > 1e-4)[0]
print(f"{name}: {len(bad)} outliers of {w*h}; samples {out[0][2]} vs {out[2][2]}")
ys, xs = bad // w, bad % w
print("rows histogram:", np.bincount(ys, minlength=h).nonzero()[0][:40], "cols range", xs.min() if len(bad) else None, xs.max() if len(bad) else None)
for i in bad[:25]:
    print(f"  ({i % w:3d},{i // w:3d}) strict {out[0][0][i,:3]} native {out[2][0][i,:3]} var {out[0][1][i]} / {out[2][1][i]}")
