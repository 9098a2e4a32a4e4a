import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
sc = O.make_c4_scene(); prims = O.flatten_scene(sc)
w, h = 640, 360
r = pkg.Renderer(w, h, conv=2); r.upload_scene(prims)
u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0, width=w, height=h, spp=4, max_depth=32)
r.render_frame(pkg.api.uniforms_from_bytes(u.tobytes()), pkg.MODE_NONE)
np.save(sys.argv[1], r.read_pixels()); print(r.kernel_name()); r.close()
