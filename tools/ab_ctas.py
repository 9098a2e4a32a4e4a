"""Frame time of the C3 frame (Cornell 3840x2160, 1024 spp, depth 32) and of one 8th of it (the rows rank 0 of 8
owns) for the resident-CTA count in TRT_B200_CTAS (set outside): python tools/ab_ctas.py [shard_count]"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
pkg, O = g.load_package(), g.load_oracle()
S = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes")
sc = json.load(open(os.path.join(S, "Cornell_Box.json")))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 8
w, h, spp, depth = 3840, 2160, 1024, 32
prims = O.flatten_scene(sc)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(sc, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=depth).tobytes())
r = pkg.Renderer(w, h, conv=int(os.environ.get("CONV", "2")), shard_index=int(os.environ.get("SHARD", "0")), shard_count=n,
                 stripe_rows=int(os.environ.get("STRIPE", "8")) if n > 1 else 1)
r.upload_scene(prims)
ts = []
for k in range(int(os.environ.get('RUNS', '4'))):
    r.render_frame(u, pkg.MODE_NONE)
    ts.append(r.timings().trace_ms)
st = r.frame_stats()
best = min(ts[1:])
print(f"{os.environ.get('TRT_B200_JIT_DEFS', '')} ctas={os.environ.get('TRT_B200_CTAS', '8')} shards={n} shard={os.environ.get('SHARD', '0')} stripe={os.environ.get('STRIPE', '8')}: trace {best:.2f} ms (runs {' '.join('%.1f' % t for t in ts)}), "
      f"{st.samples_traced / best / 1e3:.1f} Msamples/s", flush=True)
