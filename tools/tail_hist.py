"""Tail diagnostics of the trace kernel (measurement build, TRT_B200_JIT=force TRT_B200_JIT_DEFS=-DTRT_TAIL_DIAG=1, or =2 with LATE=1 for the list of the last pixels):
per frame, when warps first found the work list empty and when they left (0.5 ms bins).
python tools/tail_hist.py [shard_count] [frames]"""
import json, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import __graft_entry__ as g
from cuda.bindings import runtime as rt
pkg, O = g.load_package(), g.load_oracle()
S = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "scenes")
sc = json.load(open(os.path.join(S, "Cornell_Box.json")))
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 6
w, h, spp, depth = 3840, 2160, 1024, 32
prims = O.flatten_scene(sc)
u = pkg.api.uniforms_from_bytes(O.make_uniforms(sc, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=depth).tobytes())
r = pkg.Renderer(w, h, conv=2, shard_index=0, shard_count=n, stripe_rows=8 if n > 1 else 1)
r.upload_scene(prims)
ptr, nbytes = r.device_buffer(4)
for k in range(frames):
    r.render_frame(u, pkg.MODE_NONE)
    ms = r.timings().trace_ms
    st = np.zeros(nbytes // 8, np.uint64)
    err, = rt.cudaMemcpy(st.ctypes.data, ptr, nbytes, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost)
    assert err == rt.cudaError_t.cudaSuccess, err
    ex, dry = st[8:520].astype(np.int64), st[520:1032].astype(np.int64)
    def q(hh, f):
        c = np.cumsum(hh)
        return 0.5 * float(np.searchsorted(c, f * c[-1])) if c[-1] else -1.0
    print(f"frame {k}: trace {ms:.1f} ms; warps {ex.sum()}; dry at p1/p50/p99 {q(dry, .01):.1f}/{q(dry, .5):.1f}/{q(dry, .99):.1f} ms; "
          f"exit p1/p10/p50/p90/p99/max {q(ex, .01):.1f}/{q(ex, .1):.1f}/{q(ex, .5):.1f}/{q(ex, .9):.1f}/{q(ex, .99):.1f}/{0.5 * np.nonzero(ex)[0].max():.1f}", flush=True)
    if k == frames - 1 and os.environ.get("LATE"):
        ev = st[520:520 + 500].reshape(250, 2)
        rows = sorted(((int(b & 0xffffffff) / 10, int(b >> 32) / 10, int(a & 0xffff), int((a >> 16) & 0xffff), int(a >> 32)) for a, b in ev), reverse=True)
        print("last pixels to finish (end ms, start ms, x, y, samples):")
        for e in rows[:40]:
            print("  %.1f %.1f x=%d y=%d n=%d" % e)
    if k == frames - 1:
        nz = np.nonzero(ex)[0]
        print("exit histogram (ms: warps):", " ".join(f"{0.5 * i:.1f}:{ex[i]}" for i in nz[-40:]))
