"""Generates tests/golden/frames.npz (run in the build container, committed with its output).

The reference holds no golden vectors and cannot be executed through wgpu here, but its shader can be compiled:
oracle/_ref/libref.so is /root/reference/src/shader.wgsl transliterated to C (oracle/wgsl2c.py).  The conv 0
pixel / variance arrays below are OUTPUTS OF THAT BUILD (`source: "_ref"` in frames.json); the generator refuses to
write them unless the hand-written oracle returns the same bits.  The conv 1 arrays (fused helpers -- an arithmetic
convention, not reference text), the work counters and the stream hashes (quantiser = hand restatement of
lib.rs:503-532) come from the hand oracle.  Cases cover the three shipped scenes, both arithmetic conventions,
adaptive sampling, accumulation over frames and both stream modes."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import __graft_entry__ as g  # noqa: E402

CASES = [  # name, scene file, w, h, spp, depth, seed, frames
    ("cornell_c1", "Cornell_Box.json", 48, 24, 16, 32, 0x00C0FFEE, 1),
    ("cornell_file", "Cornell_Box.json", 40, 20, 128, 3, 5, 1),
    ("cornell_accum", "Cornell_Box.json", 32, 16, 16, 8, 0x00C0FFEE, 3),
    ("scene2", "scene2.json", 50, 25, 32, 4, 1, 1),
    ("demo", "demo.json", 40, 20, 16, 32, 2, 1),
]


def run_case(O, case, conv, from_ref=False):
    """from_ref: take pixels / variance from the transliterated reference shader (conv 0 only)."""
    name, scene, w, h, spp, depth, seed, frames = case
    sc = json.load(open(os.path.join(HERE, "scenes", scene)))
    prims = O.flatten_scene(sc)
    acc, out, acc_ref, ref = None, None, None, None
    for f in range(frames):
        s = O.c2_seed(f) if frames > 1 else seed
        u = O.make_uniforms(sc, len(prims), s, f, width=w, height=h, spp=spp, max_depth=depth)
        out = O.render(prims, u, accumulation=acc, conv=conv, want_counters=(f == 0))
        if f == 0:
            counters = out["counters"]
        acc = out["accumulation"]
        if from_ref:
            ref = O.ref_render(prims, u, accumulation=acc_ref)
            acc_ref = ref["accumulation"]
    if from_ref:
        for k in ("pixels", "accumulation"):
            if not np.array_equal(ref[k].view(np.uint32), out[k].view(np.uint32)):
                raise SystemExit(f"{name}: hand oracle and oracle/_ref disagree on {k}; fixtures not written")
        if not np.array_equal(ref["variance"].view(np.uint32), out["variance"].view(np.uint32)):
            raise SystemExit(f"{name}: hand oracle and oracle/_ref disagree on variance; fixtures not written")
        out = dict(out, pixels=ref["pixels"], variance=ref["variance"])
    streams = [hashlib.sha256(O.quantise(out["pixels"], w, h, m)[0]).hexdigest() for m in (0, 1)]
    return out["pixels"], out["variance"], counters, streams


def main():
    O = g.load_oracle()
    if not O.ref_available():
        raise SystemExit("oracle/_ref is not built (needs /root/reference): the fixtures come from it")
    data, meta = {}, {}
    for case in CASES:
        for conv in (0, 1):
            px, var, cnt, streams = run_case(O, case, conv, from_ref=(conv == 0))
            key = f"{case[0]}_conv{conv}"
            data[key + "_pixels"] = px.view(np.uint32)
            data[key + "_variance"] = var["variance"].view(np.uint32)
            meta[key] = {"case": list(case), "conv": conv, "counters": cnt, "stream_sha256": streams,
                         "source": "_ref" if conv == 0 else "oracle"}
    np.savez_compressed(os.path.join(HERE, "frames.npz"), **data)
    json.dump(meta, open(os.path.join(HERE, "frames.json"), "w"), indent=1, sort_keys=True)
    print("wrote", len(data), "arrays")


if __name__ == "__main__":
    main()
