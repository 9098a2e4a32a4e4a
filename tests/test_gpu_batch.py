"""GPU tests of the frame batch (trt_render_frames): n frames of a still camera traced by ONE launch and
blended in frame order.  Bar: in the strict conventions everything the batch leaves behind -- pixels,
accumulation, variance, every frame's byte stream -- is bit-identical to n trt_render_frame calls AND to the
CPU oracle's frame loop (lib.rs:414-560 / shader.wgsl:605-621); in the device-native convention (conv 2) the
batch equals its own sequential frames to summation-order noise (the deferred light contributions of one
sample are added in a lane-placement dependent order) and the streams differ only where the floats do."""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SCENES = os.path.join(os.path.dirname(__file__), "golden", "scenes")


def _scene(name):
    with open(os.path.join(SCENES, name)) as f:
        return json.load(f)


def _u(pkg, u_np):
    return pkg.api.uniforms_from_bytes(u_np.tobytes())


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def _frames(O, sc, n_prims, n, w, h, spp, depth, first_frame=0):
    return [O.make_uniforms(sc, n_prims, O.c2_seed(f), f, width=w, height=h, spp=spp, max_depth=depth)
            for f in range(first_frame, first_frame + n)]


@pytest.mark.parametrize("conv", [0, 1])
@pytest.mark.parametrize("name,w,h,spp,depth,n", [("Cornell_Box.json", 64, 32, 16, 8, 6), ("scene2.json", 50, 23, 8, 4, 5)])
def test_batch_equals_sequential_and_oracle(pkg, O, name, w, h, spp, depth, n, conv):
    sc = _scene(name)
    prims = O.flatten_scene(sc)
    us = _frames(O, sc, len(prims), n, w, h, spp, depth)
    # sequential frames through trt_render_frame, and the oracle's loop
    seq = pkg.Renderer(w, h, conv=conv)
    seq.upload_scene(prims)
    acc, streams = None, []
    for u in us:
        ref = O.render(prims, u, conv=conv, accumulation=acc)
        acc = ref["accumulation"]
        streams.append(seq.render_frame(_u(pkg, u), pkg.MODE_TRUECOLOR))
        assert np.array_equal(_bits(seq.read_pixels()), _bits(ref["pixels"]))
    # the same frames as one batch
    bat = pkg.Renderer(w, h, conv=conv)
    bat.upload_scene(prims)
    got = bat.render_frames([_u(pkg, u) for u in us], pkg.MODE_TRUECOLOR, emit_all=True)
    assert len(got) == n
    for f in range(n):
        assert got[f] == streams[f], f"stream of frame {f}"
    assert np.array_equal(_bits(bat.read_pixels()), _bits(seq.read_pixels()))
    assert np.array_equal(_bits(bat.read_accumulation()), _bits(seq.read_accumulation()))
    assert np.array_equal(_bits(bat.read_accumulation()), _bits(acc))
    vb, vs = bat.read_variance(), seq.read_variance()
    assert np.array_equal(np.ascontiguousarray(vb).view(np.uint8), np.ascontiguousarray(vs).view(np.uint8))
    assert bat.frame_stats().samples_traced > 0
    # only the last frame's stream
    bat2 = pkg.Renderer(w, h, conv=conv)
    bat2.upload_scene(prims)
    last = bat2.render_frames([_u(pkg, u) for u in us], pkg.MODE_ASCII)
    assert last == [O.quantise(acc, w, h, 0)[0]]
    seq.close(); bat.close(); bat2.close()


def test_batch_continues_an_accumulation(pkg, O):
    """frames 0-2 one by one, frames 3-6 as a batch: same as 7 sequential frames (the batch reads the
    accumulation buffer it finds); a frame_number-0 frame inside a batch resets it."""
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h = 48, 24
    us = _frames(O, sc, len(prims), 7, w, h, 16, 6)
    a = pkg.Renderer(w, h)
    a.upload_scene(prims)
    for u in us:
        a.render_frame(_u(pkg, u), pkg.MODE_NONE)
    b = pkg.Renderer(w, h)
    b.upload_scene(prims)
    for u in us[:3]:
        b.render_frame(_u(pkg, u), pkg.MODE_NONE)
    b.render_frames([_u(pkg, u) for u in us[3:]], pkg.MODE_NONE)
    assert np.array_equal(_bits(a.read_accumulation()), _bits(b.read_accumulation()))
    assert np.array_equal(_bits(a.read_pixels()), _bits(b.read_pixels()))
    # reset inside a batch: [f=5, f=0, f=1] leaves what [f=0, f=1] leave
    mixed = [us[5], us[0], us[1]]
    b.render_frames([_u(pkg, u) for u in mixed], pkg.MODE_NONE)
    c = pkg.Renderer(w, h)
    c.upload_scene(prims)
    for u in us[:2]:
        c.render_frame(_u(pkg, u), pkg.MODE_NONE)
    assert np.array_equal(_bits(b.read_accumulation()), _bits(c.read_accumulation()))
    a.close(); b.close(); c.close()


def test_batch_native_tracks_sequential(pkg, O):
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h, n = 80, 40, 8
    us = _frames(O, sc, len(prims), n, w, h, 16, 32)
    seq = pkg.Renderer(w, h, conv=2)
    seq.upload_scene(prims)
    for u in us:
        s_last = seq.render_frame(_u(pkg, u), pkg.MODE_TRUECOLOR)
    bat = pkg.Renderer(w, h, conv=2)
    bat.upload_scene(prims)
    b_last = bat.render_frames([_u(pkg, u) for u in us], pkg.MODE_TRUECOLOR)[0]
    ps, pb = seq.read_pixels().reshape(-1, 4)[:, :3], bat.read_pixels().reshape(-1, 4)[:, :3]
    rel = np.abs(pb - ps) / np.maximum(np.abs(ps), 1e-2)
    print(f"  conv 2 batch vs sequential: max rel {rel.max():.3g}, pixels differing in bits {(ps != pb).any(axis=1).sum()} of {w * h}")
    assert rel.max() <= 1e-5
    assert b_last == O.quantise(bat.read_pixels().reshape(h, w, 4), w, h, 1)[0]     # exact quantiser on its own pixels
    assert sum(x != y for x, y in zip(s_last, b_last)) <= 8 and abs(len(s_last) - len(b_last)) <= 4
    assert bat.frame_stats().samples_traced >= n * w * h * 4
    seq.close(); bat.close()


def test_batch_rejects_differing_frames(pkg, O):
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    r = pkg.Renderer(32, 16)
    r.upload_scene(prims)
    u0 = O.make_uniforms(sc, len(prims), 1, 0, width=32, height=16, spp=4, max_depth=2)
    u1 = O.make_uniforms(sc, len(prims), 2, 1, width=32, height=16, spp=8, max_depth=2)        # spp differs
    with pytest.raises(pkg.TrtError):
        r.render_frames([_u(pkg, u0), _u(pkg, u1)], pkg.MODE_NONE)
    r.close()
