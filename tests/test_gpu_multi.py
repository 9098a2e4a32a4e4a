"""Multi-GPU through the C ABI (trt_multi_*: one process, one row-sharded context per device), the row-addressed
stream combine it shares with the multi-process driver, and the pipelined frame loop (trt_frame_submit / collect).
On a one-GPU box the shards all live on device 0: the code path is the same, only the overlap is missing.
Bar: byte streams and float frames bit-identical to the single-context frame (conv 0), and to the oracle."""
import json
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def devices(n):
    import torch
    have = torch.cuda.device_count()
    return [i % max(1, have) for i in range(n)]


@pytest.mark.parametrize("n,stripe,mode", [(2, 4, 0), (3, 1, 1), (8, 4, 1)])
def test_multi_frame_equals_single_context_frame(pkg, O, n, stripe, mode):
    prims = O.flatten_scene(O.CORNELL)
    w, h = 96, 50
    u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), 11, 0, width=w, height=h, spp=16, max_depth=8).tobytes())
    one = pkg.Renderer(w, h, conv=0)
    one.upload_scene(prims)
    want = one.render_frame(u, mode)
    want_px, want_var = one.read_pixels(), one.read_variance()
    want_samples = int(one.frame_stats().samples_traced)
    one.close()
    m = pkg.MultiRenderer(w, h, devices(n), conv=0, stripe_rows=stripe)
    assert len(m) == n
    m.upload_scene(prims)
    got = m.render_frame(u, mode)
    assert got == want
    assert np.array_equal(bits(m.read_pixels()), bits(want_px))
    assert np.array_equal(bits(m.read_variance()["variance"]), bits(want_var["variance"]))
    assert int(m.frame_stats().samples_traced) == want_samples
    # a second frame accumulates on every shard's resident rows (frame_number 1)
    u2 = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), 12, 1, width=w, height=h, spp=16, max_depth=8).tobytes())
    one = pkg.Renderer(w, h, conv=0)
    one.upload_scene(prims)
    one.render_frame(u, mode)
    want2 = one.render_frame(u2, mode)
    one.close()
    assert m.render_frame(u2, mode) == want2
    m.close()


def test_multi_frame_against_the_oracle(pkg, O):
    sc = json.load(open(os.path.join(ROOT, "tests", "golden", "scenes", "demo.json")))
    prims = O.flatten_scene(sc)
    w, h = 80, 37
    u_np = O.make_uniforms(sc, len(prims), 5, 0, width=w, height=h, spp=8, max_depth=16)
    ref = (O.ref_render if O.ref_available() else O.render)(prims, u_np)
    m = pkg.MultiRenderer(w, h, devices(4), conv=0, stripe_rows=2)
    m.upload_scene(prims)
    stream = m.render_frame(pkg.api.uniforms_from_bytes(u_np.tobytes()), pkg.MODE_TRUECOLOR)
    assert np.array_equal(bits(m.read_pixels()), bits(ref["pixels"]))
    assert stream == O.quantise(ref["pixels"], w, h, 1)[0]
    m.close()


def test_row_lengths_and_row_addressed_copy(pkg, O):
    """The pieces the multi-process driver uses: per-row lengths of a shard, and the copy of its rows to offsets."""
    prims = O.flatten_scene(O.SCENE2)
    w, h, n, stripe = 64, 29, 3, 2
    u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.SCENE2, len(prims), 3, 0, width=w, height=h, spp=8, max_depth=4).tobytes())
    one = pkg.Renderer(w, h, conv=0)
    one.upload_scene(prims)
    want = one.render_frame(u, pkg.MODE_ASCII)
    one.close()
    from terminal_raytracer_b200 import multigpu
    shards, lens = [], []
    for r in range(n):
        c = pkg.Renderer(w, h, conv=0, shard_index=r, shard_count=n, stripe_rows=stripe)
        c.upload_scene(prims)
        c.trace_async(u)
        c.quantise_rows_async(pkg.MODE_ASCII)
        lens.append(c.row_lengths())
        assert len(lens[-1]) == len(multigpu.shard_rows(h, r, n, stripe))
        shards.append(c)
    off = multigpu.stream_row_offsets(lens, h, n, stripe)
    assert int(off[-1]) == len(want)
    rows = want.split(b"\r\n")[:-1]
    assert [len(x) + 2 for x in rows] == [int(off[y + 1] - off[y]) for y in range(h)]
    import ctypes as C
    cap = len(want) + 16
    host = pkg.api._P()
    assert pkg.api.lib().trt_alloc_host(cap, C.byref(host)) == 0
    C.memset(host, 0, cap)
    for c in shards:
        c.read_stream_rows(host, cap, off, wait=True)
    assert C.string_at(host, len(want)) == want
    # a buffer that is too small is refused, not overrun
    with pytest.raises(pkg.TrtError):
        shards[0].read_stream_rows(host, len(want) - 1, off, wait=True)
    pkg.api.lib().trt_free_host(host)
    for c in shards:
        c.close()


@pytest.mark.parametrize("conv", [0, 2])
def test_pipelined_frames_equal_serial_frames(pkg, O, conv):
    """trt_frame_submit / trt_frame_collect with a MOVING camera (every frame frame_number 0, as in C5) and with an
    accumulating one: same frames, same order, same bytes as trt_render_frame per frame."""
    scene = pkg.Scene.parse(json.dumps(O.SCENE2))
    info = scene.info
    serial = pkg.Renderer(info.width, info.height, conv=conv)
    piped = pkg.Renderer(info.width, info.height, conv=conv)
    serial.upload_scene(scene)
    piped.upload_scene(scene)
    cam = pkg.Camera()
    us = []
    for i in range(12):
        moved = cam.key(O.C5_KEYS[i % 8]) if i < 8 else False
        us.append(scene.uniforms(cam, O.c2_seed(i), 0 if moved else i - 7))
    want = [serial.render_frame(u, pkg.MODE_ASCII) for u in us]
    got = []
    piped.frame_submit(us[0], pkg.MODE_ASCII)
    for u in us[1:]:
        piped.frame_submit(u, pkg.MODE_ASCII)           # two in flight
        got.append(piped.frame_collect())               # the older one
    got.append(piped.frame_collect())
    assert got == want                                  # (scene2 has one light: conv 2 is run-to-run identical there too)
    with pytest.raises(pkg.TrtError):
        piped.frame_collect()                            # nothing in flight
    serial.close()
    piped.close()


def test_cli_pipeline_and_gpus_streams_equal_the_serial_loop(pkg, tmp_path):
    """The drop-in binary: --pipeline and --gpus 2 write the stream of the serial one-GPU loop (strict build)."""
    exe = os.path.join(ROOT, "terminal-raytracer_b200", "terminal-raytracer")
    scene = os.path.join(ROOT, "tests", "golden", "scenes", "scene2.json")
    keys = ",".join(["w", "Right", "d", "Up", "s", "Left", "a", "Down"] * 2 + ["x"] * 4)
    base = [exe, "--path", scene, "--strict", "--quiet", "--seed", "77", "--frames", "20", "--size", "120x40", "--keys", keys]
    outs = {}
    for tag, extra, env in (("serial", [], {}), ("pipeline", ["--pipeline"], {}),
                            ("gpus2", ["--gpus", "2"], {"TRT_B200_MULTI_SAME_DEVICE": "1"}),
                            ("color_pipeline", ["--pipeline", "--full-color"], {}), ("color", ["--full-color"], {})):
        out = str(tmp_path / f"{tag}.bin")
        r = subprocess.run(base + extra + ["--out", out], env=dict(os.environ, **env), capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stderr
        outs[tag] = open(out, "rb").read()
    assert len(outs["serial"]) > 20 * 120 * 40 * 15
    assert outs["pipeline"] == outs["serial"]
    assert outs["gpus2"] == outs["serial"]
    assert outs["color_pipeline"] == outs["color"]


@pytest.mark.parametrize("conv", [0, 2])
def test_frames_over_gpus_equal_sequential_frames(pkg, O, conv):
    """The second axis (C2's shape: frames of a still camera, frame_number = f): frame f on GPU f mod G, means gathered
    and blended in frame order on the first GPU -- bit-identical to the frames one trt_render_frame call at a time."""
    prims = O.flatten_scene(O.CORNELL)
    w, h, n = 100, 50, 11
    us = [pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), O.c2_seed(f), f, width=w, height=h, spp=16, max_depth=16).tobytes())
          for f in range(n)]
    one = pkg.Renderer(w, h, conv=conv)
    one.upload_scene(prims)
    for u in us:
        want = one.render_frame(u, pkg.MODE_TRUECOLOR)
    want_px, want_var = one.read_pixels(), one.read_variance()
    one.close()
    m = pkg.MultiRenderer(w, h, devices(3), conv=conv)
    m.upload_scene(prims)
    got = m.render_frames(us, pkg.MODE_TRUECOLOR)
    assert np.array_equal(bits(m.frames_pixels()), bits(want_px))
    assert np.array_equal(bits(m.frames_variance()["variance"]), bits(want_var["variance"]))
    assert got == want
    if conv == 0:                                        # and against the oracle's own loop
        acc = None
        for f in range(n):
            ref = O.render(prims, O.make_uniforms(O.CORNELL, len(prims), O.c2_seed(f), f, width=w, height=h, spp=16, max_depth=16), accumulation=acc)
            acc = ref["accumulation"]
        assert np.array_equal(bits(want_px), bits(ref["pixels"]))
    m.close()
