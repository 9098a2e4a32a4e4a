"""trt_quantise_noalloc (include/trt_b200.h): the quantiser with its scratch in a static device array, for callers that
cannot allocate.  No counterpart in the reference.

Checked through the C ABI: on two calls in a row (ASCII, then truecolor, on the same frame; then a third call on a
smaller frame, so that the shared scratch still holds the larger call's tile states) the bytes equal those of the
existing entry points (trt_render_frame / trt_quantise_async + trt_read_stream) -- byte for byte, the tolerance every
quantiser test here uses -- and those of the CPU oracle; frames beyond the static limit are rejected with a text.
The file sorts last on purpose."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 8388608          # TRT_NOALLOC_MAX_PIXELS


def noalloc(pkg, torch, dptr, w, h, mode, stream, cap=None):
    """-> (rc, bytes, error text); outputs live in torch device tensors the CALLER owns"""
    cap = w * h * 26 + 2 * h + 64 if cap is None else cap
    out = torch.full((cap,), 0xEE, dtype=torch.uint8, device="cuda:0")
    n = torch.full((1,), -1, dtype=torch.int64, device="cuda:0")
    torch.cuda.synchronize()
    err = C.create_string_buffer(512)
    rc = pkg.api.lib().trt_quantise_noalloc(0, dptr, w, h, mode, out.data_ptr(), cap, n.data_ptr(), stream, err, len(err))
    torch.cuda.synchronize()
    length = int(n.item())
    data = bytes(out[:length].cpu().numpy()) if rc == 0 and 0 <= length <= cap else b""
    return rc, data, err.value.decode(), length


def test_noalloc_quantiser_equals_the_context_quantiser(pkg, O):
    import torch
    prims = O.flatten_scene(O.CORNELL)
    w, h = 400, 200                                       # 80 000 pixels = 79 tiles: the look-back chain is exercised
    u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), 7, 0, width=w, height=h, spp=2, max_depth=6).tobytes())
    r = pkg.Renderer(w, h, conv=0)
    r.upload_scene(prims)
    want_ascii = r.render_frame(u, pkg.MODE_ASCII)        # existing entry point
    r.quantise_async(pkg.MODE_TRUECOLOR)
    r.sync()
    want_color = r.read_stream()                          # existing entry point, same pixels
    pixels = r.read_pixels()
    assert want_ascii == O.quantise(pixels, w, h, 0)[0] and want_color == O.quantise(pixels, w, h, 1)[0]
    dptr, nbytes = r.device_buffer(0)
    assert nbytes >= w * h * 16
    stream = r.cuda_stream()
    # two calls in a row, then the first one again
    for k, (mode, want) in enumerate([(pkg.MODE_ASCII, want_ascii), (pkg.MODE_TRUECOLOR, want_color), (pkg.MODE_ASCII, want_ascii)]):
        rc, got, err, length = noalloc(pkg, torch, dptr, w, h, mode, stream)
        assert rc == 0, err
        assert length == len(want), f"call {k}: length {length} != {len(want)}"
        assert got == want, f"call {k}: bytes differ from the context's quantiser"
    # a smaller frame after a larger one: the states of the tiles it uses are cleared, the rest is stale and unused
    w2, h2 = 400, 37
    rc, got, err, _ = noalloc(pkg, torch, dptr, w2, h2, pkg.MODE_TRUECOLOR, None)
    assert rc == 0, err
    assert got == O.quantise(pixels[: w2 * h2], w2, h2, 1)[0]
    # an output buffer that is too small: nothing beyond it is written, the length is still reported
    small = 1000
    out = torch.full((small + 64,), 0xEE, dtype=torch.uint8, device="cuda:0")
    n = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    rc = pkg.api.lib().trt_quantise_noalloc(0, dptr, w, h, pkg.MODE_ASCII, out.data_ptr(), small, n.data_ptr(), None, None, 0)
    torch.cuda.synchronize()
    assert rc == 0 and int(n.item()) == len(want_ascii)
    assert bool((out[small:] == 0xEE).all())
    r.close()


def test_noalloc_rejects_what_the_static_scratch_cannot_hold(pkg):
    import torch
    px = torch.zeros(64, dtype=torch.float32, device="cuda:0")
    out = torch.zeros(1024, dtype=torch.uint8, device="cuda:0")
    n = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    err = C.create_string_buffer(512)
    f = pkg.api.lib().trt_quantise_noalloc
    # beyond the limit: rejected before anything is launched (the device pointers are never touched)
    rc = f(0, px.data_ptr(), 4096, 2049, 0, out.data_ptr(), 1024, n.data_ptr(), None, err, len(err))
    assert rc == -1                                        # TRT_ERR_INVALID
    assert b"TRT_NOALLOC_MAX_PIXELS" in err.value and str(4096 * 2049).encode() in err.value
    assert 4096 * 2048 == MAX_PIXELS                       # ... and the limit itself is accepted (checked on arguments only here)
    for bad in [(0, 0, 10, 0), (0, 10, 0, 0), (0, 4, 4, 7), (99, 4, 4, 0)]:
        dev, w, h, mode = bad
        assert f(dev, px.data_ptr(), w, h, mode, out.data_ptr(), 1024, n.data_ptr(), None, err, len(err)) == -1
        assert err.value
    assert f(0, None, 4, 4, 0, out.data_ptr(), 1024, n.data_ptr(), None, err, len(err)) == -1
    torch.cuda.synchronize()
