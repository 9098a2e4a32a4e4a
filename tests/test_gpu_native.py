"""GPU tolerance tests of the device-native arithmetic convention (conv 2) against the CPU oracle.

conv 2 runs the same kernel source as the strict build (same state machine, same random numbers, same
tests in the same scan order) with FFMA contraction and the SM's reciprocal / rsqrt / sqrt / sin / cos
units -- the latitude WGSL gives the reference's own shader compiler.  It cannot be bit-exact against a
CPU; the bars are BASELINE.json north_star's floating-point tolerances, written here:

  * single-bounce (max_depth = 1) radiance within 1e-4 relative per pixel, fixed seed (relative to
    max(|ref|, 0.01): 1 % of unit radiance is the floor below which a pixel is black on any terminal);
    pixels whose value differs by more are counted and bounded by the oracle's own yardstick: the
    number of pixels on which its two exact conventions (conv 0 / conv 1) disagree with each other
    (see _assert_within_conventions: the shader is ill-conditioned at NEE self-occlusion, far-field
    sphere tests and silhouettes, where ANY two roundings of the same WGSL disagree);
  * deep paths: the frame tracks the strict frame far below the Monte-Carlo noise floor (RMSE against
    the oracle <= 5 % of the oracle's own seed-to-seed RMSE);
  * converged multi-frame images: RMSE within the noise floor of the accumulation;
  * the quantiser is the same exact kernel: the stream is byte-identical to the oracle's quantiser
    on the frame's own pixels, and against the strict frame's stream only pixels whose floats differ
    can differ;
  * row shards stay bit-identical to the whole frame (determinism of the convention).
"""
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
SCENES = os.path.join(os.path.dirname(__file__), "golden", "scenes")
REL_TOL = 1e-4          # north_star: single-bounce radiance, relative, per pixel


def _scene(name):
    with open(os.path.join(SCENES, name)) as f:
        return json.load(f)


def _u(pkg, u_np):
    return pkg.api.uniforms_from_bytes(u_np.tobytes())


def _render(pkg, prims, u, w, h, conv, mode=None):
    r = pkg.Renderer(w, h, conv=conv)
    r.upload_scene(prims)
    stream = r.render_frame(_u(pkg, u), pkg.MODE_NONE if mode is None else mode)
    px, var = r.read_pixels(), r.read_variance()
    st = r.frame_stats()
    r.close()
    return px, var, stream, st


def _rel_err(got, ref):
    """per-pixel max over RGB of |got - ref| / max(|ref|, floor); floor keeps black pixels meaningful"""
    g, r = got[:, :3].astype(np.float64), ref[:, :3].astype(np.float64)
    scale = np.maximum(np.abs(r).max(axis=1), 1e-2)
    return np.abs(g - r).max(axis=1) / scale


def _rmse(a, b):
    d = a[:, :3].astype(np.float64) - b[:, :3].astype(np.float64)
    return float(np.sqrt((d * d).mean()))


SINGLE_BOUNCE = [
    ("Cornell_Box.json", 160, 80, 16, 0x00C0FFEE),
    ("scene2.json", 200, 100, 32, 1),                  # sphere light: NEE live
    ("demo.json", 160, 80, 16, 2),                     # 21 primitives, open scene (sky)
]


def _compare(got, ref, w, h):
    """(pixels over REL_TOL, pixels with |diff| > 5e-3 [a whole sample took the other branch], L1 relative difference)"""
    err = _rel_err(got, ref)
    d_abs = np.abs(got[:, :3].astype(np.float64) - ref[:, :3].astype(np.float64)).max(axis=1)
    l1 = d_abs.sum() / max(np.abs(ref[:, :3]).max(axis=1).sum(), 1e-30)
    return int((err > REL_TOL).sum()), int((d_abs > 5e-3).sum()), float(l1), err


def _assert_within_conventions(tag, got, ref0, ref1, w, h):
    """The 1e-4 bar with its counted exceptions.  The shader itself is ill-conditioned in places: the shadow ray
    starts 0.001 along the normal but ends 0.001 before a distance measured from the un-offset point (:394-400), so
    the light's own surface sits within +-0.001-scale of the ray's end and `t < closest` flips under ANY change of
    rounding (1e-6 / 1e-3 = one NEE sample in a thousand); sphere tests cancel catastrophically for far origins
    (horizon rows); rays graze silhouettes.  The yardstick is therefore the oracle itself: its two exact conventions
    (conv 0 as written, conv 1 fused helpers -- both bit-exact against the strict GPU builds) disagree on N pixels;
    the native build may disagree with conv 0 on at most 2 N + 0.05 % of the frame, and likewise for whole-sample
    flips and for the image-level L1 difference."""
    base_out, base_gross, base_l1, _ = _compare(ref1, ref0, w, h)
    out, gross, l1, err = _compare(got, ref0, w, h)
    print(f"\n{tag}: native vs conv 0: {out} of {w * h} pixels over {REL_TOL} (max {err.max():.3g}, median {np.median(err):.3g}), "
          f"{gross} whole-sample flips, L1 {l1:.3g};  oracle conv 1 vs conv 0: {base_out} pixels, {base_gross} flips, L1 {base_l1:.3g}")
    assert np.median(err) < 2e-6
    assert out <= 2 * base_out + max(3, w * h // 2000), f"{out} pixels over {REL_TOL} (oracle's own conventions: {base_out})"
    assert gross <= 2 * base_gross + 5
    assert l1 <= 2 * base_l1 + 1e-4
    return out


@pytest.mark.parametrize("name,w,h,spp,seed", SINGLE_BOUNCE)
def test_single_bounce_radiance_within_1e4(pkg, O, name, w, h, spp, seed):
    sc = _scene(name)
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=1)
    ref0, ref1 = O.render(prims, u, conv=0), O.render(prims, u, conv=1)
    got, var, _, st = _render(pkg, prims, u, w, h, 2)
    _assert_within_conventions(name, got, ref0["pixels"], ref1["pixels"], w, h)
    assert abs(int(st.samples_traced) - int(ref0["samples_traced"])) <= w * h // 100


def test_deep_paths_track_strict_below_noise(pkg, O):
    """C1-like frame (depth 32): differences to the strict frame vs the Monte-Carlo noise floor."""
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h = 128, 64
    ua = O.make_uniforms(sc, len(prims), 0x00C0FFEE, 0, width=w, height=h, spp=16, max_depth=32)
    ub = O.make_uniforms(sc, len(prims), 0x00C0FFEF, 0, width=w, height=h, spp=16, max_depth=32)
    ref_a, ref_b = O.render(prims, ua)["pixels"], O.render(prims, ub)["pixels"]
    got, _, _, _ = _render(pkg, prims, ua, w, h, 2)
    noise = _rmse(ref_a, ref_b)
    rm = _rmse(got, ref_a)
    err = _rel_err(got, ref_a)
    frac_close = float((err <= 1e-3).mean())
    print(f"\nnoise floor (seed-to-seed RMSE) {noise:.4g}; native-vs-strict RMSE {rm:.4g} ({rm / noise:.3%} of it); "
          f"pixels within 1e-3: {frac_close:.2%}, within 1e-4: {(err <= 1e-4).mean():.2%}")
    assert rm <= 0.05 * noise
    assert frac_close >= 0.97


def test_converged_image_rmse(pkg, O):
    """24 accumulated frames (shader.wgsl:610-620): native accumulation vs the oracle's, RMSE far inside
    the noise floor of the accumulated image (estimated from two independent oracle accumulations)."""
    sc = _scene("scene2.json")
    prims = O.flatten_scene(sc)
    w, h, frames = 96, 48, 24
    r = pkg.Renderer(w, h, conv=2)
    r.upload_scene(prims)
    acc_a = acc_b = None
    for f in range(frames):
        u = O.make_uniforms(sc, len(prims), O.c2_seed(f), f, width=w, height=h, spp=8, max_depth=4)
        acc_a = O.render(prims, u, accumulation=acc_a)["accumulation"]
        ub = O.make_uniforms(sc, len(prims), O.c2_seed(f + 1000), f, width=w, height=h, spp=8, max_depth=4)
        acc_b = O.render(prims, ub, accumulation=acc_b)["accumulation"]
        r.render_frame(_u(pkg, u), pkg.MODE_NONE)
    got = r.read_pixels()
    r.close()
    noise, rm = _rmse(acc_a, acc_b), _rmse(got, acc_a)
    print(f"\nconverged: noise floor {noise:.4g}, native-vs-oracle RMSE {rm:.4g} ({rm / noise:.3%})")
    assert rm <= 0.05 * noise


@pytest.mark.parametrize("mode", [0, 1])
def test_stream_bytes_native(pkg, O, mode):
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h = 120, 50
    u = O.make_uniforms(sc, len(prims), 11, 0, width=w, height=h, spp=16, max_depth=1)
    got, _, stream, _ = _render(pkg, prims, u, w, h, 2, mode)
    own, _ = O.quantise(got, w, h, mode)
    if stream != own:           # ASCII powf boundary pixels only (same rule as the strict test)
        near = O.quantise_boundary_count(got, w, h, mode, 1e-3)
        assert mode == 0 and len(stream) == len(own)
        assert sum(a != b for a, b in zip(stream, own)) <= near
    # against the strict frame: byte differences only where the floats differ near a quantisation boundary
    ref = O.render(prims, u, conv=0)["pixels"]
    ref_stream, _ = O.quantise(ref, w, h, mode)
    changed = int((got.view(np.uint32) != ref.view(np.uint32)).any(axis=1).sum())
    if len(stream) == len(ref_stream):
        diff = sum(a != b for a, b in zip(stream, ref_stream))
    else:
        diff = abs(len(stream) - len(ref_stream)) + sum(a != b for a, b in zip(stream, ref_stream))
    near = O.quantise_boundary_count(ref, w, h, mode, 1e-3)
    outl = int((_rel_err(got, ref) > REL_TOL).sum())
    print(f"\nmode {mode}: {changed} of {w * h} pixels differ in some bit, {diff} stream bytes differ; "
          f"{near} pixels within 1e-3 of a quantisation boundary, {outl} outlier pixels")
    assert diff <= 12 * (near + outl)        # a pixel is at most 12 digit/glyph bytes of the stream


def test_row_shards_bit_identical_native(pkg, O):
    sc = _scene("scene2.json")
    prims = O.flatten_scene(sc)
    w, h = 50, 23
    u = O.make_uniforms(sc, len(prims), 3, 0, width=w, height=h, spp=8, max_depth=4)
    ref, _, _, _ = _render(pkg, prims, u, w, h, 2)
    again, _, _, _ = _render(pkg, prims, u, w, h, 2)
    assert np.array_equal(ref.view(np.uint32), again.view(np.uint32))       # run-to-run deterministic
    union = np.zeros((h, w, 4), np.float32)
    for s in range(3):
        r = pkg.Renderer(w, h, conv=2, shard_index=s, shard_count=3, stripe_rows=2)
        r.upload_scene(prims)
        r.render_frame(_u(pkg, u), pkg.MODE_NONE)
        rows = r.shard_rows(h)
        union[rows] = r.read_pixels().reshape(h, w, 4)[rows]
        r.close()
    assert np.array_equal(union.reshape(-1, 4).view(np.uint32), ref.view(np.uint32))


@pytest.mark.parametrize("which", ["cornell", "c4"])
def test_conv2_is_deterministic_with_several_lights(pkg, O, which):
    """Two and four lights, a frame larger than the GPU's resident lanes (so pixels are handed out dynamically and warp
    composition differs from run to run and between shards): light-sample sums join a sample's radiance in posting
    order (ordered_late_add, trt_native.cuh), so a conv 2 frame is run-to-run and shard-union BIT-identical."""
    if which == "cornell":
        sc, w, h, spp, depth, seed = _scene("Cornell_Box.json"), 640, 360, 4, 12, 0x00C0FFEE      # 230 400 pixels > 151 552 lanes
    else:
        sc = O.make_c4_scene(n_spheres=200, n_tris=600, width=512, height=320, spp=4, max_depth=6)
        w, h, spp, depth, seed = 512, 320, 4, 6, O.C4_SEED
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=depth)
    a, va, _, sa = _render(pkg, prims, u, w, h, 2)
    b, vb, _, sb = _render(pkg, prims, u, w, h, 2)
    assert int((a.view(np.uint32) != b.view(np.uint32)).any(axis=1).sum()) == 0      # run to run
    assert np.array_equal(va["variance"].view(np.uint32), vb["variance"].view(np.uint32))
    assert int(sa.samples_traced) == int(sb.samples_traced)
    union = np.zeros((h, w, 4), np.float32)
    for s in range(3):
        r = pkg.Renderer(w, h, conv=2, shard_index=s, shard_count=3, stripe_rows=4)
        r.upload_scene(prims)
        r.render_frame(_u(pkg, u), pkg.MODE_NONE)
        rows = r.shard_rows(h)
        union[rows] = r.read_pixels().reshape(h, w, 4)[rows]
        r.close()
    assert int((union.reshape(-1, 4).view(np.uint32) != a.view(np.uint32)).any(axis=1).sum()) == 0   # shard union


def test_large_scene_native(pkg, O):
    """> 256 primitives (geometry from HBM) and mixed plane kinds, single bounce, tolerance 1e-4."""
    sc = O.make_c4_scene(n_spheres=150, n_tris=400, width=96, height=54, spp=16, max_depth=1)
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0)
    ref0, ref1 = O.render(prims, u, conv=0)["pixels"], O.render(prims, u, conv=1)["pixels"]
    got, _, _, _ = _render(pkg, prims, u, 96, 54, 2)
    _assert_within_conventions("C4-reduced", got, ref0, ref1, 96, 54)
    # deeper paths on the same scene: below the noise floor
    sc6 = dict(sc, max_depth=6)
    u6 = O.make_uniforms(sc6, len(prims), O.C4_SEED, 0)
    u6b = O.make_uniforms(sc6, len(prims), O.C4_SEED + 1, 0)
    a, b = O.render(prims, u6)["pixels"], O.render(prims, u6b)["pixels"]
    g6, _, _, _ = _render(pkg, prims, u6, 96, 54, 2)
    assert _rmse(g6, a) <= 0.08 * _rmse(a, b)


def test_large_scene_native_deferred_kernel(pkg, O):
    """> 256 primitives with at most 3 lights: the deferred-shadow kernel reading geometry from HBM
    (trace_native_kernel<SrcGlobal>; the 4-light scene above takes the wave kernel), sphere + triangle lights."""
    sc = O.make_c4_scene(n_spheres=150, n_tris=400, width=96, height=54, spp=16, max_depth=1)
    sc["triangles"][-1]["emission"] = [0.0, 0.0, 0.0]          # second quad dark: 2 triangle lights left
    sc["triangles"][-2]["emission"] = [0.0, 0.0, 0.0]
    sc["spheres"][0].update(center=[0.0, 3.5, -9.0], radius=0.6, emission=[9.0, 9.0, 9.0], color=[0.0, 0.0, 0.0])   # + 1 sphere light
    prims = O.flatten_scene(sc)
    assert len(prims) > 256
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0)
    ref0, ref1 = O.render(prims, u, conv=0)["pixels"], O.render(prims, u, conv=1)["pixels"]
    got, _, _, _ = _render(pkg, prims, u, 96, 54, 2)
    _assert_within_conventions("C4-reduced, 3 lights", got, ref0, ref1, 96, 54)
    sc6 = dict(sc, max_depth=6)
    u6 = O.make_uniforms(sc6, len(prims), O.C4_SEED, 0)
    u6b = O.make_uniforms(sc6, len(prims), O.C4_SEED + 1, 0)
    a, b = O.render(prims, u6)["pixels"], O.render(prims, u6b)["pixels"]
    g6, _, _, _ = _render(pkg, prims, u6, 96, 54, 2)
    print(f"  depth 6: native-vs-strict RMSE {_rmse(g6, a):.4g}, noise floor {_rmse(a, b):.4g}")
    assert _rmse(g6, a) <= 0.08 * _rmse(a, b)


def test_full_c4_scene_cull_keeps_every_hit(pkg, O):
    """All 20 484 primitives, single bounce: the large-scene kernel culls triangles and spheres by a distance test on packed
    pairs before the as-written tests (scan_run / tri_cull, two primitives per step, one branch per four).  A cull that
    dropped a real hit would put sky or a farther primitive into a pixel -- a whole-sample flip.  The yardstick is the
    strict machine itself (no cull there): its two exact conventions against each other, both on the GPU."""
    sc = O.make_c4_scene()
    prims = O.flatten_scene(sc)
    assert len(prims) == 20484
    w, h = 160, 90
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0, width=w, height=h, spp=4, max_depth=1)
    s0, _, _, _ = _render(pkg, prims, u, w, h, 0)
    s1, _, _, _ = _render(pkg, prims, u, w, h, 1)
    r = pkg.Renderer(w, h, conv=2)
    r.upload_scene(prims)
    r.render_frame(_u(pkg, u), pkg.MODE_NONE)
    got, kernel = r.read_pixels(), r.kernel_name()
    r.close()
    assert "SrcTiled" in kernel
    _assert_within_conventions(f"C4 full scene, depth 1, {kernel}", got, s0, s1, w, h)
