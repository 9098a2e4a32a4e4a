"""Known-answer tests that pin the CPU oracle to the reference's source text.

The reference has no tests or golden vectors ("parity unpinned"); these values were
derived by hand from /root/reference/src/shader.wgsl, src/lib.rs, src/camera.rs and are
re-derived here independently of the oracle's C code wherever possible (pure-Python integer PCG, numpy
float32 formulas)."""
import ctypes
import math

import numpy as np
import pytest


# ---- RNG (shader.wgsl:104-115, :507, :517) -------------------------------------------------------------
def py_pcg(v):
    state = (v * 747796405 + 2891336453) & 0xFFFFFFFF
    word = (((state >> ((state >> 28) + 4)) ^ state) * 277803737) & 0xFFFFFFFF
    return ((word >> 22) ^ word) & 0xFFFFFFFF


PCG_KAT = [(0, 0x07BB2FE2), (1, 0xA8BEEA3C), (2, 0x7A7ECC88), (12345, 0xF45EAD0E), (0xFFFFFFFF, 0xE62A4902), (1973, 0xFAE9338C)]


@pytest.mark.parametrize("v,expect", PCG_KAT)
def test_pcg_hash_kat(O, v, expect):
    assert py_pcg(v) == expect          # the hand derivation itself
    assert O.pcg_hash(v) == expect


def test_pcg_hash_random_inputs_match_python(O):
    rng = np.random.default_rng(1)
    for v in rng.integers(0, 2**32, 2000):
        assert O.pcg_hash(int(v)) == py_pcg(int(v))


RNG_STREAMS = [
    ((0, 0, 400, 0, 0), 0x00000000, [0x30BE035E, 0x7FDDB461, 0x8D324821, 0x36DFF3B3]),
    ((199, 99, 400, 1, 0), 0x04AE5060, [0x81C2089A, 0xAFB9CF32, 0xBFD16465, 0x9D4F6CC9]),
    ((399, 199, 400, 0xDEADBEEF, 7), 0x87A9D74D, [0x6D21617F, 0x1DF5CE17, 0x29FEC56A, 0x2E7F71BF]),
]


@pytest.mark.parametrize("args,seed0,states", RNG_STREAMS)
def test_rng_stream_after_main_seeding(O, args, seed0, states):
    x, y, w, seed, frame = args
    assert ((y * w + x) * 1973 + seed * 9277 + frame * 12345) & 0xFFFFFFFF == seed0      # shader.wgsl:507
    st, va = O.rng_stream(x, y, w, seed, frame, 4)
    assert [int(s) for s in st] == states
    # random_f32 = f32(state) / 4294967295.0, whose literal is 2^32 in f32 (shader.wgsl:114)
    expect = (st.astype(np.float32) / np.float32(4294967296.0)).astype(np.float32)
    assert np.array_equal(va.view(np.uint32), expect.view(np.uint32))


def test_rng_first_floats_of_pixel_0(O):
    _, va = O.rng_stream(0, 0, 400, 0, 0, 4)
    assert [f"{v:.8f}" for v in va] == ["0.19039936", "0.49947670", "0.55154848", "0.21435474"]


def test_random_f32_reaches_exactly_one():
    # state >= 0xFFFFFF80 rounds up to 2^32 in f32 -> 1.0 exactly
    assert np.float32(np.uint32(0xFFFFFF80)) / np.float32(4294967296.0) == np.float32(1.0)
    assert np.float32(np.uint32(0xFFFFFF7F)) / np.float32(4294967296.0) < np.float32(1.0)


# ---- camera / uniforms (camera.rs:17-25, lib.rs:118, lib.rs:427-429) --------------------------------------
def test_default_camera_vectors(O):
    f, r, u = O.Camera().vectors()
    assert f[1] == 0.0 and f[2] == -1.0 and abs(f[0] - (-4.3711388e-8)) < 1e-14
    assert r[0] == 1.0 and r[1] == 0.0 and abs(r[2] - (-4.3711388e-8)) < 1e-14
    assert abs(u[0]) < 1e-12 and u[1] == 1.0 and abs(u[2]) < 1e-12


def test_fov_to_radians_and_tan(O):
    fov = O.to_radians_f32(45.0)
    assert f"{fov:.8f}" == "0.78539819"
    assert f"{O.lib().oracle_tan_half(ctypes.c_float(fov)):.8f}" == "0.41421357"


def test_struct_sizes(O):
    assert O.VEC3.itemsize == 16 and O.PRIM.itemsize == 180 and O.UNIFORMS.itemsize == 160 and O.VARIANCE.itemsize == 8
    # the shader-visible prefix of Uniforms ends after camera_up (shader.wgsl:49-69) = 112 bytes
    assert O.UNIFORMS.fields["grid_min"][1] == 112


def test_cornell_light_triangles_face_up(O):
    """normalize(cross(e1, e2)) = (0, +1, 0), area 0.32 for both light triangles."""
    prims = O.flatten_scene(O.CORNELL)
    lights = [p for p in prims if p["type"] == 2]
    assert len(lights) == 2
    for p in lights:
        v = [np.array([p[k]["x"], p[k]["y"], p[k]["z"]], np.float64) for k in ("v0", "v1", "v2")]
        c = np.cross(v[1] - v[0], v[2] - v[0])
        assert np.allclose(c / np.linalg.norm(c), [0, 1, 0]) and abs(0.5 * np.linalg.norm(c) - 0.32) < 1e-6


@pytest.mark.parametrize("spp,base", [(1, 4), (4, 4), (16, 4), (32, 8), (128, 32), (256, 64), (1024, 256)])
def test_base_samples(O, spp, base):
    """shader.wgsl:513 base = max(4, spp/4); the variance buffer records it (:608)."""
    prims = O.flatten_scene(O.SCENE2)
    u = O.make_uniforms(O.SCENE2, len(prims), 1, 0, width=4, height=2, spp=spp, max_depth=1)
    r = O.render(prims, u)
    assert (r["variance"]["sample_count"] == base).all()
    assert r["samples_traced"] >= base * 8


# ---- intersections on hand-computable rays (shader.wgsl:155-223) ----------------------------------------------
@pytest.mark.parametrize("conv", [0, 1])
def test_intersection_kats(O, conv):
    o, d = [0, 0, 0], [0, 0, -1]
    assert O.hit_sphere(o, d, [0, 0, -3], 1.0, 0.001, 1e10, conv) == np.float32(2.0)          # near root
    assert O.hit_sphere([0, 0, -3], d, [0, 0, -3], 1.0, 0.001, 1e10, conv) == np.float32(1.0)   # inside: far root
    assert O.hit_sphere(o, d, [0, 0, -3], 1.0, 0.001, 2.0, conv) == np.float32(-1.0)           # t_max <= root (exclusive)
    assert O.hit_sphere(o, d, [5, 0, -3], 1.0, 0.001, 1e10, conv) == np.float32(-1.0)          # miss
    assert O.hit_plane(o, [0, -1, 0], [0, -1, 0], [0, 1, 0], 0.001, 1e10, conv) == np.float32(1.0)
    assert O.hit_plane(o, [0, -1, 0], [0, -1, 0], [0, 2, 0], 0.001, 1e10, conv) == np.float32(1.0)   # un-normalised normal
    assert O.hit_plane(o, [0, -1, 0], [0, -1, 0], [0, 1, 0], 0.001, 1.0, conv) == np.float32(1.0)    # t == t_max kept (inclusive)
    assert O.hit_plane(o, [1, 0, 0], [0, -1, 0], [0, 1, 0], 0.001, 1e10, conv) == np.float32(-1.0)   # parallel
    assert O.hit_plane(o, [0, 1, 0], [0, -1, 0], [0, 1, 0], 0.001, 1e10, conv) == np.float32(-1.0)   # behind
    tri = ([-1, -1, -2], [1, -1, -2], [0, 1, -2])
    assert O.hit_triangle(o, d, *tri, 0.001, 1e10, conv) == np.float32(2.0)
    assert O.hit_triangle(o, [0, 0, 1], *tri, 0.001, 1e10, conv) == np.float32(-1.0)
    assert O.hit_triangle([3, 0, 0], d, *tri, 0.001, 1e10, conv) == np.float32(-1.0)
    assert O.hit_triangle(o, d, tri[0], tri[2], tri[1], 0.001, 1e10, conv) == np.float32(2.0)        # two-sided


# ---- pinned sin/cos against glibc (the stated arithmetic convention) ------------------------------------------
def test_pinned_sincos_within_2ulp_of_libm(O):
    libm = ctypes.CDLL("libm.so.6")
    libm.sinf.restype = libm.cosf.restype = ctypes.c_float
    libm.sinf.argtypes = libm.cosf.argtypes = [ctypes.c_float]
    rng = np.random.default_rng(7)
    phis = np.concatenate([rng.random(20000, dtype=np.float32) * np.float32(6.2831855),
                           np.array([0.0, 6.2831855, 1.5707964, 3.1415927, 4.712389, 0.7853982, 1e-6], np.float32)])
    worst = 0.0
    for phi in phis:
        s, c = O.sincos_pinned(float(phi))
        for got, ref in ((s, libm.sinf(float(phi))), (c, libm.cosf(float(phi)))):
            ulp = np.spacing(np.float32(max(abs(ref), 2.0**-20)))      # below 2^-20 compare absolutely
            worst = max(worst, abs(float(got) - ref) / float(ulp))
    assert worst <= 2.0, worst


# ---- conventions agree within the north_star tolerance (single bounce, 1e-4 relative) -----------------------------
@pytest.mark.parametrize("scene_name", ["CORNELL", "SCENE2"])
def test_single_bounce_conventions_within_1e4(O, scene_name):
    sc = getattr(O, scene_name)
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), 9, 0, width=64, height=32, spp=16, max_depth=1)
    base = O.render(prims, u, conv=0, math_mode=0)["pixels"][:, :3]
    for conv, mm in ((1, 0), (0, 1)):
        other = O.render(prims, u, conv=conv, math_mode=mm)["pixels"][:, :3]
        rel = np.abs(other - base) / np.maximum(np.abs(base), 1e-6)
        # a sample that flips hit/miss, lit/unlit at a silhouette, shadow edge or horizon is allowed on a
        # handful of pixels (0 vs non-0 contributions), counted here: at most 0.5 % of the frame
        n_edge = int((rel.max(axis=1) > 1e-4).sum())
        assert n_edge <= len(base) // 200, f"conv {conv} math {mm}: {n_edge} pixels beyond 1e-4"


# ---- quantiser byte grammar (lib.rs:509-527) -------------------------------------------------------------------------
def _px(*rgb):
    return np.array([[rgb[0], rgb[1], rgb[2], 0.0]], np.float32)


def test_quantiser_grammar(O):
    assert O.quantise(_px(1, 1, 1), 1, 1, 1)[0] == b"\x1b[38;2;255;255;255m\xe2\x96\x88\x1b[0m\r\n"
    assert O.quantise(_px(0, 0, 0), 1, 1, 1)[0] == b"\x1b[38;2;0;0;0m\xe2\x96\x88\x1b[0m\r\n"
    assert O.quantise(_px(0.25, 0.0625, 4.0), 1, 1, 1)[0] == b"\x1b[38;2;127;63;255m\xe2\x96\x88\x1b[0m\r\n"   # sqrt, truncation, clamp
    assert O.quantise(_px(1, 1, 1), 1, 1, 0)[0] == b"\x1b[38;2;255;255;255m$\x1b[0m\r\n"    # brightest glyph
    assert O.quantise(_px(0, 0, 0), 1, 1, 0)[0] == b"\x1b[38;2;0;0;0m \x1b[0m\r\n"          # darkest glyph = space
    # NaN: channels -> 0 (clamp keeps NaN, `as u8` -> 0) but the glyph index -> 67 ('$'): f32::min(NaN, 67) = 67
    assert O.quantise(_px(np.nan, np.nan, np.nan), 1, 1, 0)[0] == b"\x1b[38;2;0;0;0m$\x1b[0m\r\n"
    assert O.quantise(_px(-1, -1, -1), 1, 1, 1)[0] == b"\x1b[38;2;0;0;0m\xe2\x96\x88\x1b[0m\r\n"   # sqrt(-1) = NaN -> 0


def test_quantiser_ramp_and_lengths(O):
    ramp = " .`^\",:;Il!i><~+_-?][}{1)(|\\tfjrxnuvczXYUJCLQ0OZmwqpdbkhao*#MW&8%B@$"
    assert len(ramp) == 68
    # every glyph index is reachable and maps through x^0.3 * 67 (lib.rs:522)
    seen = set()
    for k in range(68):
        v = np.float32(((k + 0.5) / 67.0) ** (1 / 0.3)) if k < 67 else np.float32(1.0)
        out = O.quantise(_px(v, v, v), 1, 1, 0)[0]
        seen.add(chr(out[out.index(b"m") + 1]))
    assert seen == set(ramp)
    # 18..24 bytes per ASCII pixel, 20..26 per truecolor pixel, + "\r\n" per row
    rng = np.random.default_rng(3)
    px = rng.random((6 * 5, 4), dtype=np.float32) * 1.5
    for mode, lo, hi in ((0, 18, 24), (1, 20, 26)):
        data, rows = O.quantise(px, 6, 5, mode)
        assert len(data) == int(rows.sum()) and data.count(b"\r\n") == 5
        assert all(6 * lo + 2 <= r <= 6 * hi + 2 for r in rows)
