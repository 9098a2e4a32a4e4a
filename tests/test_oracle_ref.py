"""Pins the hand-written oracle to the reference's own source text.

oracle/_ref/libref.so is /root/reference/src/shader.wgsl transliterated to C by oracle/wgsl2c.py at build time
(one C statement per WGSL statement) and compiled with -ffp-contract=off: "the reference compiled here".  These
tests show that the hand restatement (oracle/trt_oracle_core.inc, conv 0 -- the thing every GPU parity test
compares against) returns the SAME BITS as that build: pixels, accumulation and variance, on the three shipped
scenes, spp 1...128, depth 0...32, accumulated frames 0...3, and a reduced C4 scene (spheres + triangles +
live NEE).  They also test the transliterator itself on small WGSL fragments.

Where neither /root/reference nor a prebuilt oracle/_ref/libref.so exists the _ref tests are skipped (the GPU
box receives the prebuilt file with the snapshot)."""
import json
import os
import subprocess
import sys
import textwrap

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import wgsl2c  # noqa: E402


@pytest.fixture(scope="module")
def R(O):
    if not O.ref_available():
        pytest.skip("oracle/_ref/libref.so not built and /root/reference absent")
    O.ref_lib()
    return O


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def same(a, b):
    return (np.array_equal(bits(a["pixels"]), bits(b["pixels"]))
            and np.array_equal(bits(a["accumulation"]), bits(b["accumulation"]))
            and np.array_equal(bits(a["variance"]["variance"]), bits(b["variance"]["variance"]))
            and np.array_equal(a["variance"]["sample_count"], b["variance"]["sample_count"]))


def scene(name):
    return json.load(open(os.path.join(SCENES, name)))


# ---- the transliterated shader's own KATs (SURVEY 8(c) pins, now run on the reference's text) ----------------
PCG_KAT = [(0, 0x07BB2FE2), (1, 0xA8BEEA3C), (2, 0x7A7ECC88), (12345, 0xF45EAD0E), (0xFFFFFFFF, 0xE62A4902),
           (1973, 0xFAE9338C)]


def test_ref_pcg_hash(R):
    for v, expect in PCG_KAT:
        assert R.ref_pcg_hash(v) == expect == R.pcg_hash(v)
    rng = np.random.default_rng(7)
    for v in rng.integers(0, 2**32, 3000):
        assert R.ref_pcg_hash(int(v)) == R.pcg_hash(int(v))


@pytest.mark.parametrize("args", [(0, 0, 400, 0, 0), (199, 99, 400, 1, 0), (399, 199, 400, 0xDEADBEEF, 7),
                                  (3839, 2159, 3840, 0x00C0FFEE, 99)])
def test_ref_rng_stream(R, args):
    st_r, va_r = R.ref_rng_stream(*args, 64)
    st_o, va_o = R.rng_stream(*args, 64)
    assert np.array_equal(st_r, st_o) and np.array_equal(bits(va_r), bits(va_o))


def test_ref_random_f32_reaches_one(R):
    # f32(state) / 4294967295.0: the AbstractFloat literal rounds to 2^32 (checked in the generated text below)
    inc = open(os.path.join(ROOT, "oracle", "_ref", "shader_wgsl.inc")).read() \
        if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "shader_wgsl.inc")) else None
    if inc is not None:
        assert "wgsl_f32_from_u32(g_rand_state) / 0x1.0000000000000p+32f" in inc


def test_ref_intersections_match_oracle(R):
    """hit_sphere / hit_plane / hit_triangle of the transliterated text vs the hand restatement on random rays."""
    rng = np.random.default_rng(11)
    prims = R.flatten_scene(scene("Cornell_Box.json"))
    extra = R.flatten_scene(scene("demo.json"))
    n_bad = 0
    for p in list(prims) + list(extra[:6]):
        for _ in range(300):
            o = rng.uniform(-1, 1, 3).astype(np.float32) * np.float32(0.9) + np.array([0, 0, -2], np.float32)
            d = rng.normal(size=3).astype(np.float32)
            d = (d / np.float32(np.sqrt(np.float32(d @ d)))).astype(np.float32)
            tmax = np.float32(rng.choice([1e10, 3.0, 0.7]))
            got = R.ref_hit(p, o, d, 0.001, float(tmax))
            if p["type"] == 0:
                want = R.hit_sphere(o, d, [p["sphere_center"][k] for k in "xyz"], float(p["sphere_radius"]), 0.001, float(tmax))
            elif p["type"] == 1:
                want = R.hit_plane(o, d, [p["plane_point"][k] for k in "xyz"], [p["plane_normal"][k] for k in "xyz"], 0.001, float(tmax))
            else:
                want = R.hit_triangle(o, d, *[[p[v][k] for k in "xyz"] for v in ("v0", "v1", "v2")], 0.001, float(tmax))
            n_bad += int(np.float32(got).view(np.uint32) != np.float32(want).view(np.uint32))
    assert n_bad == 0


# ---- frames: the hand oracle == the transliterated reference, bit for bit ------------------------------------
FRAME_CASES = [  # scene file, w, h, spp, depth, seed
    ("Cornell_Box.json", 64, 32, 16, 32, 0x00C0FFEE),      # C1's parameters on a small frame
    ("Cornell_Box.json", 40, 20, 128, 3, 5),               # the file's own spp / depth
    ("Cornell_Box.json", 33, 17, 1, 1, 9),                 # spp < 4: base = 4 > spp
    ("Cornell_Box.json", 24, 12, 4, 0, 3),                 # depth 0: no bounce at all
    ("Cornell_Box.json", 16, 8, 128, 32, 77),              # adaptive extra samples at depth 32
    ("scene2.json", 50, 25, 32, 4, 1),
    ("scene2.json", 31, 19, 2, 2, 12345),
    ("scene2.json", 20, 10, 64, 32, 0xFFFFFFFF),
    ("demo.json", 40, 20, 16, 32, 2),
    ("demo.json", 25, 13, 8, 1, 0xDEADBEEF),
    ("demo.json", 12, 6, 128, 8, 4),
]


@pytest.mark.parametrize("case", FRAME_CASES, ids=lambda c: f"{c[0][:-5]}-{c[1]}x{c[2]}-spp{c[3]}-d{c[4]}")
def test_frame_bits_equal_reference_build(R, case):
    name, w, h, spp, depth, seed = case
    sc = scene(name)
    prims = R.flatten_scene(sc)
    u = R.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=depth)
    assert same(R.render(prims, u), R.ref_render(prims, u))


@pytest.mark.parametrize("name", ["Cornell_Box.json", "scene2.json", "demo.json"])
def test_accumulated_frames_bits_equal_reference_build(R, name):
    """frames 0...3 with the C2 seed schedule: the temporal blend shader.wgsl:610-621 on both sides."""
    sc = scene(name)
    prims = R.flatten_scene(sc)
    acc_o = acc_r = None
    for f in range(4):
        u = R.make_uniforms(sc, len(prims), R.c2_seed(f), f, width=36, height=18, spp=16, max_depth=8)
        a = R.render(prims, u, accumulation=acc_o)
        b = R.ref_render(prims, u, accumulation=acc_r)
        assert same(a, b), f"frame {f}"
        acc_o, acc_r = a["accumulation"], b["accumulation"]


def test_camera_pose_and_odd_sizes(R):
    sc = scene("scene2.json")
    prims = R.flatten_scene(sc)
    cam = R.Camera()
    for k in ("w", "Right", "d", "Up", "Up", "a", "Left", "s"):
        cam.key(k)
    for (w, h) in ((2, 2), (9, 5), (17, 3), (8, 8)):
        u = R.make_uniforms(sc, len(prims), 42, 0, camera=cam, width=w, height=h, spp=8, max_depth=4)
        assert same(R.render(prims, u), R.ref_render(prims, u))


def test_c4_reduced_scene_bits_equal_reference_build(R):
    """Spheres + triangles + 4 live emissive triangles (NEE contributes, unlike Cornell: SURVEY Q7)."""
    sc = R.make_c4_scene(n_spheres=96, n_tris=384, width=36, height=20, spp=16, max_depth=32)
    prims = R.flatten_scene(sc)
    assert len(prims) == 96 + 384 + 4
    u = R.make_uniforms(sc, len(prims), R.C4_SEED, 0)
    a, b = R.render(prims, u), R.ref_render(prims, u)
    assert same(a, b)
    assert float(a["pixels"][:, :3].max()) > 0.0


def test_idx_and_row_modes_agree(R):
    """The pixel-subset entry (what bench.py uses on C3) returns the same bits as the full dispatch."""
    sc = scene("Cornell_Box.json")
    prims = R.flatten_scene(sc)
    u = R.make_uniforms(sc, len(prims), 1, 0, width=48, height=24, spp=16, max_depth=16)
    full = R.ref_render(prims, u)
    idx = np.arange(5, 48 * 24, 37, dtype=np.uint32)
    sub = R.ref_render(prims, u, idx=idx)
    assert np.array_equal(bits(sub["pixels"]), bits(full["pixels"][idx]))
    assert np.array_equal(bits(sub["variance"]["variance"]), bits(full["variance"]["variance"][idx]))
    rows = R.ref_render(prims, u, rows=(8, 16))
    assert np.array_equal(bits(rows["pixels"][8 * 48:16 * 48]), bits(full["pixels"][8 * 48:16 * 48]))
    assert not rows["pixels"][:8 * 48].any() and not rows["pixels"][16 * 48:].any()
    hand = R.render(prims, u, idx=idx)
    assert np.array_equal(bits(hand["pixels"]), bits(sub["pixels"]))


def test_glibc_sincos_mode_also_equal(R):
    sc = scene("demo.json")
    prims = R.flatten_scene(sc)
    u = R.make_uniforms(sc, len(prims), 3, 0, width=30, height=16, spp=16, max_depth=16)
    assert same(R.render(prims, u, math_mode=1), R.ref_render(prims, u, math_mode=1))


def test_fast_build_of_the_oracle_has_the_same_bits(O):
    """liboracle_fast.so (-O3 -march=native, contraction still off) is what the TIMED cpu_baseline legs run."""
    sc = scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    for conv in (0, 1):
        u = O.make_uniforms(sc, len(prims), 5, 0, width=40, height=20, spp=32, max_depth=32)
        assert same(O.render(prims, u, conv=conv), O.render(prims, u, conv=conv, fast=True))


def test_contracted_reference_build_is_the_yardstick(R):
    """_ref/libref_fma.so = the same text compiled with FMA contraction allowed (legal for WGSL).  Its distance
    from the strict build is what 'a different legal rounding of the reference' looks like; the conv 2 kernels
    are held to a multiple of it (tests/test_gpu_native.py)."""
    sc = scene("Cornell_Box.json")
    prims = R.flatten_scene(sc)
    u = R.make_uniforms(sc, len(prims), R.C1_SEED, 0, width=80, height=40, spp=16, max_depth=1)
    a = R.ref_render(prims, u)["pixels"][:, :3]
    b = R.ref_render(prims, u, fma=True)["pixels"][:, :3]
    rel = np.abs(a - b) / np.maximum(np.abs(a), 1e-2)
    over = int((rel.max(axis=1) > 1e-4).sum())
    assert not np.array_equal(bits(a), bits(b))         # it IS a different rounding
    assert over <= 0.02 * len(a)                        # and stays a rounding-noise distance away


# ---- the transliterator on small fragments -----------------------------------------------------------------
def tr(src):
    probe = wgsl2c.Translator(textwrap.dedent(src), "<test>")
    probe.translate()
    return wgsl2c.Translator(textwrap.dedent(src), "<test>", impure=probe.effects()).translate()


def test_wgsl2c_abstract_constants_fold_in_f64_and_round_once():
    c = tr("""
        fn f(r: f32) -> f32 { return 2.0 * 3.14159265359 * r; }
        fn g(s: u32) -> f32 { return f32(s) / 4294967295.0; }
        fn h() -> f32 { let t = -1.0; var n = 1; return t; }
    """)
    two_pi = np.float32(2.0 * 3.14159265359)
    assert f"({float(two_pi).hex()}f" in c
    assert "/ 0x1.0000000000000p+32f" in c
    assert "const float t = -0x1.0000000000000p+0f" in c and "int32_t n = 1;" in c


def test_wgsl2c_hoists_calls_with_side_effects_in_source_order():
    c = tr("""
        var<private> s: u32;
        fn next() -> f32 { s = s + 1u; return f32(s); }
        fn pure(a: f32, b: f32) -> f32 { return a - b; }
        fn use2() -> f32 { return pure(next(), next()); }
    """)
    body = c[c.index("static float w_use2(void) {"):]
    i1, i2, i3 = body.index("_w1 = w_next()"), body.index("_w2 = w_next()"), body.index("w_pure(_w1, _w2)")
    assert i1 < i2 < i3


def test_wgsl2c_rejects_what_it_does_not_understand():
    for bad in ("fn f() -> f32 { return pow(2.0, 3.0); }",                 # unsupported builtin
                "fn f(a: f32, b: u32) -> f32 { return a + b; }",           # WGSL has no implicit conversion
                "var<private> s: u32;\nfn n() -> bool { s = 1u; return true; }\n"
                "fn f(a: bool) -> bool { return a && n(); }",              # effect behind a short circuit
                "fn f() { let v = vec3<f32>(1.0, 2.0, 3.0); }"):
        with pytest.raises(wgsl2c.WgslError):
            tr(bad)


def test_wgsl2c_output_is_not_committed():
    """The generated C is derived from reference text: it must stay out of the history."""
    tracked = subprocess.run(["git", "ls-files", "oracle/_ref"], cwd=ROOT, capture_output=True, text=True)
    if tracked.returncode == 0:
        assert tracked.stdout.strip() == ""
    gi = open(os.path.join(ROOT, ".gitignore")).read()
    assert "oracle/_ref/" in gi
