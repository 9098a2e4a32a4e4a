"""Host logic of the drop-in (no GPU): the native scene parser / camera / uniforms in libtrt_b200.so
against the numpy restatement of the reference's host code in oracle/oracle.py, the C ABI surface,
and the no-CPU-fallback rule."""
import json
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(os.path.dirname(__file__), "golden", "scenes")


def _load(name):
    with open(os.path.join(SCENES, name)) as f:
        return f.read()


# ---- C ABI surface -------------------------------------------------------------------------------------
def test_library_exports_every_declared_symbol(pkg):
    hdr = open(os.path.join(ROOT, "include", "trt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(trt_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 35
    lib = pkg.api.lib()
    missing = [n for n in sorted(declared) if not hasattr(lib, n)]
    assert not missing, f"not exported: {missing}"
    # and the ctypes table covers the header, so tests and bench bind the real signatures
    assert declared == set(pkg.api.SYMBOLS), declared ^ set(pkg.api.SYMBOLS)


def test_rust_binding_declares_every_symbol():
    """bindings/rust/trt-b200-sys (source only: no Rust toolchain in this image) must declare every entry point
    of the header, with the same number of parameters."""
    hdr = open(os.path.join(ROOT, "include", "trt_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    rs = open(os.path.join(ROOT, "bindings", "rust", "trt-b200-sys", "src", "lib.rs")).read()
    rs = re.sub(r"//.*", "", rs)
    ext = rs[rs.index('extern "C" {'):]
    ext = ext[:ext.index("\n}\n")]
    rust = {m.group(1): m.group(2) for m in re.finditer(r"pub fn (trt_[a-z0-9_]+)\s*\((.*?)\)\s*(?:->[^;]*)?;", ext, flags=re.S)}
    cdecl = {m.group(1): m.group(2) for m in re.finditer(r"\b(trt_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", hdr)}
    assert set(cdecl) == set(rust), set(cdecl) ^ set(rust)

    def arity(params):
        params = params.strip()
        return 0 if params in ("", "void") else params.count(",") + 1

    bad = [n for n in cdecl if arity(cdecl[n]) != arity(rust[n])]
    assert not bad, bad


def test_wire_struct_sizes(pkg):
    import ctypes as C
    a = pkg.api
    assert C.sizeof(a.Vec3) == 16 and C.sizeof(a.Primitive) == 180 and C.sizeof(a.Uniforms) == 160
    assert a.Uniforms.grid_min.offset == 112                       # shader-visible prefix (shader.wgsl:49-69)
    assert C.sizeof(a.Counters) == 19 * 8 and C.sizeof(a.Config) == 64


def test_no_cpu_fallback(pkg):
    """Without a CUDA device the product must refuse, not compute on the CPU."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.TrtError) as e:
        pkg.Renderer(16, 16)
    assert e.value.code == -2 and "no CPU fallback" in str(e.value)


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/."""
    pkg_dir = os.path.join(ROOT, "terminal-raytracer_b200")
    for dirpath, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")) or f == "Makefile":
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "import oracle" not in text and "trt_oracle" not in text, f


# ---- scene schema (lib.rs:52-98) and flattening order (lib.rs:120-154) ------------------------------------------
@pytest.mark.parametrize("name", ["Cornell_Box.json", "scene2.json", "demo.json"])
def test_scene_primitives_match_numpy_restatement(pkg, O, name):
    text = _load(name)
    s = pkg.Scene.parse(text)
    ref = O.flatten_scene(json.loads(text))
    assert np.array_equal(s.primitives(), ref.view(np.uint8).reshape(-1, 180))
    i, j = s.info, json.loads(text)
    assert (i.width, i.height, i.samples_per_pixel, i.max_depth, i.frames_to_accumulate) == (
        j["width"], j["height"], j["samples_per_pixel"], j["max_depth"], j["frames_to_accumulate"])
    assert (i.n_spheres, i.n_planes, i.n_triangles) == (len(j["spheres"]), len(j["planes"]), len(j.get("triangles", [])))


def test_embedded_default_scene_is_cornell(pkg, O):
    """lib.rs:107: no path -> the embedded Cornell box."""
    assert np.array_equal(pkg.Scene.load(None).primitives(), O.flatten_scene(O.CORNELL).view(np.uint8).reshape(-1, 180))


def test_scene_schema_rules(pkg):
    base = json.loads(_load("scene2.json"))
    # triangles is #[serde(default)]; unknown keys are ignored; numbers are f64 -> f32 casts
    d = dict(base); d.pop("triangles"); d["unknown_key"] = {"x": [1, 2, 3]}
    s = pkg.Scene.parse(json.dumps(d))
    assert s.info.n_triangles == 0 and len(s.primitives()) == 3
    # spheres and planes are mandatory (may be empty)
    for key in ("spheres", "planes", "width", "camera", "samples_per_pixel"):
        d = dict(base); d.pop(key)
        with pytest.raises(pkg.TrtError) as e:
            pkg.Scene.parse(json.dumps(d))
        assert e.value.code == -6 and key in str(e.value)
    d = dict(base); d["spheres"] = []; d["planes"] = []
    assert len(pkg.Scene.parse(json.dumps(d)).primitives()) == 1
    for bad in ("", "{", '{"width": 1,}', "[1, 2]", '{"width": -3}'):
        with pytest.raises(pkg.TrtError):
            pkg.Scene.parse(bad)
    with pytest.raises(pkg.TrtError) as e:
        pkg.Scene.load("/nonexistent/scene.json")
    assert e.value.code == -7


def test_f64_to_f32_cast(pkg):
    d = json.loads(_load("scene2.json"))
    d["spheres"][0]["center"] = [0.1, 1e-50, 16777217.0]
    d["spheres"][0]["radius"] = 0.30000000000000004
    p = pkg.Scene.parse(json.dumps(d)).primitives()[0].view(np.float32)
    assert p[16] == np.float32(0.1) and p[17] == 0.0 and p[18] == np.float32(16777216.0) and p[20] == np.float32(0.3)


def test_terminal_clamp(pkg):
    """lib.rs:114-115: width = min(cols, w), height = min(rows - 2, h)."""
    s = pkg.Scene.load(None)
    s.clamp(120, 40)
    assert (s.info.width, s.info.height) == (120, 38)
    s.clamp(500, 500)
    assert (s.info.width, s.info.height) == (120, 38)


# ---- camera (camera.rs:17-25) and keys (lib.rs:392-405) ---------------------------------------------------------------
def test_camera_and_keys_match_oracle(pkg, O):
    a, b = pkg.Camera(), O.Camera()
    keys = (O.C5_KEYS * 13) + ["Up"] * 40 + ["Down"] * 75 + ["w", "a", "Left", "s", "d", "Right"] * 5
    for k in keys:
        assert a.key(k) == b.key(k)
        assert np.array_equal(a.position.view(np.uint32), b.position.view(np.uint32)), k
        for va, vb in zip(a.vectors(), b.vectors()):
            assert np.array_equal(va.view(np.uint32), vb.view(np.uint32)), k
    assert float(a.c.pitch) == -1.5                      # clamp(-1.5, 1.5), lib.rs:405
    assert not a.key(ord("x"))                           # other keys do not move the camera


@pytest.mark.parametrize("name", ["Cornell_Box.json", "scene2.json"])
def test_uniforms_match_oracle(pkg, O, name):
    text = _load(name)
    s, sc = pkg.Scene.parse(text), json.loads(text)
    a, b = pkg.Camera(), O.Camera()
    for step, k in enumerate(["w", "Right", "d", "Up", None, "s"]):
        if k:
            a.key(k); b.key(k)
        u = s.uniforms(a, 0x1234 + step, step)
        ref = O.make_uniforms(sc, len(s.primitives()), 0x1234 + step, step, camera=b)
        assert bytes(u) == ref.tobytes()
    s.set_size(123, 45)
    u = s.uniforms(a, 1, 0)
    assert (u.width, u.height) == (123, 45) and u.aspect_ratio == np.float32(123) / np.float32(45)


# ---- row sharding (multi-GPU layout) ------------------------------------------------------------------------------------
@pytest.mark.parametrize("h,n,stripe", [(2160, 8, 4), (200, 3, 2), (7, 4, 1), (23, 2, 8), (5, 8, 1), (16, 1, 4)])
def test_shard_rows_partition(pkg, h, n, stripe):
    from terminal_raytracer_b200 import multigpu
    seen = np.zeros(h, int)
    for r in range(n):
        rows = pkg.api.shard_rows_of(r, n, stripe, h)
        assert np.array_equal(rows, multigpu.shard_rows(h, r, n, stripe))      # C and Python statements agree
        assert (np.diff(rows.astype(int)) > 0).all()
        seen[rows] += 1
    assert (seen == 1).all()
    assert multigpu.max_rows_per_shard(h, n, stripe) == max(len(pkg.api.shard_rows_of(r, n, stripe, h)) for r in range(n))
