"""GPU tests of the run-time specialised trace kernel (trt_jit.cuh): a canonical small scene uploaded to a conv 2
context gets a kernel compiled by NVRTC for its shape (the counterpart of the reference's create_shader_module +
create_compute_pipeline, lib.rs:321-352).  Checked here: which kernel runs for which scene, that the specialised
kernel tracks the oracle within conv 2's tolerance (the tolerance tests of test_gpu_native.py run scene2 through it as
well), and that it equals the statically compiled run-time-count kernel (a subprocess with TRT_B200_JIT=0) to
rounding noise."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCENES = os.path.join(ROOT, "tests", "golden", "scenes")


def _u(pkg, u_np):
    return pkg.api.uniforms_from_bytes(u_np.tobytes())


def _render(pkg, O, sc, w, h, conv, spp=8, depth=4, seed=7):
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=depth)
    r = pkg.Renderer(w, h, conv=conv)
    r.upload_scene(prims)
    r.render_frame(_u(pkg, u), pkg.MODE_NONE)
    px, name = r.read_pixels().copy(), r.kernel_name()
    r.close()
    return px, name, prims, u


def test_which_kernel_runs(pkg, O):
    scene2 = json.load(open(os.path.join(SCENES, "scene2.json")))
    cornell = json.load(open(os.path.join(SCENES, "Cornell_Box.json")))
    _, name, _, _ = _render(pkg, O, scene2, 40, 20, 2)
    if os.environ.get("TRT_B200_JIT") in ("0", "off"):
        assert name == "trace_native_kernel<SrcConst, ShapeAny>"
    else:
        assert name.startswith("trt_jit_trace<SrcConst, ShapeFixed<"), name
        sph, pl, tri = len(scene2["spheres"]), len(scene2["planes"]), len(scene2["triangles"])
        assert f"ShapeFixed<{sph}, {pl}, {tri}, " in name
    _, name, _, _ = _render(pkg, O, cornell, 40, 20, 2)
    assert name in ("trace_native_kernel<SrcConst, ShapeCornell>",) or name.startswith("trt_jit_trace<SrcConst, ShapeFixed<3, 6, 2, 0xa05u, 2, 0x0u>")
    _, name, _, _ = _render(pkg, O, cornell, 40, 20, 0)
    assert name == "trace_native_kernel<CONV, SrcConst, ShapeCornell> (strict)"      # the same machine, strict arithmetic
    big = O.make_c4_scene(n_spheres=200, n_tris=300, width=40, height=20, spp=8, max_depth=2)
    _, name, _, _ = _render(pkg, O, big, 40, 20, 2, depth=2)
    assert name == "trace_native_kernel<SrcTiled, ShapeAny>"


def test_specialised_kernel_within_tolerance_of_the_oracle(pkg, O):
    """a shape no static build exists for: 2 spheres (one a light), 3 planes of mixed kinds, 1 triangle light"""
    sc = {"width": 64, "height": 32, "samples_per_pixel": 16, "max_depth": 1, "frames_to_accumulate": 1,
          "camera": {"fov_degrees": 55.0, "char_aspect_ratio": 0.55},
          "spheres": [{"center": [0.6, -0.3, -3.0], "radius": 0.7, "color": [0.8, 0.4, 0.3], "emission": [0, 0, 0], "reflectivity": 0.3},
                      {"center": [-1.2, 1.4, -3.5], "radius": 0.3, "color": [1, 1, 1], "emission": [7, 7, 6], "reflectivity": 0.0}],
          "planes": [{"point": [0, -1, 0], "normal": [0, 1, 0], "color": [0.6, 0.6, 0.6], "emission": [0, 0, 0], "reflectivity": 0.1},
                     {"point": [0, 0, -6], "normal": [0.1, 0.0, 1.0], "color": [0.4, 0.5, 0.8], "emission": [0, 0, 0], "reflectivity": 0.0},
                     {"point": [-2.5, 0, 0], "normal": [1, 0, 0], "color": [0.8, 0.3, 0.3], "emission": [0, 0, 0], "reflectivity": 0.0}],
          "triangles": [{"v0": [0.5, 2.0, -2.5], "v1": [1.5, 2.0, -2.5], "v2": [0.5, 2.0, -3.5], "color": [1, 1, 1], "emission": [9, 9, 9], "reflectivity": 0.0}]}
    got, name, prims, u = _render(pkg, O, sc, 64, 32, 2, spp=16, depth=1)
    assert name.startswith("trt_jit_trace") or os.environ.get("TRT_B200_JIT") in ("0", "off")
    refs = [O.render(prims, u, conv=c)["pixels"].reshape(-1, 4)[:, :3] for c in (0, 1)]
    g = got.reshape(-1, 4)[:, :3]

    def bad(a, b):
        return int(((np.abs(a - b) / np.maximum(np.abs(b), 1e-2)).max(axis=1) > 1e-4).sum())

    yard = bad(refs[1], refs[0])
    assert min(bad(g, refs[0]), bad(g, refs[1])) <= max(1, 2 * yard + int(0.002 * 64 * 32)), (bad(g, refs[0]), yard)


def test_specialised_equals_runtime_count_build(pkg, O, tmp_path):
    scene2 = json.load(open(os.path.join(SCENES, "scene2.json")))
    got, name, _, _ = _render(pkg, O, scene2, 60, 30, 2, spp=16, depth=4)
    out = str(tmp_path / "px.npy")
    code = ("import sys, json, os, numpy as np; sys.path.insert(0, %r); import __graft_entry__ as g; pkg, O = g.load_package(), g.load_oracle();"
            "sc = json.load(open(%r)); prims = O.flatten_scene(sc);"
            "u = O.make_uniforms(sc, len(prims), 7, 0, width=60, height=30, spp=16, max_depth=4);"
            "r = pkg.Renderer(60, 30, conv=2); r.upload_scene(prims); r.render_frame(pkg.api.uniforms_from_bytes(u.tobytes()), pkg.MODE_NONE);"
            "assert r.kernel_name() == 'trace_native_kernel<SrcConst, ShapeAny>', r.kernel_name(); np.save(%r, r.read_pixels())"
            % (ROOT, os.path.join(SCENES, "scene2.json"), out))
    env = dict(os.environ, TRT_B200_JIT="0")
    subprocess.run([sys.executable, "-c", code], check=True, env=env, timeout=300)
    ref = np.load(out)
    a, b = got.reshape(-1, 4)[:, :3], ref.reshape(-1, 4)[:, :3]
    rel = np.abs(a - b) / np.maximum(np.abs(b), 1e-2)
    print(f"  {name}: vs run-time-count build max rel {rel.max():.3g}, pixels differing in bits {(a != b).any(axis=1).sum()} of {60 * 30}")
    assert (rel.max(axis=1) > 1e-4).sum() <= 2
