"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bar: float buffers bit-exact in both arithmetic conventions (the oracle and the kernels execute the
same IEEE operation sequence); byte streams bit-exact except pixels within eps of a quantisation
boundary (counted)."""
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


CASES = [
    ("Cornell_Box.json", 96, 48, 16, 32, 0x00C0FFEE),
    ("Cornell_Box.json", 64, 32, 128, 3, 7),          # the file's own spp/depth
    ("scene2.json", 100, 50, 32, 4, 1),               # sphere light: NEE live
    ("demo.json", 80, 40, 16, 32, 2),                 # 21 primitives, open scene (sky)
]


@pytest.mark.parametrize("conv", [0, 1])
@pytest.mark.parametrize("name,w,h,spp,depth,seed", CASES)
def test_frame_bit_exact(pkg, O, name, w, h, spp, depth, seed, conv):
    sc = _scene(name)
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), seed, 0, width=w, height=h, spp=spp, max_depth=depth)
    ref = O.render(prims, u, conv=conv)
    r = pkg.Renderer(w, h, conv=conv)
    r.upload_scene(prims)
    r.render_frame(_u(pkg, u), pkg.MODE_NONE)
    got, var = r.read_pixels(), r.read_variance()
    acc = r.read_accumulation()
    r.close()
    bad = int((_bits(got) != _bits(ref["pixels"])).any(axis=1).sum())
    assert bad == 0, f"{bad}/{w*h} pixels differ"
    assert np.array_equal(_bits(acc), _bits(ref["accumulation"]))
    assert np.array_equal(_bits(var["variance"]), _bits(ref["variance"]["variance"]))
    assert np.array_equal(var["sample_count"], ref["variance"]["sample_count"])


@pytest.mark.parametrize("conv", [0, 1])
def test_counters_match_oracle(pkg, O, conv):
    sc = _scene("scene2.json")
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), 5, 0, width=64, height=32)
    ref = O.render(prims, u, conv=conv, want_counters=True)
    r = pkg.Renderer(64, 32, conv=conv, count_work=True)
    r.upload_scene(prims)
    r.render_frame(_u(pkg, u), pkg.MODE_NONE)
    got = r.read_pixels()
    cnt = r.counters()
    r.close()
    assert np.array_equal(_bits(got), _bits(ref["pixels"]))       # instrumented kernel == oracle
    assert cnt == ref["counters"]


def test_temporal_accumulation(pkg, O):
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h = 64, 32
    r = pkg.Renderer(w, h)
    r.upload_scene(prims)
    acc = None
    for f in range(4):
        u = O.make_uniforms(sc, len(prims), O.c2_seed(f), f, width=w, height=h, spp=16, max_depth=8)
        ref = O.render(prims, u, accumulation=acc)
        acc = ref["accumulation"]
        r.render_frame(_u(pkg, u), pkg.MODE_NONE)
        assert np.array_equal(_bits(r.read_pixels()), _bits(ref["pixels"])), f"frame {f}"
    r.close()


@pytest.mark.parametrize("mode", [0, 1])
def test_stream_bytes(pkg, O, mode):
    sc = _scene("Cornell_Box.json")
    prims = O.flatten_scene(sc)
    w, h = 100, 37
    u = O.make_uniforms(sc, len(prims), 11, 0, width=w, height=h, spp=16, max_depth=16)
    r = pkg.Renderer(w, h)
    r.upload_scene(prims)
    stream = r.render_frame(_u(pkg, u), mode)
    px = r.read_pixels()
    r.close()
    ref_stream, _ = O.quantise(px, w, h, mode)
    near = O.quantise_boundary_count(px, w, h, mode, 1e-3)
    if stream != ref_stream:
        # only boundary pixels may differ, and only in ASCII mode (powf); count them
        assert mode == 0 and len(stream) == len(ref_stream)
        diff = sum(a != b for a, b in zip(stream, ref_stream))
        assert diff <= near, f"{diff} differing bytes but only {near} boundary pixels"
    else:
        assert stream == ref_stream


def test_stream_edge_values(pkg, O):
    """NaN / negative / huge / zero pixels through the quantiser (Rust cast semantics, lib.rs:510-524)."""
    w, h = 8, 3
    px = np.zeros((w * h, 4), np.float32)
    vals = [0.0, -0.0, 1.0, 1e-9, 0.5, 2.0, np.inf, -1.0, np.nan, 1e30, 0.00390625, 0.999999, 0.25, 1e-30, 3.0, 0.1]
    for i in range(w * h):
        px[i, 0], px[i, 1], px[i, 2] = vals[i % 16], vals[(i * 3 + 1) % 16], vals[(i * 7 + 2) % 16]
    r = pkg.Renderer(w, h)
    prims = O.flatten_scene(O.CORNELL)
    r.upload_scene(prims)
    u = O.make_uniforms(O.CORNELL, len(prims), 1, 1, width=w, height=h, spp=1, max_depth=0)
    # max_depth 0: every sample is black, frame_number 1: acc = acc*(1-1/2) + 0 -> halves what we wrote
    r.write_accumulation(px * 2.0)
    for mode in (0, 1):
        r.write_accumulation(px * 2.0)
        stream = r.render_frame(_u(pkg, u), mode)
        got = r.read_pixels()
        ref_stream, _ = O.quantise(got, w, h, mode)
        assert stream == ref_stream
    r.close()


def test_row_shards_union_equals_whole(pkg, O):
    sc = _scene("scene2.json")
    prims = O.flatten_scene(sc)
    w, h = 50, 23
    u = O.make_uniforms(sc, len(prims), 3, 0, width=w, height=h, spp=8, max_depth=4)
    whole = pkg.Renderer(w, h)
    whole.upload_scene(prims)
    whole.render_frame(_u(pkg, u), pkg.MODE_NONE)
    ref = whole.read_pixels()
    whole.close()
    union = np.full((w * h, 4), np.nan, np.float32)
    seen = np.zeros(h, bool)
    for s in range(3):
        r = pkg.Renderer(w, h, shard_index=s, shard_count=3, stripe_rows=2)
        r.upload_scene(prims)
        r.render_frame(_u(pkg, u), pkg.MODE_NONE)
        rows = r.shard_rows(h)
        part = r.read_pixels().reshape(h, w, 4)
        union.reshape(h, w, 4)[rows] = part[rows]
        assert not seen[rows].any()
        seen[rows] = True
        r.close()
    assert seen.all()
    assert np.array_equal(_bits(union), _bits(ref))


@pytest.mark.parametrize("conv", [0, 1])
def test_large_scene_global_memory_path(pkg, O, conv):
    """> 256 primitives: geometry comes from the SoA arrays in HBM, not from __constant__ memory.
    A reduced C4 (BASELINE.json configs[3]): random spheres + triangles + 2 emissive quads, NEE live."""
    sc = O.make_c4_scene(n_spheres=150, n_tris=400, width=64, height=36, spp=16, max_depth=6)
    prims = O.flatten_scene(sc)
    assert len(prims) > 256
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0)
    ref = O.render(prims, u, conv=conv, want_counters=True)
    r = pkg.Renderer(64, 36, conv=conv)
    r.upload_scene(prims)
    stream = r.render_frame(_u(pkg, u), pkg.MODE_ASCII)
    got = r.read_pixels()
    r.close()
    bad = int((_bits(got) != _bits(ref["pixels"])).any(axis=1).sum())
    assert bad == 0, f"{bad} pixels differ"
    assert stream == O.quantise(got, 64, 36, 0)[0]
    rc = pkg.Renderer(64, 36, conv=conv, count_work=True)
    rc.upload_scene(prims)
    rc.render_frame(_u(pkg, u), pkg.MODE_NONE)
    assert rc.counters() == ref["counters"]
    rc.close()


def test_mixed_plane_kinds_and_run_order(pkg, O):
    """Scan-order corner cases for the production scan: tilted planes, single axis-aligned planes,
    slab pairs in both orders, coincident planes, a ray origin within t_min of a wall, unknown types."""
    sc = {"width": 72, "height": 40, "samples_per_pixel": 16, "max_depth": 8, "frames_to_accumulate": 1,
          "camera": {"fov_degrees": 60.0, "char_aspect_ratio": 0.55},
          "spheres": [{"center": [0.3, -0.2, -2.0], "radius": 0.5, "color": [0.9, 0.5, 0.5], "emission": [0, 0, 0], "reflectivity": 0.4},
                      {"center": [-0.8, 0.6, -2.6], "radius": 0.3, "color": [1, 1, 1], "emission": [6, 6, 6], "reflectivity": 0.0}],
          "planes": [
              {"point": [0, 1.5, 0], "normal": [0, -1, 0], "color": [0.8, 0.8, 0.8], "emission": [0, 0, 0], "reflectivity": 0.0},   # pair (hi first)
              {"point": [0, -1.0, 0], "normal": [0, 1, 0], "color": [0.7, 0.7, 0.9], "emission": [0, 0, 0], "reflectivity": 0.2},
              {"point": [0, -1.0005, 0], "normal": [0, 1, 0], "color": [0.1, 0.9, 0.1], "emission": [0, 0, 0], "reflectivity": 0.0},  # single, 0.5 mm below the floor
              {"point": [0, 0, -4], "normal": [0.2, 0.1, 1.0], "color": [0.9, 0.9, 0.3], "emission": [0, 0, 0], "reflectivity": 0.0},  # tilted, un-normalised
              {"point": [-2, 0, 0], "normal": [1, 0, 0], "color": [0.9, 0.3, 0.3], "emission": [0, 0, 0], "reflectivity": 0.0},      # pair (lo first)
              {"point": [2, 0, 0], "normal": [-1, 0, 0], "color": [0.3, 0.3, 0.9], "emission": [0, 0, 0], "reflectivity": 0.6},
              {"point": [0, 0, 1], "normal": [0, 0, -2], "color": [0.5, 0.5, 0.5], "emission": [0, 0, 0], "reflectivity": 0.0},      # axis direction but |n| = 2: general plane
              {"point": [0, 0, 3], "normal": [0, 0, -1], "color": [0.5, 0.5, 0.5], "emission": [0, 0, 0], "reflectivity": 0.0},      # coincident pair -> not paired
              {"point": [0, 0, 3], "normal": [0, 0, -1], "color": [0.9, 0.1, 0.9], "emission": [0, 0, 0], "reflectivity": 0.0}],
          "triangles": [{"v0": [-1, 1.49, -3], "v1": [1, 1.49, -3], "v2": [0, 1.49, -1.5], "color": [1, 1, 1], "emission": [9, 9, 9], "reflectivity": 0.0},
                        {"v0": [-1.5, -1, -3.5], "v1": [-0.5, -1, -3.5], "v2": [-1, 0.5, -3.2], "color": [0.2, 0.8, 0.8], "emission": [0, 0, 0], "reflectivity": 0.3}]}
    prims = O.flatten_scene(sc)
    weird = prims.copy()
    weird = np.concatenate([weird[:3], weird[2:3], weird[3:]])      # an unknown-type record in the middle of the scan order
    weird[3]["type"] = 7
    for P in (prims, weird):
        for conv in (0, 1):
            u = O.make_uniforms(sc, len(P), 21, 0)
            ref = O.render(P, u, conv=conv)
            r = pkg.Renderer(72, 40, conv=conv)
            r.upload_scene(P)
            r.render_frame(_u(pkg, u), pkg.MODE_NONE)
            got = r.read_pixels()
            r.close()
            bad = int((_bits(got) != _bits(ref["pixels"])).any(axis=1).sum())
            assert bad == 0, f"conv {conv}: {bad} pixels differ"


def test_deferred_machine_is_bit_exact_on_every_scene(tmp_path):
    """The production kernel's machine (deferred NEE, parking, ordered late sums) in the strict conventions, forced onto
    every scene (by default it serves the default Cornell layout, where its scan is specialised; the wave kernel serves the
    others): the strict frame tests of this file and the golden fixtures, in a process with TRT_B200_STRICT=deferred."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TRT_B200_STRICT="deferred")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-x",
                        os.path.join(root, "tests", "test_gpu_parity.py"), os.path.join(root, "tests", "test_golden.py"),
                        os.path.join(root, "tests", "test_gpu_edges.py"), "-k", "not deferred_machine"],
                       env=env, capture_output=True, text=True, timeout=900, cwd=root)
    assert r.returncode == 0, r.stdout[-3000:]
    probe = ("import json,os,sys; sys.path.insert(0, %r); import __graft_entry__ as g; pkg, O = g.load_package(), g.load_oracle(); "
             "prims = O.flatten_scene(O.SCENE2); u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.SCENE2, len(prims), 1, 0, width=32, height=16).tobytes()); "
             "r = pkg.Renderer(32, 16, conv=0); r.upload_scene(prims); r.render_frame(u, pkg.MODE_NONE); print(r.kernel_name())" % root)
    name = subprocess.run([sys.executable, "-c", probe], env=env, capture_output=True, text=True, timeout=300, cwd=root).stdout
    assert "trace_native_kernel<CONV, SrcConst, ShapeAny> (strict)" in name, name


def test_chain_pool_machine_is_bit_exact(tmp_path):
    """trace_pool_kernel (trt_pool.cuh: a pool of pixel chains per warp, regeneration at SIMT width) serves strict frames
    that hold several pools' worth of pixels (C3; checked there inside the bench run, config.parity).  Here it is forced
    onto the small frames of the strict tests and of the named configurations: TRT_B200_POOL=1."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, TRT_B200_POOL="1")
    r = subprocess.run([sys.executable, "-m", "pytest", "-q", "-m", "gpu", "-x",
                        os.path.join(root, "tests", "test_gpu_parity.py"), os.path.join(root, "tests", "test_golden.py"),
                        os.path.join(root, "tests", "test_gpu_batch.py"), os.path.join(root, "tests", "test_gpu_named_configs.py"),
                        "-k", "not deferred_machine and not chain_pool and not c5_all and not c4_full"],
                       env=env, capture_output=True, text=True, timeout=900, cwd=root)
    assert r.returncode == 0, r.stdout[-3000:]
    probe = ("import json,os,sys; sys.path.insert(0, %r); import __graft_entry__ as g; pkg, O = g.load_package(), g.load_oracle(); "
             "prims = O.flatten_scene(O.CORNELL); u = pkg.api.uniforms_from_bytes(O.make_uniforms(O.CORNELL, len(prims), 1, 0, width=32, height=16).tobytes()); "
             "r = pkg.Renderer(32, 16, conv=0); r.upload_scene(prims); r.render_frame(u, pkg.MODE_NONE); print(r.kernel_name())" % root)
    name = subprocess.run([sys.executable, "-c", probe], env=env, capture_output=True, text=True, timeout=300, cwd=root).stdout
    assert "trace_pool_kernel<CONV, SrcConst, ShapeCornell> (strict)" in name, name
