"""The opt-in acceleration structure (trt_config.accel = TRT_ACCEL_GRID, `--accel grid`): the uniform grid the reference
builds and never uses (lib.rs:181-244).  It is OUTSIDE the parity contract (the reference executes a linear scan), so it
is held to the conv 2 bars against the oracle's linear scan, not to bits -- but since it returns the same (t, index) per
ray, the frames in fact agree with the linear-scan conv 2 frame far more closely than the bars ask."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frame(pkg, prims, u_np, w, h, **kw):
    r = pkg.Renderer(w, h, conv=2, **kw)
    r.upload_scene(prims)
    r.render_frame(pkg.api.uniforms_from_bytes(u_np.tobytes()), pkg.MODE_NONE)
    px, name, st = r.read_pixels(), r.kernel_name(), r.frame_stats()
    ms = r.timings().trace_ms
    r.close()
    return px, name, st, ms


def test_grid_frame_tracks_the_linear_scan(pkg, O):
    sc = O.make_c4_scene(n_spheres=300, n_tris=900, width=96, height=54, spp=16, max_depth=1)
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0)
    ref = O.render(prims, u)["pixels"]
    lin, name_lin, st_lin, _ = _frame(pkg, prims, u, 96, 54)
    grid, name_grid, st_grid, _ = _frame(pkg, prims, u, 96, 54, accel=pkg.api.ACCEL_GRID)
    assert "SrcGrid" in name_grid and "SrcGrid" not in name_lin
    rel = lambda a, b: np.abs(a[:, :3].astype(np.float64) - b[:, :3]) / np.maximum(np.abs(b[:, :3]), 1e-2)
    over_grid = int((rel(grid, ref).max(axis=1) > 1e-4).sum())
    over_lin = int((rel(lin, ref).max(axis=1) > 1e-4).sum())
    same = int((grid.view(np.uint32) == lin.view(np.uint32)).all(axis=1).sum())
    print(f"\ngrid vs oracle: {over_grid} pixels over 1e-4 (linear scan conv 2: {over_lin}); {same} of {len(grid)} pixels bit-equal to the linear-scan frame")
    assert over_grid <= 2 * over_lin + 30
    assert int(st_grid.samples_traced) == int(st_lin.samples_traced)


def test_grid_on_the_full_c4_scene_and_planes(pkg, O):
    """20 484 primitives + a ground plane (planes are unbounded: tested outside the grid); deep paths; speed-up reported."""
    sc = O.make_c4_scene(width=160, height=90, spp=4, max_depth=8)
    sc["planes"] = [{"point": [0.0, -1.5, 0.0], "normal": [0.0, 1.0, 0.0], "color": [0.5, 0.5, 0.5], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.2}]
    prims = O.flatten_scene(sc)
    u = O.make_uniforms(sc, len(prims), O.C4_SEED, 0)
    lin, _, st_lin, ms_lin = _frame(pkg, prims, u, 160, 90)
    grid, name, st_grid, ms_grid = _frame(pkg, prims, u, 160, 90, accel=pkg.api.ACCEL_GRID)
    d = np.abs(grid[:, :3].astype(np.float64) - lin[:, :3])
    rmse = float(np.sqrt((d ** 2).mean()))
    print(f"\nfull C4 + plane, 160x90 x 4 spp depth 8: linear scan {ms_lin:.1f} ms, grid {ms_grid:.1f} ms ({ms_lin / ms_grid:.1f}x); RMSE between the frames {rmse:.3g}")
    assert np.isfinite(grid).all()
    assert abs(int(st_grid.samples_traced) - int(st_lin.samples_traced)) <= int(st_lin.samples_traced) // 100
    assert rmse < 0.05 * float(np.sqrt((lin[:, :3].astype(np.float64) ** 2).mean()) + 1e-9) + 0.02


def test_accel_is_refused_outside_conv2(pkg):
    with pytest.raises(pkg.TrtError):
        pkg.Renderer(32, 16, conv=0, accel=pkg.api.ACCEL_GRID)
