"""
oracle.py -- Python face of the CPU ORACLE (test infrastructure only; conv 0 is pinned bit for bit to
oracle/_ref, the reference's shader.wgsl transliterated and compiled here -- see trt_oracle.h).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this module.  It holds
  * ctypes bindings to oracle/liboracle.so (the C restatement of src/shader.wgsl and of the
    quantiser src/lib.rs:503-532), and
  * a numpy restatement of the reference's host-side data path: scene JSON -> primitive array
    (src/lib.rs:52-154, src/vec3.rs:15-17, src/primitive.rs:37-92), per-frame uniforms
    (src/lib.rs:417-439), camera (src/camera.rs:17-25) and key handling (src/lib.rs:392-412).
All file:line citations are relative to /root/reference/.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
FAST_LIB_PATH = os.path.join(HERE, "liboracle_fast.so")     # -O3 -march=native: the timed cpu_baseline leg
REF_DIR = os.path.join(HERE, "_ref")
REF_LIB_PATH = os.path.join(REF_DIR, "libref.so")           # the reference's shader.wgsl, transliterated + compiled
REF_FMA_LIB_PATH = os.path.join(REF_DIR, "libref_fma.so")   # the same text with FMA contraction allowed
REF_SOURCE = "/root/reference/src/shader.wgsl"

# ---- wire formats ------------------------------------------------------------------------------
VEC3 = np.dtype([("x", "<f4"), ("y", "<f4"), ("z", "<f4"), ("pad", "<f4")])
# primitive.rs:4-34 -- 180 bytes
PRIM = np.dtype([
    ("type", "<u4"), ("pad1", "<u4"), ("pad2", "<u4"), ("pad3", "<u4"),
    ("color", VEC3), ("emission", VEC3),
    ("reflectivity", "<f4"), ("pad4", "<f4"), ("pad5", "<f4"), ("pad6", "<f4"),
    ("sphere_center", VEC3), ("sphere_radius", "<f4"), ("sphere_pad", "<f4", (3,)),
    ("plane_point", VEC3), ("plane_normal", VEC3),
    ("v0", VEC3), ("v1", VEC3), ("v2", VEC3), ("tri_pad", "<f4"),
])
assert PRIM.itemsize == 180
# lib.rs:23-50 -- 160 bytes (the shader reads the first 112, shader.wgsl:49-69)
UNIFORMS = np.dtype([
    ("width", "<u4"), ("height", "<u4"), ("samples_per_pixel", "<u4"), ("max_depth", "<u4"),
    ("seed", "<u4"), ("frame_number", "<u4"), ("primitive_count", "<u4"), ("pad0", "<u4"),
    ("aspect_ratio", "<f4"), ("char_aspect_ratio", "<f4"), ("fov_rad", "<f4"), ("pad1", "<f4"),
    ("camera_pos", VEC3), ("camera_forward", VEC3), ("camera_right", VEC3), ("camera_up", VEC3),
    ("grid_min", VEC3), ("inv_cell_size", VEC3), ("grid_dims", "<u4", (3,)), ("pad3", "<u4"),
])
assert UNIFORMS.itemsize == 160
VARIANCE = np.dtype([("variance", "<f4"), ("sample_count", "<u4")])

COUNTER_FIELDS = [
    "pixels", "samples_base", "samples_extra", "closest_scans", "hits", "misses", "shadow_scans",
    "shadow_unoccluded", "shadow_lit", "shadow_needed", "mirror", "diffuse", "roulette_tests",
    "roulette_kills", "prim_tests", "prim_tests_min", "flops_written", "flops_min", "rng_draws",
]
COUNTERS = np.dtype([(n, "<u8") for n in COUNTER_FIELDS])

_lib = None


def build(force: bool = False) -> str:
    """Compile liboracle.so with oracle/Makefile (gcc only; no reference sources involved)."""
    srcs = [os.path.join(HERE, f) for f in
            ("trt_oracle.c", "trt_oracle_quant.c", "trt_oracle_core.inc", "trt_oracle.h", "trt_sincos_pinned.h",
             "Makefile")]
    stale = force or not os.path.exists(LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in srcs)
    if stale:
        subprocess.run(["make", "-C", HERE, "-s", "liboracle.so"], check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        try:
            build()
        except Exception:
            if not os.path.exists(LIB_PATH):
                raise
        L = C.CDLL(LIB_PATH)
        f32p, u32p, u8p = C.POINTER(C.c_float), C.POINTER(C.c_uint32), C.POINTER(C.c_uint8)
        L.oracle_pcg_hash.restype = C.c_uint32
        L.oracle_pcg_hash.argtypes = [C.c_uint32]
        L.oracle_rng_stream.restype = None
        L.oracle_rng_stream.argtypes = [C.c_uint32] * 6 + [u32p, f32p]
        L.oracle_sincos_pinned.restype = None
        L.oracle_sincos_pinned.argtypes = [C.c_float, f32p, f32p]
        L.oracle_camera_vectors.restype = None
        L.oracle_camera_vectors.argtypes = [C.c_float, C.c_float, f32p, f32p, f32p]
        L.oracle_tan_half.restype = C.c_float
        L.oracle_tan_half.argtypes = [C.c_float]
        L.oracle_hit_sphere.restype = C.c_float
        L.oracle_hit_sphere.argtypes = [C.c_int, f32p, f32p, f32p, C.c_float, C.c_float, C.c_float]
        L.oracle_hit_plane.restype = C.c_float
        L.oracle_hit_plane.argtypes = [C.c_int, f32p, f32p, f32p, f32p, C.c_float, C.c_float]
        L.oracle_hit_triangle.restype = C.c_float
        L.oracle_hit_triangle.argtypes = [C.c_int, f32p, f32p, f32p, f32p, f32p, C.c_float, C.c_float]
        L.oracle_render.restype = C.c_int
        L.oracle_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_uint32,
                                    C.c_uint32, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                    C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64)]
        L.oracle_quantise.restype = C.c_size_t
        L.oracle_quantise.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_void_p,
                                      C.c_size_t, C.c_void_p]
        L.oracle_quantise_boundary_count.restype = C.c_uint64
        L.oracle_quantise_boundary_count.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32, C.c_int, C.c_float]
        L.oracle_max_threads.restype = C.c_int
        _lib = L
    return _lib


_fast = None


def fast_lib() -> C.CDLL:
    """liboracle_fast.so (-O3 -march=x86-64-v3 -ffp-contract=off): same sources, same bits (tested), used by the
    TIMED cpu_baseline / --impl reference legs only."""
    global _fast
    if _fast is None:
        srcs = [os.path.join(HERE, f) for f in ("trt_oracle.c", "trt_oracle_quant.c", "trt_oracle_core.inc",
                                                "trt_oracle.h", "trt_sincos_pinned.h", "Makefile")]
        if not os.path.exists(FAST_LIB_PATH) or any(os.path.getmtime(x) > os.path.getmtime(FAST_LIB_PATH)
                                                     for x in srcs):
            try:
                subprocess.run(["make", "-C", HERE, "-s", "liboracle_fast.so"], check=True,
                               stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
            except Exception:
                if not os.path.exists(FAST_LIB_PATH):
                    raise
        L = C.CDLL(FAST_LIB_PATH)
        L.oracle_render.restype = C.c_int
        L.oracle_render.argtypes = lib().oracle_render.argtypes
        _fast = L
    return _fast


# ---- oracle/_ref: the reference's own shader text, compiled here -------------------------------------
_ref = {}


def ref_build() -> bool:
    """(Re)build oracle/_ref/*.so from /root/reference/src/shader.wgsl where that file exists (this container).
    On the GPU box the prebuilt files travel with the snapshot and nothing is rebuilt."""
    if not os.path.exists(REF_SOURCE):
        return os.path.exists(REF_LIB_PATH)
    subprocess.run(["make", "-C", HERE, "-s", "ref"], check=True, stdout=subprocess.DEVNULL,
                   stderr=subprocess.DEVNULL)
    return True


def ref_available() -> bool:
    try:
        return ref_build() and os.path.exists(REF_LIB_PATH)
    except Exception:
        return os.path.exists(REF_LIB_PATH)


def ref_lib(fma: bool = False) -> C.CDLL:
    path = REF_FMA_LIB_PATH if fma else REF_LIB_PATH
    if path not in _ref:
        if not ref_available() or not os.path.exists(path):
            raise FileNotFoundError(f"{path}: the transliterated reference shader is not built "
                                    f"(needs {REF_SOURCE} at build time)")
        L = C.CDLL(path)
        f32p, u32p = C.POINTER(C.c_float), C.POINTER(C.c_uint32)
        L.ref_pcg_hash.restype = C.c_uint32
        L.ref_pcg_hash.argtypes = [C.c_uint32]
        L.ref_rng_stream.restype = None
        L.ref_rng_stream.argtypes = [C.c_uint32] * 6 + [u32p, f32p]
        L.ref_hit.restype = C.c_float
        L.ref_hit.argtypes = [C.c_void_p, f32p, f32p, C.c_float, C.c_float]
        L.ref_render.restype = C.c_int
        L.ref_render.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p,
                                 C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ref_max_threads.restype = C.c_int
        _ref[path] = L
    return _ref[path]


def ref_render(prims: np.ndarray, u: np.ndarray, accumulation: np.ndarray | None = None, math_mode=0, threads=0,
               rows=None, idx=None, fma=False):
    """One dispatch of the transliterated reference shader; same arguments and result layout as render()."""
    w, h = int(u["width"]), int(u["height"])
    prims = np.ascontiguousarray(prims)
    assert prims.dtype == PRIM and len(prims) >= int(u["primitive_count"])
    ub = np.ascontiguousarray(u.reshape(1))
    n = len(idx) if idx is not None else w * h
    if accumulation is None:
        accumulation = np.zeros((n, 4), np.float32)
    else:
        accumulation = np.ascontiguousarray(accumulation, np.float32).reshape(n, 4).copy()
    pixels = np.zeros((n, 4), np.float32)
    var = np.zeros(n, VARIANCE)
    y0, y1 = rows if rows is not None else (0, h)
    if idx is not None:
        idx = np.ascontiguousarray(idx, np.uint32)
    rc = ref_lib(fma).ref_render(prims.ctypes.data, ub.ctypes.data, math_mode, threads, y0, y1,
                                 idx.ctypes.data if idx is not None else None, n if idx is not None else 0,
                                 accumulation.ctypes.data, pixels.ctypes.data, var.ctypes.data)
    if rc != 0:
        raise ValueError("ref_render: bad arguments")
    return {"accumulation": accumulation, "pixels": pixels, "variance": var}


def samples_from_variance(var: np.ndarray, spp: int) -> int:
    """ray_color calls a frame made, re-derived from the variance buffer it left behind: every pixel traces
    base = max(4, spp / 4) samples (shader.wgsl:513) plus, when variance > 10 and base < spp, min(spp - base,
    u32(variance * 50)) extra ones (:561-562).  The transliterated shader has no counter; the hand oracle's own
    `samples_traced` equals this (tests/test_oracle_ref.py)."""
    base = max(4, int(spp) // 4)
    v = np.ascontiguousarray(var["variance"], np.float32)
    extra = np.zeros(len(v), np.int64)
    if base < spp:
        m = v > np.float32(10.0)
        prod = (v[m] * np.float32(50.0)).astype(np.float32)
        u = np.where(prod >= np.float32(4294967296.0), np.float32(4294967040.0), prod).astype(np.uint32).astype(np.int64)
        extra[m] = np.minimum(int(spp) - base, u)
    return int(len(v) * base + extra.sum())


def ref_pcg_hash(v: int) -> int:
    return int(ref_lib().ref_pcg_hash(C.c_uint32(v & 0xFFFFFFFF)))


def ref_rng_stream(x, y, w, seed, frame, n):
    st = np.zeros(n, np.uint32)
    va = np.zeros(n, np.float32)
    ref_lib().ref_rng_stream(x, y, w, seed & 0xFFFFFFFF, frame, n,
                             st.ctypes.data_as(C.POINTER(C.c_uint32)), va.ctypes.data_as(C.POINTER(C.c_float)))
    return st, va


def ref_hit(prim: np.ndarray, o, d, tmin, tmax):
    """hit_sphere / hit_plane / hit_triangle of the transliterated shader on one 180-byte primitive record."""
    rec = np.ascontiguousarray(prim.reshape(1))
    (_, po), (_, pd) = _f3(o), _f3(d)
    return np.float32(ref_lib().ref_hit(rec.ctypes.data, po, pd, tmin, tmax))


def _f3(v):
    a = np.ascontiguousarray(v, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


# ---- small entry points --------------------------------------------------------------------------
def pcg_hash(v: int) -> int:
    return int(lib().oracle_pcg_hash(C.c_uint32(v & 0xFFFFFFFF)))


def rng_stream(x, y, w, seed, frame, n):
    st = np.zeros(n, np.uint32)
    va = np.zeros(n, np.float32)
    lib().oracle_rng_stream(x, y, w, seed & 0xFFFFFFFF, frame, n,
                            st.ctypes.data_as(C.POINTER(C.c_uint32)), va.ctypes.data_as(C.POINTER(C.c_float)))
    return st, va


def sincos_pinned(phi):
    s, c = C.c_float(), C.c_float()
    lib().oracle_sincos_pinned(C.c_float(phi), C.byref(s), C.byref(c))
    return np.float32(s.value), np.float32(c.value)


def camera_vectors(yaw, pitch):
    f, r, u = (np.zeros(3, np.float32) for _ in range(3))
    p = C.POINTER(C.c_float)
    lib().oracle_camera_vectors(C.c_float(yaw), C.c_float(pitch), f.ctypes.data_as(p), r.ctypes.data_as(p),
                                u.ctypes.data_as(p))
    return f, r, u


def hit_sphere(o, d, c, r, tmin, tmax, conv=0):
    (_, po), (_, pd), (_, pc) = _f3(o), _f3(d), _f3(c)
    return np.float32(lib().oracle_hit_sphere(conv, po, pd, pc, r, tmin, tmax))


def hit_plane(o, d, p, n, tmin, tmax, conv=0):
    ks = [_f3(v) for v in (o, d, p, n)]
    return np.float32(lib().oracle_hit_plane(conv, *[k[1] for k in ks], tmin, tmax))


def hit_triangle(o, d, v0, v1, v2, tmin, tmax, conv=0):
    ks = [_f3(v) for v in (o, d, v0, v1, v2)]
    return np.float32(lib().oracle_hit_triangle(conv, *[k[1] for k in ks], tmin, tmax))


# ---- numpy restatement of the reference's host data path -----------------------------------------
def _v3(dst, xyz):
    """Vec3::new(f64, f64, f64) -> f32 casts (vec3.rs:15-17)."""
    dst["x"], dst["y"], dst["z"], dst["pad"] = np.float32(xyz[0]), np.float32(xyz[1]), np.float32(xyz[2]), 0.0


def flatten_scene(scene: dict) -> np.ndarray:
    """lib.rs:120-154: spheres, then planes, then triangles, in file order.
    `spheres` and `planes` are mandatory keys, `triangles` defaults to [] (lib.rs:60-63)."""
    spheres, planes = scene["spheres"], scene["planes"]
    tris = scene.get("triangles", [])
    prims = np.zeros(len(spheres) + len(planes) + len(tris), PRIM)
    k = 0
    for s in spheres:                                   # primitive.rs:37-54
        p = prims[k]; k += 1
        p["type"] = 0
        _v3(p["color"], s["color"]); _v3(p["emission"], s["emission"])
        p["reflectivity"] = np.float32(s["reflectivity"])
        _v3(p["sphere_center"], s["center"]); p["sphere_radius"] = np.float32(s["radius"])
    for s in planes:                                    # primitive.rs:56-73
        p = prims[k]; k += 1
        p["type"] = 1
        _v3(p["color"], s["color"]); _v3(p["emission"], s["emission"])
        p["reflectivity"] = np.float32(s["reflectivity"])
        _v3(p["plane_point"], s["point"]); _v3(p["plane_normal"], s["normal"])
    for s in tris:                                      # primitive.rs:75-92
        p = prims[k]; k += 1
        p["type"] = 2
        _v3(p["color"], s["color"]); _v3(p["emission"], s["emission"])
        p["reflectivity"] = np.float32(s["reflectivity"])
        _v3(p["v0"], s["v0"]); _v3(p["v1"], s["v1"]); _v3(p["v2"], s["v2"])
    return prims


def load_scene(path: str) -> dict:
    with open(path) as f:
        return json.load(f)


def to_radians_f32(deg) -> np.float32:
    """Rust f32::to_radians: self * (PI_f32 / 180_f32) (lib.rs:429)."""
    return np.float32(np.float32(deg) * (np.float32(np.pi) / np.float32(180.0)))


class Camera:
    """camera.rs:4-26 + key semantics lib.rs:392-405."""

    def __init__(self, position=(0.0, 0.0, 0.0), yaw=-np.float32(np.pi) / np.float32(2.0), pitch=0.0):
        self.position = np.array(position, np.float32)       # lib.rs:118
        self.yaw = np.float32(yaw)
        self.pitch = np.float32(pitch)

    def vectors(self):
        return camera_vectors(float(self.yaw), float(self.pitch))

    def key(self, k: str) -> bool:
        """Apply one key; returns moved_this_frame (lib.rs:392-405)."""
        f, r, _ = self.vectors()
        s01, s005 = np.float32(0.1), np.float32(0.05)
        moved = True
        if k == "w": self.position = (self.position + f * s01).astype(np.float32)
        elif k == "s": self.position = (self.position - f * s01).astype(np.float32)
        elif k == "a": self.position = (self.position - r * s01).astype(np.float32)
        elif k == "d": self.position = (self.position + r * s01).astype(np.float32)
        elif k == "Up": self.pitch = np.float32(self.pitch + s005)
        elif k == "Down": self.pitch = np.float32(self.pitch - s005)
        elif k == "Left": self.yaw = np.float32(self.yaw - s005)
        elif k == "Right": self.yaw = np.float32(self.yaw + s005)
        else: moved = False
        self.pitch = np.float32(min(max(self.pitch, np.float32(-1.5)), np.float32(1.5)))   # lib.rs:405
        return moved


def make_uniforms(scene: dict, n_prims: int, seed: int, frame_number: int, camera: Camera | None = None,
                  width=None, height=None, spp=None, max_depth=None) -> np.ndarray:
    """lib.rs:417-439 (grid fields are dead and left zero)."""
    cam = camera or Camera()
    f, r, up = cam.vectors()
    u = np.zeros((), UNIFORMS)
    w = int(width if width is not None else scene["width"])
    h = int(height if height is not None else scene["height"])
    u["width"], u["height"] = w, h
    u["samples_per_pixel"] = int(spp if spp is not None else scene["samples_per_pixel"])
    u["max_depth"] = int(max_depth if max_depth is not None else scene["max_depth"])
    u["seed"] = seed & 0xFFFFFFFF
    u["frame_number"] = frame_number
    u["primitive_count"] = n_prims
    u["aspect_ratio"] = np.float32(w) / np.float32(h)
    u["char_aspect_ratio"] = np.float32(scene["camera"]["char_aspect_ratio"])
    u["fov_rad"] = to_radians_f32(scene["camera"]["fov_degrees"])
    for name, v in (("camera_pos", cam.position), ("camera_forward", f), ("camera_right", r), ("camera_up", up)):
        u[name]["x"], u[name]["y"], u[name]["z"] = v[0], v[1], v[2]
    return u


# ---- render / quantise -----------------------------------------------------------------------------
def render(prims: np.ndarray, u: np.ndarray, accumulation: np.ndarray | None = None, conv=0, math_mode=0,
           threads=0, rows=None, idx=None, want_counters=False, fast=False):
    """One frame of shader.wgsl main().  Returns dict(accumulation, pixels, variance[, counters])."""
    w, h = int(u["width"]), int(u["height"])
    prims = np.ascontiguousarray(prims)
    assert prims.dtype == PRIM and len(prims) >= int(u["primitive_count"])
    ub = np.ascontiguousarray(u.reshape(1))
    n = len(idx) if idx is not None else w * h
    if accumulation is None:
        accumulation = np.zeros((n, 4), np.float32)
    else:
        accumulation = np.ascontiguousarray(accumulation, np.float32).reshape(n, 4).copy()
    pixels = np.zeros((n, 4), np.float32)
    var = np.zeros(n, VARIANCE)
    cnt = np.zeros((), COUNTERS) if want_counters else None
    y0, y1 = rows if rows is not None else (0, h)
    traced = C.c_uint64(0)
    if idx is not None:
        idx = np.ascontiguousarray(idx, np.uint32)
    rc = (fast_lib() if fast else lib()).oracle_render(prims.ctypes.data, ub.ctypes.data, conv, math_mode, threads, y0, y1,
                             idx.ctypes.data if idx is not None else None, n if idx is not None else 0,
                             accumulation.ctypes.data, pixels.ctypes.data, var.ctypes.data,
                             cnt.ctypes.data if cnt is not None else None, C.byref(traced))
    if rc != 0:
        raise ValueError("oracle_render: bad arguments")
    out = {"accumulation": accumulation, "pixels": pixels, "variance": var, "samples_traced": int(traced.value)}
    if cnt is not None:
        out["counters"] = {k: int(cnt[k]) for k in COUNTER_FIELDS}
    return out


def quantise(pixels: np.ndarray, w: int, h: int, mode: int, threads=0):
    """lib.rs:503-532 rows block.  mode 0 ASCII, 1 truecolor.  Returns (bytes, row_lengths)."""
    px = np.ascontiguousarray(pixels, np.float32).reshape(w * h, 4)
    cap = w * h * 26 + 2 * h
    out = np.zeros(cap, np.uint8)
    rl = np.zeros(h, np.uint32)
    n = lib().oracle_quantise(px.ctypes.data, w, h, mode, threads, out.ctypes.data, cap, rl.ctypes.data)
    return out[:n].tobytes(), rl


def quantise_boundary_count(pixels, w, h, mode, eps=1e-3) -> int:
    px = np.ascontiguousarray(pixels, np.float32).reshape(w * h, 4)
    return int(lib().oracle_quantise_boundary_count(px.ctypes.data, w, h, mode, C.c_float(eps)))


def max_threads() -> int:
    return int(lib().oracle_max_threads())


# ---- the BASELINE.json configurations --------------------------------------------------
CORNELL = {
    "width": 400, "height": 200, "samples_per_pixel": 128, "max_depth": 3, "frames_to_accumulate": 100000000,
    "camera": {"fov_degrees": 45.0, "char_aspect_ratio": 0.55},
    "spheres": [
        {"center": [-0.5, -0.4, -2.5], "radius": 0.6, "color": [0.9, 0.9, 0.9], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.5},
        {"center": [0.5, -0.7, -3.2], "radius": 0.3, "color": [0.6, 0.8, 0.9], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"center": [0.0, 0.1, -3.2], "radius": 0.4, "color": [0.1, 0.2, 0.4], "emission": [0.0, 0.0, 0.0], "reflectivity": 1.0},
    ],
    "planes": [
        {"point": [0.0, -1.0, 0.0], "normal": [0.0, 1.0, 0.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"point": [0.0, 1.0, 0.0], "normal": [0.0, -1.0, 0.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"point": [1.0, 0.0, 0.0], "normal": [-1.0, 0.0, 0.0], "color": [0.25, 0.75, 0.25], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"point": [-1.0, 0.0, 0.0], "normal": [1.0, 0.0, 0.0], "color": [0.75, 0.25, 0.25], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"point": [0.0, 0.0, 1.0], "normal": [0.0, 0.0, -1.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
        {"point": [0.0, 0.0, -4.0], "normal": [0.0, 0.0, 1.0], "color": [1.0, 1.0, 1.0], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.0},
    ],
    "triangles": [
        {"v0": [-0.4, 0.99, -2.4], "v1": [0.4, 0.99, -2.4], "v2": [-0.4, 0.99, -3.2], "color": [1.0, 1.0, 1.0], "emission": [8.0, 8.0, 8.0], "reflectivity": 0.0},
        {"v0": [0.4, 0.99, -2.4], "v1": [0.4, 0.99, -3.2], "v2": [-0.4, 0.99, -3.2], "color": [1.0, 1.0, 1.0], "emission": [8.0, 8.0, 8.0], "reflectivity": 0.0},
    ],
}
"""The data of src/scenes/Cornell_Box.json (the reference's embedded default scene, lib.rs:107),
re-stated as data so that tests and the bench run on the GPU box where /root/reference is absent."""

SCENE2 = {
    "width": 200, "height": 100, "samples_per_pixel": 32, "max_depth": 4, "frames_to_accumulate": 100000000,
    "camera": {"fov_degrees": 45.0, "char_aspect_ratio": 0.55},
    "spheres": [
        {"center": [0.0, 0.0, -3.0], "radius": 0.5, "color": [0.8, 0.2, 0.2], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.3},
        {"center": [0.0, 2.0, -3.0], "radius": 0.5, "color": [0.0, 0.0, 0.0], "emission": [5.0, 5.0, 5.0], "reflectivity": 0.3},
    ],
    "planes": [
        {"point": [0.0, -1.0, 0.0], "normal": [0.0, 1.0, 0.0], "color": [0.6, 0.6, 0.6], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.1},
    ],
    "triangles": [
        {"v0": [-1.0, 1.0, -4.0], "v1": [1.0, 1.0, -4.0], "v2": [0.0, 2.0, -4.0], "color": [0.2, 0.8, 0.2], "emission": [0.0, 0.0, 0.0], "reflectivity": 0.5},
    ],
}
"""The data of src/scenes/scene2.json."""

C1_SEED = 0x00C0FFEE
C4_SEED = 0x5EED0004
C5_KEYS = ["w", "Right", "d", "Up", "s", "Left", "a", "Down"]


def c2_seed(f: int) -> int:
    return (C1_SEED + 0x9E3779B9 * f) & 0xFFFFFFFF


class _SplitMix64:
    def __init__(self, seed):
        self.s = seed & 0xFFFFFFFFFFFFFFFF

    def next(self):
        self.s = (self.s + 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF
        z = self.s
        z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & 0xFFFFFFFFFFFFFFFF
        z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & 0xFFFFFFFFFFFFFFFF
        return z ^ (z >> 31)

    def uni(self, lo=0.0, hi=1.0):
        """uniform float from the top 24 bits"""
        return lo + (hi - lo) * ((self.next() >> 40) / 16777216.0)


def make_c4_scene(n_spheres=4096, n_tris=16384, width=1920, height=1080, spp=256, max_depth=32, seed=C4_SEED) -> dict:
    """Synthetic linear-scan scene of BASELINE.json configs[3] (definition: SURVEY 8(d) C4)."""
    g = _SplitMix64(seed)

    def material():
        color = [g.uni(0.1, 0.9) for _ in range(3)]
        refl = 0.0 if g.uni() < 0.5 else g.uni()
        return color, refl

    spheres, tris = [], []
    for _ in range(n_spheres):
        c = [g.uni(-8, 8), g.uni(-1, 3), g.uni(-20, -4)]
        r = g.uni(0.05, 0.25)
        color, refl = material()
        spheres.append({"center": c, "radius": r, "color": color, "emission": [0.0, 0.0, 0.0], "reflectivity": refl})
    for _ in range(n_tris):
        v0 = [g.uni(-8, 8), g.uni(-1, 3), g.uni(-20, -4)]
        v1 = [v0[i] + g.uni(-0.3, 0.3) for i in range(3)]
        v2 = [v0[i] + g.uni(-0.3, 0.3) for i in range(3)]
        color, refl = material()
        tris.append({"v0": v0, "v1": v1, "v2": v2, "color": color, "emission": [0.0, 0.0, 0.0], "reflectivity": refl})
    for (x0, x1) in ((-6.0, -2.0), (2.0, 6.0)):          # two emissive quads facing down (-y)
        y, z0, z1 = 5.0, -14.0, -10.0
        e = {"color": [1.0, 1.0, 1.0], "emission": [12.0, 12.0, 12.0], "reflectivity": 0.0}
        tris.append({"v0": [x0, y, z0], "v1": [x1, y, z0], "v2": [x0, y, z1], **e})
        tris.append({"v0": [x1, y, z0], "v1": [x1, y, z1], "v2": [x0, y, z1], **e})
    return {"width": width, "height": height, "samples_per_pixel": spp, "max_depth": max_depth,
            "frames_to_accumulate": 1, "camera": {"fov_degrees": 50.0, "char_aspect_ratio": 0.55},
            "spheres": spheres, "planes": [], "triangles": tris}
