"""
ctypes face of libtrt_b200.so (include/trt_b200.h) -- the thin Python host layer used by the tests,
bench.py and the multi-GPU driver.  It mirrors the reference's host-side names: `Scene` (SceneConfig,
src/lib.rs:52-98), `Camera` (src/camera.rs), `Uniforms` (src/lib.rs:23-50), `Renderer.render_frame`
(the body of the frame loop, src/lib.rs:415-545).  No arithmetic happens here: every call goes to the
native library, and the library has no CPU fallback -- a missing .so or a missing GPU raises.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
# TRT_B200_LIB selects an experimental build of the same library (kernel tuning A/B runs)
LIB_PATH = os.environ.get("TRT_B200_LIB") or os.path.join(HERE, "libtrt_b200.so")

TRT_ABI_VERSION = 1
CONV_AS_WRITTEN, CONV_FUSED, CONV_NATIVE = 0, 1, 2
MODE_ASCII, MODE_TRUECOLOR, MODE_NONE = 0, 1, 2
ACCEL_NONE, ACCEL_GRID = 0, 1
KEY_UP, KEY_DOWN, KEY_LEFT, KEY_RIGHT = 1001, 1002, 1003, 1004
KEYS = {"w": ord("w"), "a": ord("a"), "s": ord("s"), "d": ord("d"),
        "Up": KEY_UP, "Down": KEY_DOWN, "Left": KEY_LEFT, "Right": KEY_RIGHT}
STATUS = {0: "TRT_OK", -1: "TRT_ERR_INVALID", -2: "TRT_ERR_NO_DEVICE", -3: "TRT_ERR_CUDA", -4: "TRT_ERR_NO_SCENE",
          -5: "TRT_ERR_TOO_SMALL", -6: "TRT_ERR_PARSE", -7: "TRT_ERR_IO", -8: "TRT_ERR_NOMEM"}


class TrtError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"{STATUS.get(code, code)}: {msg}")
        self.code = code


class Vec3(C.Structure):
    _fields_ = [("x", C.c_float), ("y", C.c_float), ("z", C.c_float), ("_padding", C.c_float)]


class Primitive(C.Structure):                        # primitive.rs:4-34
    _fields_ = [("primitive_type", C.c_uint32), ("_padding1", C.c_uint32), ("_padding2", C.c_uint32), ("_padding3", C.c_uint32),
                ("color", Vec3), ("emission", Vec3),
                ("reflectivity", C.c_float), ("_padding4", C.c_float), ("_padding5", C.c_float), ("_padding6", C.c_float),
                ("sphere_center", Vec3), ("sphere_radius", C.c_float), ("_sphere_padding", C.c_float * 3),
                ("plane_point", Vec3), ("plane_normal", Vec3),
                ("triangle_v0", Vec3), ("triangle_v1", Vec3), ("triangle_v2", Vec3), ("_triangle_padding", C.c_float)]


class Uniforms(C.Structure):                         # lib.rs:23-50
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("samples_per_pixel", C.c_uint32), ("max_depth", C.c_uint32),
                ("seed", C.c_uint32), ("frame_number", C.c_uint32), ("primitive_count", C.c_uint32), ("_padding1", C.c_uint32),
                ("aspect_ratio", C.c_float), ("char_aspect_ratio", C.c_float), ("fov_rad", C.c_float), ("_padding2", C.c_float),
                ("camera_pos", Vec3), ("camera_forward", Vec3), ("camera_right", Vec3), ("camera_up", Vec3),
                ("grid_min", Vec3), ("inv_cell_size", Vec3), ("grid_dims", C.c_uint32 * 3), ("_padding3", C.c_uint32)]


class Config(C.Structure):
    _fields_ = [("abi_version", C.c_uint32), ("device", C.c_int32), ("max_width", C.c_uint32), ("max_height", C.c_uint32),
                ("conv", C.c_uint32), ("count_work", C.c_uint32),
                ("shard_index", C.c_uint32), ("shard_count", C.c_uint32), ("stripe_rows", C.c_uint32),
                ("accel", C.c_uint32), ("reserved", C.c_uint32 * 6)]


COUNTER_FIELDS = ["pixels", "samples_base", "samples_extra", "closest_scans", "hits", "misses", "shadow_scans",
                  "shadow_unoccluded", "shadow_lit", "shadow_needed", "mirror", "diffuse", "roulette_tests",
                  "roulette_kills", "prim_tests", "prim_tests_min", "flops_written", "flops_min", "rng_draws"]


class Counters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in COUNTER_FIELDS]


class Timings(C.Structure):
    _fields_ = [("trace_ms", C.c_float), ("quantise_ms", C.c_float), ("d2h_ms", C.c_float), ("total_ms", C.c_float),
                ("kernel_launches", C.c_uint32), ("_pad", C.c_uint32)]


class FrameStats(C.Structure):
    _fields_ = [("samples_traced", C.c_uint64), ("bounces", C.c_uint64), ("pixels", C.c_uint64), ("reserved", C.c_uint64)]


class CameraStruct(C.Structure):
    _fields_ = [("position", Vec3), ("yaw", C.c_float), ("pitch", C.c_float)]


class SceneInfo(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("samples_per_pixel", C.c_uint32), ("max_depth", C.c_uint32),
                ("frames_to_accumulate", C.c_uint32), ("fov_degrees", C.c_float), ("char_aspect_ratio", C.c_float),
                ("n_spheres", C.c_uint32), ("n_planes", C.c_uint32), ("n_triangles", C.c_uint32)]


class RunOptions(C.Structure):
    _fields_ = [("width", C.c_uint32), ("height", C.c_uint32), ("frames", C.c_uint32), ("seed", C.c_uint32),
                ("has_seed", C.c_uint32), ("keys", C.c_char_p), ("out_path", C.c_char_p), ("device", C.c_int32),
                ("conv", C.c_uint32), ("quiet", C.c_uint32), ("batch_frames", C.c_uint32), ("gpus", C.c_uint32),
                ("accel", C.c_uint32), ("pipeline", C.c_uint32)]


assert C.sizeof(Vec3) == 16 and C.sizeof(Primitive) == 180 and C.sizeof(Uniforms) == 160

# every symbol include/trt_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
SYMBOLS = {
    "trt_create": (C.c_int, [C.POINTER(Config), C.POINTER(_P)]),
    "trt_quantise_noalloc": (C.c_int, [C.c_int, _P, C.c_uint32, C.c_uint32, C.c_int, _P, C.c_size_t, _P, _P, C.c_char_p, C.c_size_t]),
    "trt_destroy": (None, [_P]),
    "trt_last_error": (C.c_char_p, [_P]),
    "trt_upload_scene": (C.c_int, [_P, _P, C.c_uint32]),
    "trt_render_frame": (C.c_int, [_P, C.POINTER(Uniforms), C.c_int, _P, C.c_size_t, C.POINTER(C.c_size_t)]),
    "trt_render_frames": (C.c_int, [_P, C.POINTER(Uniforms), C.c_uint32, C.c_int, C.c_int, _P, C.c_size_t, C.POINTER(C.c_size_t),
                                    C.POINTER(C.c_size_t)]),
    "trt_trace_async": (C.c_int, [_P, C.POINTER(Uniforms)]),
    "trt_quantise_async": (C.c_int, [_P, C.c_int, C.POINTER(C.c_size_t)]),
    "trt_sync": (C.c_int, [_P]),
    "trt_read_stream": (C.c_int, [_P, _P, C.c_size_t, C.POINTER(C.c_size_t)]),
    "trt_read_pixels": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_read_accumulation": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_read_variance": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_write_accumulation": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_get_counters": (C.c_int, [_P, C.POINTER(Counters)]),
    "trt_trace_kernel_name": (C.c_char_p, [_P]),
    "trt_get_timings": (C.c_int, [_P, C.POINTER(Timings)]),
    "trt_get_frame_stats": (C.c_int, [_P, C.POINTER(FrameStats)]),
    "trt_measure_fp32_peak": (C.c_int, [C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "trt_selftest_div3": (C.c_int, [C.c_int, C.c_uint64, C.c_uint32, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_float)]),
    "trt_device_buffer": (C.c_int, [_P, C.c_int, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "trt_cuda_stream": (C.c_int, [_P, C.POINTER(_P)]),
    "trt_shard_rows": (C.c_uint32, [_P, C.c_uint32, _P, C.c_uint32]),
    "trt_shard_rows_of": (C.c_uint32, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, _P, C.c_uint32]),
    "trt_pack_shard_pixels": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_quantise_gathered": (C.c_int, [_P, _P, C.c_size_t, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int,
                                        C.POINTER(C.c_size_t)]),
    "trt_alloc_host": (C.c_int, [C.c_size_t, C.POINTER(_P)]),
    "trt_free_host": (None, [_P]),
    "trt_host_register": (C.c_int, [_P, C.c_size_t]),
    "trt_host_unregister": (C.c_int, [_P]),
    "trt_quantise_rows_async": (C.c_int, [_P, C.c_int]),
    "trt_row_lengths": (C.c_int, [_P, _P, C.c_uint32, C.POINTER(C.c_uint32)]),
    "trt_read_stream_rows": (C.c_int, [_P, _P, C.c_size_t, _P, C.c_uint32, C.c_int]),
    "trt_frame_submit": (C.c_int, [_P, C.POINTER(Uniforms), C.c_int]),
    "trt_frame_collect": (C.c_int, [_P, _P, C.c_size_t, C.POINTER(C.c_size_t)]),
    "trt_multi_create": (C.c_int, [C.POINTER(C.c_int32), C.c_uint32, C.POINTER(Config), C.POINTER(_P)]),
    "trt_multi_destroy": (None, [_P]),
    "trt_multi_last_error": (C.c_char_p, [_P]),
    "trt_multi_size": (C.c_uint32, [_P]),
    "trt_multi_context": (_P, [_P, C.c_uint32]),
    "trt_multi_upload_scene": (C.c_int, [_P, _P, C.c_uint32]),
    "trt_multi_render_frame": (C.c_int, [_P, C.POINTER(Uniforms), C.c_int, _P, C.c_size_t, C.POINTER(C.c_size_t)]),
    "trt_multi_read_pixels": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_multi_read_variance": (C.c_int, [_P, _P, C.c_size_t]),
    "trt_multi_get_frame_stats": (C.c_int, [_P, C.POINTER(FrameStats)]),
    "trt_multi_get_timings": (C.c_int, [_P, C.POINTER(Timings)]),
    "trt_multi_render_frames": (C.c_int, [_P, C.POINTER(Uniforms), C.c_uint32, C.c_int, _P, C.c_size_t, C.POINTER(C.c_size_t)]),
    "trt_multi_frames_context": (_P, [_P]),
    "trt_trace_frames_async": (C.c_int, [_P, C.POINTER(Uniforms), C.c_uint32]),
    "trt_frame_means": (C.c_int, [_P, C.POINTER(_P), C.POINTER(C.c_size_t)]),
    "trt_blend_frames_async": (C.c_int, [_P, _P, C.POINTER(Uniforms), C.c_uint32]),
    "trt_camera_default": (None, [C.POINTER(CameraStruct)]),
    "trt_camera_vectors": (None, [C.POINTER(CameraStruct), C.POINTER(Vec3), C.POINTER(Vec3), C.POINTER(Vec3)]),
    "trt_camera_key": (C.c_int, [C.POINTER(CameraStruct), C.c_int]),
    "trt_scene_parse": (C.c_int, [C.c_char_p, C.c_size_t, C.POINTER(_P), C.c_char_p, C.c_size_t]),
    "trt_scene_load": (C.c_int, [C.c_char_p, C.POINTER(_P), C.c_char_p, C.c_size_t]),
    "trt_scene_free": (None, [_P]),
    "trt_scene_get_info": (C.c_int, [_P, C.POINTER(SceneInfo)]),
    "trt_scene_clamp": (None, [_P, C.c_uint32, C.c_uint32]),
    "trt_scene_set_size": (None, [_P, C.c_uint32, C.c_uint32]),
    "trt_scene_primitives": (_P, [_P, C.POINTER(C.c_uint32)]),
    "trt_fill_uniforms": (None, [_P, C.POINTER(CameraStruct), C.c_uint32, C.c_uint32, C.POINTER(Uniforms)]),
    "trt_run": (C.c_int, [C.c_int, C.c_int, C.c_char_p, C.POINTER(RunOptions)]),
    "trt_version": (C.c_char_p, []),
}

_lib = None


def lib() -> C.CDLL:
    """Load the native library; raises if it has not been built (there is no Python fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                              "(the CUDA library is the product; there is no fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


# ---- host logic ------------------------------------------------------------------------------------
class Camera:
    """camera.rs:4-26 with the key semantics of lib.rs:392-405."""

    def __init__(self):
        self.c = CameraStruct()
        lib().trt_camera_default(C.byref(self.c))

    def vectors(self):
        f, r, u = Vec3(), Vec3(), Vec3()
        lib().trt_camera_vectors(C.byref(self.c), C.byref(f), C.byref(r), C.byref(u))
        return tuple(np.array([v.x, v.y, v.z], np.float32) for v in (f, r, u))

    def key(self, k) -> bool:
        return bool(lib().trt_camera_key(C.byref(self.c), KEYS[k] if isinstance(k, str) else int(k)))

    @property
    def position(self):
        p = self.c.position
        return np.array([p.x, p.y, p.z], np.float32)


class Scene:
    """SceneConfig (lib.rs:52-98) parsed natively; primitives flattened in lib.rs:120-154 order."""

    def __init__(self, handle):
        self.h = handle

    @classmethod
    def parse(cls, text: str | bytes) -> "Scene":
        b = text.encode() if isinstance(text, str) else text
        h, err = _P(), C.create_string_buffer(512)
        rc = lib().trt_scene_parse(b, len(b), C.byref(h), err, 512)
        if rc:
            raise TrtError(rc, err.value.decode())
        return cls(h)

    @classmethod
    def load(cls, path: str | None) -> "Scene":
        h, err = _P(), C.create_string_buffer(512)
        rc = lib().trt_scene_load(path.encode() if path else None, C.byref(h), err, 512)
        if rc:
            raise TrtError(rc, err.value.decode())
        return cls(h)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.trt_scene_free(self.h)
            self.h = None

    @property
    def info(self) -> SceneInfo:
        i = SceneInfo()
        lib().trt_scene_get_info(self.h, C.byref(i))
        return i

    def set_size(self, w, h):
        lib().trt_scene_set_size(self.h, w, h)

    def clamp(self, cols, rows):
        lib().trt_scene_clamp(self.h, cols, rows)

    def primitives(self) -> np.ndarray:
        """180-byte records as a (n, 180) uint8 array (a copy)."""
        n = C.c_uint32()
        p = lib().trt_scene_primitives(self.h, C.byref(n))
        if not n.value:
            return np.zeros((0, 180), np.uint8)
        return np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n.value, 180)).copy()

    def uniforms(self, camera: Camera, seed: int, frame_number: int) -> Uniforms:
        u = Uniforms()
        lib().trt_fill_uniforms(self.h, C.byref(camera.c), seed & 0xFFFFFFFF, frame_number, C.byref(u))
        return u


def uniforms_from_bytes(raw: bytes | np.ndarray) -> Uniforms:
    b = raw.tobytes() if isinstance(raw, np.ndarray) else bytes(raw)
    assert len(b) == 160
    return Uniforms.from_buffer_copy(b)


# ---- device path -------------------------------------------------------------------------------------
class Renderer:
    """One GPU context.  `render_frame` = upload uniforms + trace + accumulate + quantise + copy the
    byte stream back (the wgpu section of the reference's frame loop, lib.rs:442-477 + 503-532)."""

    def __init__(self, max_width, max_height, device=0, conv=CONV_AS_WRITTEN, count_work=False,
                 shard_index=0, shard_count=1, stripe_rows=1, accel=0):
        cfg = Config(abi_version=TRT_ABI_VERSION, device=device, max_width=max_width, max_height=max_height,
                     conv=conv, count_work=1 if count_work else 0, shard_index=shard_index,
                     shard_count=shard_count, stripe_rows=stripe_rows, accel=accel)
        self.h = _P()
        self.max_pixels = max_width * max_height
        rc = lib().trt_create(C.byref(cfg), C.byref(self.h))
        if rc:
            self.h = None
            raise TrtError(rc, lib().trt_last_error(None).decode())
        self._out = None
        self.size = (0, 0)

    def close(self):
        if getattr(self, "h", None):
            if self._out is not None:
                lib().trt_free_host(self._out[0])
                self._out = None
            lib().trt_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise TrtError(rc, lib().trt_last_error(self.h).decode())

    def upload_scene(self, prims):
        """prims: Scene, or any buffer of 180-byte primitive records (numpy array)."""
        if isinstance(prims, Scene):
            prims = prims.primitives()
        a = np.ascontiguousarray(prims)
        n = a.nbytes // 180
        assert a.nbytes == n * 180
        self._ck(lib().trt_upload_scene(self.h, a.ctypes.data, n))

    def _host_out(self, cap):
        if self._out is None or self._out[1] < cap:
            if self._out is not None:
                lib().trt_free_host(self._out[0])
            p = _P()
            rc = lib().trt_alloc_host(cap, C.byref(p))
            if rc:
                raise TrtError(rc, lib().trt_last_error(None).decode())
            self._out = (p, cap)
        return self._out

    def render_frame(self, u: Uniforms, mode=MODE_ASCII) -> bytes:
        self.size = (u.width, u.height)
        if mode == MODE_NONE:
            self._ck(lib().trt_render_frame(self.h, C.byref(u), mode, None, 0, None))
            return b""
        cap = u.width * u.height * 26 + 2 * u.height + 64
        p, _ = self._host_out(cap)
        n = C.c_size_t()
        self._ck(lib().trt_render_frame(self.h, C.byref(u), mode, p, cap, C.byref(n)))
        return C.string_at(p, n.value)

    def render_frames(self, us, mode=MODE_ASCII, emit_all=False):
        """Frame batch (trt_render_frames): `us` = uniforms of the frames of a still camera.  Returns the list of
        byte streams of all frames (emit_all) or [last frame's stream]; [] for MODE_NONE."""
        n = len(us)
        arr = (Uniforms * n)(*us)
        self.size = (us[0].width, us[0].height)
        if mode == MODE_NONE:
            self._ck(lib().trt_render_frames(self.h, arr, n, mode, 0, None, 0, None, None))
            return []
        per = us[0].width * us[0].height * 26 + 2 * us[0].height + 64
        cap = per * (n if emit_all else 1)
        p, _ = self._host_out(cap)
        total = C.c_size_t()
        lens = (C.c_size_t * n)()
        self._ck(lib().trt_render_frames(self.h, arr, n, mode, 1 if emit_all else 0, p, cap, C.byref(total), lens))
        if not emit_all:
            return [C.string_at(p, total.value)]
        out, off = [], 0
        for i in range(n):
            out.append(C.string_at(p.value + off, lens[i]))
            off += lens[i]
        assert off == total.value
        return out

    def render_frame_into(self, u: Uniforms, mode, host_ptr, cap) -> int:
        """Zero-copy variant for the benchmark: bytes land in the caller's (pinned) buffer."""
        self.size = (u.width, u.height)
        n = C.c_size_t()
        self._ck(lib().trt_render_frame(self.h, C.byref(u), mode, host_ptr, cap, C.byref(n)))
        return n.value

    def kernel_name(self) -> str:
        """name of the trace kernel the last frame ran"""
        return (lib().trt_trace_kernel_name(self.h) or b"").decode()

    def trace_async(self, u: Uniforms):
        self.size = (u.width, u.height)
        self._ck(lib().trt_trace_async(self.h, C.byref(u)))

    def quantise_async(self, mode, want_len=False):
        n = C.c_size_t()
        self._ck(lib().trt_quantise_async(self.h, mode, C.byref(n) if want_len else None))
        return n.value

    def sync(self):
        self._ck(lib().trt_sync(self.h))

    def read_stream(self) -> bytes:
        w, h = self.size
        cap = w * h * 26 + 2 * h + 64
        p, _ = self._host_out(cap)
        n = C.c_size_t()
        self._ck(lib().trt_read_stream(self.h, p, cap, C.byref(n)))
        return C.string_at(p, n.value)

    def read_stream_into(self, host_ptr, cap) -> int:
        """Copy the last quantised stream into the caller's (pinned) buffer; returns its length."""
        n = C.c_size_t()
        self._ck(lib().trt_read_stream(self.h, host_ptr, cap, C.byref(n)))
        return n.value

    def _read(self, fn, dtype, cols):
        w, h = self.size
        a = np.zeros((w * h, cols), dtype) if cols else np.zeros(w * h, dtype)
        self._ck(fn(self.h, a.ctypes.data, w * h))
        return a

    def read_pixels(self):
        return self._read(lib().trt_read_pixels, np.float32, 4)

    def read_accumulation(self):
        return self._read(lib().trt_read_accumulation, np.float32, 4)

    def read_variance(self):
        return self._read(lib().trt_read_variance, np.dtype([("variance", "<f4"), ("sample_count", "<u4")]), 0)

    def write_accumulation(self, a):
        a = np.ascontiguousarray(a, np.float32)
        self._ck(lib().trt_write_accumulation(self.h, a.ctypes.data, a.size // 4))

    def counters(self) -> dict:
        c = Counters()
        self._ck(lib().trt_get_counters(self.h, C.byref(c)))
        return {k: int(getattr(c, k)) for k in COUNTER_FIELDS}

    def timings(self) -> Timings:
        t = Timings()
        self._ck(lib().trt_get_timings(self.h, C.byref(t)))
        return t

    def frame_stats(self) -> FrameStats:
        st = FrameStats()
        self._ck(lib().trt_get_frame_stats(self.h, C.byref(st)))
        return st

    def device_buffer(self, which):
        p, n = _P(), C.c_size_t()
        self._ck(lib().trt_device_buffer(self.h, which, C.byref(p), C.byref(n)))
        return p.value, n.value

    def cuda_stream(self) -> int:
        s = _P()
        self._ck(lib().trt_cuda_stream(self.h, C.byref(s)))
        return s.value or 0

    def shard_rows(self, height) -> np.ndarray:
        rows = np.zeros(height, np.uint32)
        n = lib().trt_shard_rows(self.h, height, rows.ctypes.data, height)
        return rows[:n]

    # ---- row-addressed streams (sharded frames without moving a pixel) ----
    def quantise_rows_async(self, mode):
        """Quantise the rows this context owns and compute every row's byte length (device buffer 5)."""
        self._ck(lib().trt_quantise_rows_async(self.h, mode))

    def row_lengths(self) -> np.ndarray:
        """Byte length of each owned row of the last quantise_rows_async (waits for it)."""
        out = np.zeros(max(1, self.size[1]), np.uint32)
        n = C.c_uint32()
        self._ck(lib().trt_row_lengths(self.h, out.ctypes.data, len(out), C.byref(n)))
        return out[:n.value]

    def read_stream_rows(self, host_ptr, cap, row_off: np.ndarray, wait=True):
        """Copy the owned rows to host_ptr + row_off[y] (row_off: height + 1 uint64 offsets of the whole frame)."""
        ro = np.ascontiguousarray(row_off, np.uint64)
        self._ck(lib().trt_read_stream_rows(self.h, host_ptr, cap, ro.ctypes.data, len(ro) - 1, 1 if wait else 0))

    # ---- pipelined frames ----
    def frame_submit(self, u: Uniforms, mode):
        self.size = (u.width, u.height)
        self._ck(lib().trt_frame_submit(self.h, C.byref(u), mode))

    def frame_collect(self) -> bytes:
        w, h = self.size
        cap = w * h * 26 + 2 * h + 64
        p, _ = self._host_out(cap)
        n = C.c_size_t()
        self._ck(lib().trt_frame_collect(self.h, p, cap, C.byref(n)))
        return C.string_at(p, n.value)

    def pack_shard_pixels(self, dptr, nbytes):
        self._ck(lib().trt_pack_shard_pixels(self.h, dptr, nbytes))

    def quantise_gathered(self, dptr, block_stride_bytes, width, height, shard_count, stripe_rows, mode, want_len=True):
        n = C.c_size_t()
        self._ck(lib().trt_quantise_gathered(self.h, dptr, block_stride_bytes, width, height, shard_count, stripe_rows,
                                             mode, C.byref(n) if want_len else None))
        self.size = (width, height)
        return n.value


class MultiRenderer:
    """One frame over several GPUs of THIS process (trt_multi_*): one row-sharded context per device; render_frame
    returns the assembled byte stream, byte-identical to a single-GPU frame."""

    def __init__(self, max_width, max_height, devices, conv=CONV_AS_WRITTEN, stripe_rows=4):
        cfg = Config(abi_version=TRT_ABI_VERSION, device=0, max_width=max_width, max_height=max_height, conv=conv,
                     stripe_rows=stripe_rows)
        devs = (C.c_int32 * len(devices))(*devices)
        self.h = _P()
        rc = lib().trt_multi_create(devs, len(devices), C.byref(cfg), C.byref(self.h))
        if rc:
            self.h = None
            raise TrtError(rc, lib().trt_multi_last_error(None).decode())
        self._out = None
        self.size = (0, 0)

    def close(self):
        if getattr(self, "h", None):
            if self._out is not None:
                lib().trt_free_host(self._out[0])
                self._out = None
            lib().trt_multi_destroy(self.h)
            self.h = None

    __del__ = close

    def _ck(self, rc):
        if rc:
            raise TrtError(rc, lib().trt_multi_last_error(self.h).decode())

    def __len__(self):
        return int(lib().trt_multi_size(self.h))

    def upload_scene(self, prims):
        if isinstance(prims, Scene):
            prims = prims.primitives()
        a = np.ascontiguousarray(prims)
        self._ck(lib().trt_multi_upload_scene(self.h, a.ctypes.data, a.nbytes // 180))

    def render_frame(self, u: Uniforms, mode=MODE_ASCII) -> bytes:
        self.size = (u.width, u.height)
        if mode == MODE_NONE:
            self._ck(lib().trt_multi_render_frame(self.h, C.byref(u), mode, None, 0, None))
            return b""
        cap = u.width * u.height * 26 + 2 * u.height + 64
        if self._out is None or self._out[1] < cap:
            if self._out is not None:
                lib().trt_free_host(self._out[0])
            p = _P()
            rc = lib().trt_alloc_host(cap, C.byref(p))
            if rc:
                raise TrtError(rc, lib().trt_last_error(None).decode())
            self._out = (p, cap)
        n = C.c_size_t()
        self._ck(lib().trt_multi_render_frame(self.h, C.byref(u), mode, self._out[0], cap, C.byref(n)))
        return C.string_at(self._out[0], n.value)

    def render_frames(self, us, mode=MODE_ASCII) -> bytes:
        """Frames of a still camera dealt over the GPUs (frame f on GPU f mod G), blended in order on the first one;
        returns the last frame's stream.  `frames_pixels()` / `frames_variance()` read the blended result."""
        n = len(us)
        arr = (Uniforms * n)(*us)
        self.size = (us[0].width, us[0].height)
        if mode == MODE_NONE:
            self._ck(lib().trt_multi_render_frames(self.h, arr, n, mode, None, 0, None))
            return b""
        cap = us[0].width * us[0].height * 26 + 2 * us[0].height + 64
        buf = C.create_string_buffer(cap)
        ln = C.c_size_t()
        self._ck(lib().trt_multi_render_frames(self.h, arr, n, mode, C.cast(buf, _P), cap, C.byref(ln)))
        return buf.raw[:ln.value]

    def _frames_read(self, fn, dtype, cols):
        ctx = lib().trt_multi_frames_context(self.h)
        w, h = self.size
        a = np.zeros((w * h, cols), dtype) if cols else np.zeros(w * h, dtype)
        rc = fn(ctx, a.ctypes.data, w * h)
        if rc:
            raise TrtError(rc, lib().trt_last_error(ctx).decode())
        return a

    def frames_pixels(self):
        return self._frames_read(lib().trt_read_pixels, np.float32, 4)

    def frames_variance(self):
        return self._frames_read(lib().trt_read_variance, np.dtype([("variance", "<f4"), ("sample_count", "<u4")]), 0)

    def read_pixels(self):
        w, h = self.size
        a = np.zeros((w * h, 4), np.float32)
        self._ck(lib().trt_multi_read_pixels(self.h, a.ctypes.data, w * h))
        return a

    def read_variance(self):
        w, h = self.size
        a = np.zeros(w * h, np.dtype([("variance", "<f4"), ("sample_count", "<u4")]))
        self._ck(lib().trt_multi_read_variance(self.h, a.ctypes.data, w * h))
        return a

    def frame_stats(self) -> FrameStats:
        st = FrameStats()
        self._ck(lib().trt_multi_get_frame_stats(self.h, C.byref(st)))
        return st

    def timings(self) -> Timings:
        t = Timings()
        self._ck(lib().trt_multi_get_timings(self.h, C.byref(t)))
        return t


def shard_rows_of(shard_index, shard_count, stripe_rows, height) -> np.ndarray:
    rows = np.zeros(max(1, height), np.uint32)
    n = lib().trt_shard_rows_of(shard_index, shard_count, stripe_rows, height, rows.ctypes.data, height)
    return rows[:n]


def selftest_div3(n_cases, seed=1, distribution=0, device=0):
    """(mismatching components, (a, s, a/s expected, got) of the first one)"""
    bad, first = C.c_uint64(), (C.c_float * 4)()
    rc = lib().trt_selftest_div3(device, n_cases, seed, distribution, C.byref(bad), first)
    if rc:
        raise TrtError(rc, lib().trt_last_error(None).decode())
    return bad.value, tuple(first)


def measure_fp32_peak(device=0):
    """(TFLOP/s measured by a register-resident FFMA loop, SM count, clock kHz)."""
    tf, sm, khz = C.c_float(), C.c_int(), C.c_int()
    rc = lib().trt_measure_fp32_peak(device, C.byref(tf), C.byref(sm), C.byref(khz))
    if rc:
        raise TrtError(rc, lib().trt_last_error(None).decode())
    return tf.value, sm.value, khz.value


def run(full_color: bool, verbose: bool, scene_path: str | None, **opts) -> int:
    """The reference's `run(full_color, verbose, scene_path)` (lib.rs:100); keyword extensions map to
    trt_run_options (width, height, frames, seed, keys, out_path, device, conv, quiet, batch_frames, gpus, pipeline)."""
    o = RunOptions()
    for k, v in opts.items():
        if k == "seed":
            o.seed, o.has_seed = v & 0xFFFFFFFF, 1
        elif k in ("keys", "out_path"):
            setattr(o, k, v.encode() if isinstance(v, str) else v)
        else:
            setattr(o, k, v)
    return lib().trt_run(1 if full_color else 0, 1 if verbose else 0, scene_path.encode() if scene_path else None, C.byref(o))
