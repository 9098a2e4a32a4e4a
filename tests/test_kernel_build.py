"""CPU checks of the built CUDA code (no GPU needed): resource budgets of the production trace kernels, the sources
embedded for the run-time compiler, and an NVRTC compilation of those sources for the shapes of the shipped scenes.

The production kernel's occupancy is a design input: 8 resident CTAs of 128 threads per SM need
<= 64 registers per thread and, including the 1 KB the hardware reserves per CTA, <= 228 KB / 8 of shared memory; the
tiled large-scene kernel runs 6 CTAs at <= 80 registers.  A change that silently costs a resident CTA shows up here,
before any GPU time is spent."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "terminal-raytracer_b200", "csrc")
LIB = os.path.join(ROOT, "terminal-raytracer_b200", "libtrt_b200.so")
SM_SHARED = 228 * 1024          # shared memory of one B200 SM (carve-out 100 %), bytes
# cuobjdump's SHARED figure already contains the 1 KB the hardware reserves per resident CTA (Cornell kernel: 11 264 B
# of queue rings + 128 x 124 B of lane records = 27 136 B declared, 28 160 B reported)
JIT_HEADERS = ["trt_device.cuh", "trt_trace.cuh", "trt_native.cuh"]


def _res_usage(path):
    """{mangled kernel name: {REG, STACK, SHARED}} from cuobjdump -res-usage"""
    out = subprocess.run(["cuobjdump", "-res-usage", path], capture_output=True, text=True, check=True).stdout
    res, name = {}, None
    for line in out.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            name = m.group(1)
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+)", line)
        if m and name:
            res[name] = dict(REG=int(m.group(1)), STACK=int(m.group(2)), SHARED=int(m.group(3)))
            name = None
    return res


def test_production_kernels_keep_their_resident_ctas(pkg):
    allk = {k: v for k, v in _res_usage(LIB).items() if "trace_native_kernel" in k}
    res = {k: v for k, v in allk.items() if "trace_native_kernelILi2E" in k}          # conv 2: the production builds
    assert len(res) == 4, sorted(res)          # SrcConst x {ShapeCornell, ShapeAny}, SrcGlobal, SrcTiled
    for name, r in res.items():
        tiled = "SrcTiled" in name
        ctas, regs = (6, 80) if tiled else (8, 64)
        assert r["REG"] <= regs, (name, r)
        assert 128 * r["REG"] * ctas <= 65536, (name, r)
        assert ctas * r["SHARED"] <= SM_SHARED, (name, r, "would lose a resident CTA")
        assert r["STACK"] <= 40, (name, r, "spills in the trace loop")
    cornell = next(v for k, v in res.items() if "ShapeFixed" in k)
    assert cornell["STACK"] <= 8                # one counter that is touched once per pixel
    # the same machine in the strict conventions (conv 0 / 1): 5 resident CTAs, <= 102 registers; its stack frame is the
    # ABI frame of the out-of-line fallbacks (scan_as_written, div3_slow), not spills of the loop (ptxas -v: 0 spill bytes
    # at 5 CTAs; 60 at 6)
    strict = {k: v for k, v in allk.items() if "trace_native_kernelILi0E" in k or "trace_native_kernelILi1E" in k}
    assert len(strict) == 6, sorted(strict)     # {conv 0, conv 1} x {SrcConst ShapeAny, SrcConst ShapeCornell + run list, SrcGlobal}
    for name, r in strict.items():
        assert 128 * r["REG"] * 5 <= 65536, (name, r)
        assert 5 * r["SHARED"] <= SM_SHARED, (name, r, "would lose a resident CTA")


def test_embedded_jit_sources_are_the_headers(pkg):
    """trt_jit.cuh hands NVRTC the headers embedded at build time (build/trt_jit_src.inc): they must be the headers
    in the tree, or a run-time specialised kernel would not be the kernel the static build runs."""
    inc = os.path.join(CSRC, "build", "trt_jit_src.inc")
    if not os.path.exists(inc):
        pytest.skip("library was not built in this tree")
    text = open(inc).read()
    for h in JIT_HEADERS:
        m = re.search(r'\{"%s", R"TRTSRC\((.*?)\)TRTSRC"\},' % re.escape(h), text, re.S)
        assert m, h
        assert m.group(1) == open(os.path.join(CSRC, h)).read(), f"{h}: embedded copy is stale (run make)"
    assert os.path.getmtime(LIB) >= os.path.getmtime(inc)


SHAPES = [
    ("ShapeFixed<3, 6, 2, 0xa05u, 2, 0x0u, 0x4deu>", "Cornell_Box.json"),
    ("ShapeFixed<2, 1, 1, 0x1u, 1, 0x1u, 0x26u>", "scene2.json"),
    ("ShapeFixed<20, 1, 0, 0x1u, 1, 0x1u, 0x0u>", "demo.json"),
    ("ShapeFixed<24, 8, 8, 0xffffu, 16, 0xffu, 0x0u>", "the largest shape the specialisation accepts: general planes, 16 lights"),
]


@pytest.mark.parametrize("shape,what", SHAPES)
def test_runtime_specialisation_compiles(shape, what, tmp_path):
    """the kernel body as trt_jit.cuh compiles it (same sources, options and entry point), per scene shape"""
    nvrtc = pytest.importorskip("cuda.bindings.nvrtc")
    src = ('#include "trt_native.cuh"\nnamespace trt {\ntypedef %s JitShape;\nextern "C" __global__ void '
           '__launch_bounds__(TRT_NV_THREADS, TRT_NV_CTAS_PER_SM)\ntrt_jit_trace(SceneView sc, LightView lv, FrameParams fp) '
           '{ trace_native_body<2, SrcConst, JitShape>(sc, lv, fp); }\n}\n' % shape).encode()
    texts = [open(os.path.join(CSRC, h), "rb").read() for h in JIT_HEADERS]
    err, prog = nvrtc.nvrtcCreateProgram(src, b"trt_jit.cu", len(JIT_HEADERS), texts, [h.encode() for h in JIT_HEADERS])
    assert err == nvrtc.nvrtcResult.NVRTC_SUCCESS
    opts = [b"--gpu-architecture=sm_100a", b"--use_fast_math", b"--fmad=true", b"-std=c++17", b"-lineinfo", b"-default-device",
            b"-diag-suppress=177"]
    err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    _, n = nvrtc.nvrtcGetProgramLogSize(prog)
    log = bytes(n + 16)        # NVRTC writes into it: never a length-0 / length-1 bytes object, CPython shares those
    nvrtc.nvrtcGetProgramLog(prog, log)
    assert err == nvrtc.nvrtcResult.NVRTC_SUCCESS, log.decode(errors="replace")
    _, n = nvrtc.nvrtcGetCUBINSize(prog)
    cubin = bytes(n + 16)
    nvrtc.nvrtcGetCUBIN(prog, cubin)
    path = tmp_path / "jit.cubin"
    path.write_bytes(cubin[:n])
    res = _res_usage(str(path))
    assert list(res) == ["trt_jit_trace"]
    r = res["trt_jit_trace"]
    assert r["REG"] <= 64 and 8 * r["SHARED"] <= SM_SHARED, (what, r)
    # the module's own scene constants are found by name at load time (trt_jit.cuh: cudaLibraryGetGlobal)
    syms = subprocess.run(["cuobjdump", "-elf", str(path)], capture_output=True, text=True).stdout
    assert "c_scene" in syms and "c_small" in syms
