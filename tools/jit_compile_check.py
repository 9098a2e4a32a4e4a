"""Compile the run-time specialised trace kernel with NVRTC on the host (no GPU needed): the same sources, options
and entry point trt_jit.cuh uses.  python tools/jit_compile_check.py ["ShapeFixed<...>"] [extra options...]
Prints registers / spills from the cubin (cuobjdump -res-usage) so a variant can be judged before GPU time is spent."""
import os, subprocess, sys, tempfile
from cuda.bindings import nvrtc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "terminal-raytracer_b200", "csrc")
HDRS = ["trt_device.cuh", "trt_trace.cuh", "trt_native.cuh"]


def compile_shape(shape="ShapeFixed<3, 6, 2, 0xa05u, 2, 0x0u, 0x4deu>", extra=()):
    src = ('#include "trt_native.cuh"\nnamespace trt {\ntypedef %s JitShape;\nextern "C" __global__ void '
           '__launch_bounds__(TRT_NV_THREADS, TRT_NV_CTAS_PER_SM)\ntrt_jit_trace(SceneView sc, LightView lv, FrameParams fp) '
           '{ trace_native_body<2, SrcConst, JitShape>(sc, lv, fp); }\n}\n' % shape).encode()
    texts = [open(os.path.join(CSRC, h), "rb").read() for h in HDRS]
    err, prog = nvrtc.nvrtcCreateProgram(src, b"trt_jit.cu", len(HDRS), texts, [h.encode() for h in HDRS])
    assert err == nvrtc.nvrtcResult.NVRTC_SUCCESS, err
    opts = [b"--gpu-architecture=sm_100a", b"--use_fast_math", b"--fmad=true", b"-std=c++17", b"-lineinfo", b"-default-device",
            b"-diag-suppress=177"] + [e.encode() for e in extra]
    err, = nvrtc.nvrtcCompileProgram(prog, len(opts), opts)
    _, n = nvrtc.nvrtcGetProgramLogSize(prog)
    log = bytes(n + 16)        # NVRTC writes into it: never a length-0 / length-1 bytes object, CPython shares those
    nvrtc.nvrtcGetProgramLog(prog, log)
    if err != nvrtc.nvrtcResult.NVRTC_SUCCESS:
        raise RuntimeError(log.decode(errors="replace"))
    _, n = nvrtc.nvrtcGetCUBINSize(prog)
    cubin = bytes(n + 16)
    nvrtc.nvrtcGetCUBIN(prog, cubin)
    return cubin[:n], log.decode(errors="replace").strip("\x00 \n")


if __name__ == "__main__":
    args = sys.argv[1:]
    shape = args.pop(0) if args and args[0].startswith("Shape") else "ShapeFixed<3, 6, 2, 0xa05u, 2, 0x0u, 0x4deu>"
    cubin, log = compile_shape(shape, args)
    if log:
        print(log)
    with tempfile.NamedTemporaryFile(suffix=".cubin") as f:
        f.write(cubin); f.flush()
        out = subprocess.run(["cuobjdump", "-res-usage", f.name], capture_output=True, text=True).stdout
        n = subprocess.run(["cuobjdump", "-sass", f.name], capture_output=True, text=True).stdout
    print([l.strip() for l in out.splitlines() if "REG" in l])
    print("SASS instructions:", sum(1 for l in n.splitlines() if l.strip().startswith("/*") and ";" in l))
