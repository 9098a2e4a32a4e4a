"""
terminal-raytracer_b200 -- B200-native drop-in for Terminal-Raytracer's per-pixel path-tracing hot path.

The directory name carries a hyphen (it mirrors the reference's crate name), so it is imported through
`__graft_entry__.load_package()` which registers it as the module `terminal_raytracer_b200`.
The product is the native library `libtrt_b200.so` (CUDA kernels for sm_100a behind the C ABI of
include/trt_b200.h); `api` is its ctypes face and `multigpu` the torch.distributed/NCCL driver.
"""
from . import api  # noqa: F401
from .api import (Camera, MultiRenderer, Renderer, Scene, TrtError, Uniforms, run,  # noqa: F401
                  CONV_AS_WRITTEN, CONV_FUSED, CONV_NATIVE, MODE_ASCII, MODE_TRUECOLOR, MODE_NONE)
