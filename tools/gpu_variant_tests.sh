#!/bin/bash
# GPU parity tests of a variant build: tools/gpu_variant_tests.sh <lib suffix> "<-D switches>"
#   (make BUILD=build/<sfx> LIB=libtrt_b200_<sfx>.so VARIANTS="<-D..>" lib   was run here first)
S=$1; D=$2
TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_$S.so TRT_B200_JIT_DEFS="$D" timeout 600 python -m pytest -m gpu -q tests/test_gpu_native.py tests/test_gpu_jit.py tests/test_gpu_batch.py > gpurun_out/tests_variant_$S.log 2>&1
tail -n 15 gpurun_out/tests_variant_$S.log
