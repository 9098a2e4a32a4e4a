#!/bin/bash
# A/B of the chain-pool machine (trt_pool.cuh) against trace_native_kernel: tools/ab_pool.sh [lib]
# frame times on the Cornell 1920x1080 x 256 spp frame in both arithmetic conventions, then the parity tests with the pool forced on
L=${1:-}
[ -n "$L" ] && export TRT_B200_LIB=$PWD/terminal-raytracer_b200/$L
for conv in 2 0; do
  for p in 0 1; do
    echo "### conv $conv TRT_B200_POOL=$p"; TRT_B200_POOL=$p timeout 120 python tools/profile_frame.py 1920 1080 256 $conv cornell 2>&1 | tail -1
  done
done
echo "### tests with TRT_B200_POOL=1"
TRT_B200_POOL=1 timeout 900 python -m pytest -q -x -m gpu tests/test_gpu_parity.py tests/test_gpu_native.py tests/test_gpu_named_configs.py tests/test_gpu_batch.py 2>&1 | tail -15
