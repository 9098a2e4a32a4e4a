#!/bin/bash
# One GPU call: A/B of the guarded kernel variants through the run-time compiler, then the GPU tests and the bench
# line on the variant build (libtrt_b200_next.so).  Everything lands in gpurun_out/ as it goes.
mkdir -p gpurun_out; O=gpurun_out/final_ab.log; : > $O
export AB_REF=/tmp/ab_ref.npy
ab() { echo "### DEFS='$1'" >> $O; TRT_B200_JIT=force TRT_B200_JIT_DEFS="$1" timeout 120 python tools/ab_jit.py cornell 256 >> $O 2>&1; }
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv,noheader >> $O
ab ""
ab "-DTRT_V_CAMFOLD=1"
ab "-DTRT_V_EXTRA_NOWAIT=1"
ab "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1"
ab "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1 -DTRT_NV_WAIT_W=1"
ab "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1 -DTRT_NV_WAIT_W=3"
ab "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1 -DTRT_NV_WAIT_W=4"
ab ""
abs() { echo "### $1 DEFS='$2'" >> $O; AB_REF=/tmp/ab_ref_$1.npy TRT_B200_JIT_DEFS="$2" timeout 120 python tools/ab_jit.py $1 64 >> $O 2>&1; }
abs scene2 ""
abs scene2 "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1"
abs demo ""
abs demo "-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1"
export TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_next.so
export TRT_B200_JIT_DEFS="-DTRT_V_CAMFOLD=1 -DTRT_V_EXTRA_NOWAIT=1"
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --no-strict > gpurun_out/final_bench_next.json 2> gpurun_out/final_bench_next.err
timeout 420 python -m pytest -m gpu -x -q tests/test_gpu_native.py tests/test_gpu_jit.py tests/test_gpu_edges.py tests/test_gpu_batch.py > gpurun_out/final_tests_next_a.log 2>&1
timeout 420 python -m pytest tests -m gpu -x -q --deselect tests/test_gpu_native.py --deselect tests/test_gpu_jit.py --deselect tests/test_gpu_edges.py --deselect tests/test_gpu_batch.py > gpurun_out/final_tests_next_b.log 2>&1
tail -3 gpurun_out/final_tests_next_a.log gpurun_out/final_tests_next_b.log; cat $O
