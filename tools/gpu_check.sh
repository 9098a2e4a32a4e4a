#!/bin/bash
# GPU tests, the A/B frame (Cornell 1920x1080 256 spp) and the C3 bench line of the current build: tools/gpu_check.sh <tag>
T=${1:-cur}
timeout 300 python -m pytest -m gpu -x -q tests > gpurun_out/tests_$T.log 2>&1; tail -n 3 gpurun_out/tests_$T.log
timeout 100 python tools/ab_jit.py cornell 256 2>&1 | tee gpurun_out/ab_$T.log
TRT_B200_CTAS=8 timeout 100 python tools/ab_jit.py cornell 256 2>&1 | tee -a gpurun_out/ab_$T.log
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --no-strict > gpurun_out/bench_$T.json 2> gpurun_out/bench_$T.err; cut -c1-130 gpurun_out/bench_$T.json
