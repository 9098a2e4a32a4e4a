#!/bin/bash
# GPU tests + C3 bench line on an A/B build of the library: tools/gpu_check_next.sh <lib.so> "<variant defines>"
export TRT_B200_LIB=$PWD/terminal-raytracer_b200/$1
export TRT_B200_JIT_DEFS="$2"
timeout 300 python -m pytest -m gpu -x -q tests > gpurun_out/tests_${1%.so}.log 2>&1; tail -n 3 gpurun_out/tests_${1%.so}.log
timeout 200 python bench.py --steps 3 --warmup 3 --no-cpu --no-strict > gpurun_out/bench_${1%.so}.json 2> gpurun_out/bench_${1%.so}.err; cut -c1-120 gpurun_out/bench_${1%.so}.json
