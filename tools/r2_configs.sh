#!/bin/bash
# Final-build lines of the other BASELINE configurations (one GPU):  bash tools/r2_configs.sh   -> gpurun_out/r2_bench_*.json
python bench.py --workload c1 --steps 20 --warmup 5 > gpurun_out/r2_bench_c1.json 2> gpurun_out/r2_bench_c1.err; cut -c1-120 gpurun_out/r2_bench_c1.json
python bench.py --workload c2 --steps 100 --warmup 5 > gpurun_out/r2_bench_c2.json 2> gpurun_out/r2_bench_c2.err; cut -c1-120 gpurun_out/r2_bench_c2.json
python tools/bench_c2_batch.py > gpurun_out/r2_bench_c2_batch.json 2>&1; tail -2 gpurun_out/r2_bench_c2_batch.json | cut -c1-300
python tools/bench_c5.py 600 2 > gpurun_out/r2_bench_c5_native.json 2>&1; tail -1 gpurun_out/r2_bench_c5_native.json | cut -c1-300
python tools/bench_c5.py 600 0 > gpurun_out/r2_bench_c5_strict.json 2>&1; tail -1 gpurun_out/r2_bench_c5_strict.json | cut -c1-300
python tools/bench_c5_pipeline.py > gpurun_out/r2_bench_c5_pipeline.json 2>&1; tail -1 gpurun_out/r2_bench_c5_pipeline.json | cut -c1-400
timeout 900 python bench.py --workload c4 --steps 1 --warmup 3 --no-strict --count-stride 216 --cpu-seconds 10 > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err; cut -c1-200 gpurun_out/r2_bench_c4.json
python tools/bench_c4_accel.py > gpurun_out/r2_bench_c4_accel.json 2>&1; tail -2 gpurun_out/r2_bench_c4_accel.json | cut -c1-300
