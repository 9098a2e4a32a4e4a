#!/bin/bash
# Final one-GPU measurements of the round: C3 bench line (default command), the other configurations, ncu evidence
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; tail -c 400 gpurun_out/bench_final_n1.json; echo
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_final_ref.json 2> gpurun_out/bench_final_ref.err; tail -c 300 gpurun_out/bench_final_ref.json; echo
bash tools/r2_configs.sh
