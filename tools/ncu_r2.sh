#!/bin/bash
# Round-2 ncu evidence in one GPU call (every command first runs plain and must exit 0):  bash tools/ncu_r2.sh
#   1. --set full of the production trace kernel, Cornell 1920x1080 x 256 spp (2nd frame)      -> prof_r2.ncu-rep
#   2. --set full of the one-pass quantiser on the same frame                                  -> prof_r2_quant.ncu-rep
#   3. DRAM bytes of the trace kernel at C3 (3840x2160 x 1024 spp), metrics only               -> traffic_r2_c3.csv
#   4. launch list of the bench command                                                        -> launches_r2.csv
set -x
python tools/profile_frame.py 1920 1080 256 2 cornell > gpurun_out/plain_r2.log 2>&1 || exit 1
cat gpurun_out/plain_r2.log
ncu --set full --clock-control none --import-source on -k regex:trace_native -s 1 -c 1 -o gpurun_out/prof_r2 -f python tools/profile_frame.py 1920 1080 256 2 cornell > gpurun_out/ncu_r2.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:quant_fused -s 1 -c 1 -o gpurun_out/prof_r2_quant -f python tools/profile_frame.py 3840 2160 4 2 cornell > gpurun_out/ncu_r2_quant.log 2>&1
ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:trace_native -s 1 -c 1 --csv --log-file gpurun_out/traffic_r2_c3.csv python tools/profile_frame.py 3840 2160 1024 2 cornell > gpurun_out/ncu_r2_c3.log 2>&1
python bench.py --steps 2 --warmup 3 --no-cpu --no-strict --no-parity > gpurun_out/bench_r2_plain.json 2> gpurun_out/bench_r2_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r2.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-strict --no-parity > gpurun_out/ncu_bench_r2.log 2>&1
tail -2 gpurun_out/ncu_r2.log gpurun_out/ncu_r2_quant.log gpurun_out/ncu_r2_c3.log
