#!/bin/bash
# usage (here, no GPU): tools/ncu_report.sh <tag> <kernel-substring> [top_n]  -> per-line table + key raw metrics
tag=$1; k=$2; top=${3:-40}
d=$(mktemp -d); ( cd $d && cuobjdump -xelf all /root/repo/terminal-raytracer_b200/libtrt_b200.so >/dev/null && for c in trt_api trt_native; do nvdisasm -g -c $c.sm_100a.cubin; done > all.dis 2>&1 )
ncu -i gpurun_out/prof_$tag.ncu-rep --page source --csv 2>/dev/null > $d/src.csv
ncu -i gpurun_out/prof_$tag.ncu-rep --page raw --csv 2>/dev/null > $d/raw.csv
python - $d/raw.csv <<'PY'
import csv,sys
rows=list(csv.reader(open(sys.argv[1])))
hdr,units,vals=rows[0],rows[1],rows[2]
want=['gpu__time_duration.sum','smsp__inst_executed.sum ','smsp__thread_inst_executed_per_inst_executed.ratio','smsp__issue_active.avg.pct_of_peak_sustained_active','sm__warps_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread ','sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active','sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active','sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum ','dram__bytes_write.sum ','smsp__inst_executed.sum','sm__throughput.avg.pct_of_peak_sustained_elapsed']
for h,u,v in zip(hdr,units,vals):
    if any(h==w.strip() for w in want): print(f"{h} [{u}] = {v}")
PY
python /root/repo/tools/ncu_by_line.py $d/src.csv $d/all.dis "$k" $top
rm -rf $d
