#!/bin/bash
# usage (on the GPU box, via gpurun): tools/ncu_trace.sh <tag> [w h spp conv scene]
# plain run first (must exit 0), then one ncu --set full capture of the trace kernel -> gpurun_out/prof_<tag>.ncu-rep
tag=$1; shift
python tools/profile_frame.py "$@" > gpurun_out/plain_$tag.log 2>&1 && cat gpurun_out/plain_$tag.log &&
ncu --set full --clock-control none --import-source on -k regex:trace_ -s 1 -c 1 -o gpurun_out/prof_$tag python tools/profile_frame.py "$@" > gpurun_out/ncu_$tag.log 2>&1
tail -2 gpurun_out/ncu_$tag.log
