#!/bin/bash
# A/B of the strict (conv 0) kernels on the Cornell 1920x1080 x 256 spp frame: tools/ab_strict.sh
for v in "TRT_B200_STRICT=wave" "TRT_B200_STRICT=generic" "X=1" "TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_s5.so" "TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_s4.so"; do
  echo "### $v"; env $v timeout 120 python tools/ab_conv.py 1920 1080 256 0 2>&1 | tail -1
done
