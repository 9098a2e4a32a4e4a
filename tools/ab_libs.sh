#!/bin/bash
# A/B of library builds on one frame: tools/ab_libs.sh "<w h spp conv scene>" lib1.so lib2.so ...  ("-" = the default library)
F=$1; shift
for L in "$@"; do
  if [ "$L" = "-" ]; then unset TRT_B200_LIB; else export TRT_B200_LIB=$PWD/terminal-raytracer_b200/$L; fi
  echo "### $L"; timeout 300 python tools/profile_frame.py $F 2>&1 | tail -1
done
