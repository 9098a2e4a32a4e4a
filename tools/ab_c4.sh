#!/bin/bash
# A/B of the tiled large-scene scan on the C4 scene (1920x1080, 4 spp): tools/ab_c4.sh lib1.so lib2.so ...  ("-" = the default library)
for L in "$@"; do
  if [ "$L" = "-" ]; then unset TRT_B200_LIB; else export TRT_B200_LIB=$PWD/terminal-raytracer_b200/$L; fi
  echo "### $L"; timeout 300 python tools/profile_frame.py 1920 1080 4 2 c4 2>&1 | tail -1
done
