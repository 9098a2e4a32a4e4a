for l in "" q26 q28 q46 q18; do
  echo "### ${l:-default}"; L=""; [ -n "$l" ] && L="TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_$l.so"
  env $L python tools/profile_frame.py 3840 2160 64 2 cornell | cut -c1-110
  env $L ncu --metrics gpu__time_duration.sum --clock-control none -k regex:quant_fused -s 1 -c 1 python tools/profile_frame.py 3840 2160 64 2 cornell 2>&1 | grep -i "duration"
done
