for v in shift r1; do
  D="-DTRT_V_KEY=0 -DTRT_V_SHIFT_T=1 -DTRT_V_NO_RENORM=1"; [ $v = r1 ] && D="-DTRT_V_KEY=0 -DTRT_V_SHIFT_T=0 -DTRT_V_NO_RENORM=0"
  TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_$v.so TRT_B200_JIT_DEFS="$D" timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu --no-strict > gpurun_out/bench_r2_$v.json 2> gpurun_out/bench_r2_$v.err
  TRT_B200_LIB=$PWD/terminal-raytracer_b200/libtrt_b200_$v.so TRT_B200_JIT_DEFS="$D" timeout 300 python -m pytest -m gpu -q tests/test_gpu_native.py tests/test_gpu_jit.py tests/test_gpu_edges.py 2>&1 | tail -3
done
timeout 900 python -m pytest -m gpu -x -q tests/test_gpu_named_configs.py 2>&1 | tail -15
