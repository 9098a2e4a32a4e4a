set -x
timeout 900 python -m pytest -q tests/test_gpu_multi.py -m gpu > gpurun_out/multi8_tests.log 2>&1; tail -3 gpurun_out/multi8_tests.log
E=terminal-raytracer_b200/terminal-raytracer; S=tests/golden/scenes/scene2.json
timeout 120 $E --path $S --strict --quiet --seed 77 --frames 20 --size 120x40 --out /tmp/serial.bin && timeout 120 $E --path $S --strict --quiet --seed 77 --frames 20 --size 120x40 --gpus 2 --out /tmp/g2.bin && timeout 120 $E --path $S --strict --quiet --seed 77 --frames 20 --size 120x40 --gpus 8 --out /tmp/g8.bin && cmp /tmp/serial.bin /tmp/g2.bin && cmp /tmp/serial.bin /tmp/g8.bin && echo "CLI_REAL_MULTI_GPU_STREAMS_EQUAL $(stat -c %s /tmp/serial.bin) bytes" | tee gpurun_out/cli_real_multi.txt
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2950$n bench.py --gpus $n --steps 5 --warmup 3 > gpurun_out/bench_final_n$n.json 2> gpurun_out/bench_final_n$n.err
  tail -c 300 gpurun_out/bench_final_n$n.json; echo
done
