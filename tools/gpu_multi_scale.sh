#!/bin/bash
# 8-GPU refresh: probe test + matrices incl. all-at-once, N=4 and N=8 bench
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_probes.py -m gpu -x -q ) > gpurun_out/m8b_pytest_probes.log 2>&1
echo "pytest rc=$?" >> gpurun_out/m8b_pytest_probes.log
timeout 900 python tools/probe.py --gpus 8 --reps 6 --hbm-bytes 2147483648 > gpurun_out/m8b_probe.json 2> gpurun_out/m8b_probe.err
for N in 4 8; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2955$N \
  bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/m8b_bench_n$N.json 2> gpurun_out/m8b_bench_n$N.err
echo "bench rc=$?" >> gpurun_out/m8b_bench_n$N.err
done
