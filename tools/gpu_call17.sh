#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c17_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c17_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29556 \
  bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c17_bench_n2.json 2> gpurun_out/c17_bench_n2.err
echo "bench rc=$?" >> gpurun_out/c17_bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29557 \
  bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/c17_bench_ref_n2.json 2> gpurun_out/c17_bench_ref_n2.err
echo "ref rc=$?" >> gpurun_out/c17_bench_ref_n2.err
