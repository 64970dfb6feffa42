#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c3_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c3_pytest.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c3_quick.json 2> gpurun_out/c3_quick.err
echo "quick rc=$?" >> gpurun_out/c3_quick.err
timeout 300 ./tools/bin/latency > gpurun_out/c3_latency.json 2> gpurun_out/c3_latency.err
( time timeout 1200 python bench.py ) > gpurun_out/c3_bench.json 2> gpurun_out/c3_bench.err
echo "bench rc=$?" >> gpurun_out/c3_bench.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c3_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c3_ncu_records.log 2>&1
nproc > gpurun_out/c3_nproc.txt; nvidia-smi topo -m > gpurun_out/c3_topo.txt 2>&1
