#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_subset.py -m gpu -x -q ) > gpurun_out/c22_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c22_pytest.log
for r in 4 16 32; do timeout 300 python tools/subset_bench.py --replicas $r --reps 4 > gpurun_out/c22_subset_$r.log 2>&1; done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:subset_batch -s 2 -c 1 -f -o gpurun_out/c22_subset_batch \
  python tools/subset_bench.py > gpurun_out/c22_ncu_subset.log 2>&1
