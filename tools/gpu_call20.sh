#!/bin/bash
mkdir -p gpurun_out
for v in 0 1; do
  GPA_TAB_VARIANT=$v timeout 600 python bench.py --quick --steps 30 --warmup 3 > gpurun_out/c20_quick_v$v.json 2> gpurun_out/c20_quick_v$v.err
done
GPA_TAB_VARIANT=1 timeout 900 python -m pytest tests/test_gpu_table.py tests/test_gpu_besteffort.py -m gpu -x -q > gpurun_out/c20_pytest_v1.log 2>&1
