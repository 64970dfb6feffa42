#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c12_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c12_pytest.log
for v in 0 1 2 3; do
  GPA_TAB_VARIANT=$v timeout 600 python bench.py --quick --steps 30 --warmup 3 > gpurun_out/c12_quick_v$v.json 2> gpurun_out/c12_quick_v$v.err
done
GPA_TAB_VARIANT=2 timeout 900 python -m pytest tests/test_gpu_table.py tests/test_gpu_besteffort.py -m gpu -x -q > gpurun_out/c12_pytest_v2.log 2>&1
