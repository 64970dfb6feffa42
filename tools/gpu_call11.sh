#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests/test_gpu_besteffort.py -m gpu -x -q ) > gpurun_out/c11_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c11_pytest.log
timeout 900 python tools/search_sweep.py --oracle > gpurun_out/c11_search_sweep.jsonl 2> gpurun_out/c11_search_sweep.err
for k in 5 6; do
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c11_dp_n16k${k}.csv python tools/search_one.py 16 $k dp 2 > /dev/null 2>&1
done
