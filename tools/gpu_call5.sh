#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c5_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c5_pytest.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c5_quick.json 2> gpurun_out/c5_quick.err
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/c5_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c5_smoke.log
ncu --query-metrics 2>/dev/null | grep -i -E "nvl|fabric|pcie" > gpurun_out/c5_ncu_metric_names.txt
