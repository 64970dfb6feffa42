#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c9_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c9_pytest.log
timeout 900 python tools/search_sweep.py --oracle > gpurun_out/c9_search_sweep.jsonl 2> gpurun_out/c9_search_sweep.err
