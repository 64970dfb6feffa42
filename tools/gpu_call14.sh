#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_plugin.py -m gpu -x -q ) > gpurun_out/c14_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c14_pytest.log
for r in 1 4; do ./tools/bin/stream $r 100000 > gpurun_out/c14_stream_$r.json 2>&1; done
