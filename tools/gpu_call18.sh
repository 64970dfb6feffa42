#!/bin/bash
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_probes.py -m gpu -x -q ) > gpurun_out/c18_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c18_pytest.log
timeout 600 python tools/probe.py --gpus 2 --reps 6 --hbm-bytes 1073741824 > gpurun_out/c18_probe.json 2> gpurun_out/c18_probe.err
