#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c6_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c6_pytest.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c6_quick.json 2> gpurun_out/c6_quick.err
GPA_ZERO_COPY=1 timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c6_quick_zerocopy.json 2> gpurun_out/c6_quick_zerocopy.err
