#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c4_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c4_pytest.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c4_quick.json 2> gpurun_out/c4_quick.err
echo "quick rc=$?" >> gpurun_out/c4_quick.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c4_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c4_ncu_records.log 2>&1
timeout 120 ./tools/bin/nvml_reenum 30 > gpurun_out/c4_nvml_reenum.json 2> gpurun_out/c4_nvml_reenum.err
