#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c16_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c16_pytest.log
for z in 0 1 2; do
  GPA_ZERO_COPY=$z timeout 600 python bench.py --quick --steps 30 --warmup 3 > gpurun_out/c16_quick_z$z.json 2> gpurun_out/c16_quick_z$z.err
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c16_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c16_ncu_records.log 2>&1
