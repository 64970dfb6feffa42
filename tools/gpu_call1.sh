#!/bin/bash
# round-2 GPU call 1: GPU suite, quick bench, launch list, ncu --set full of the table kernels
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/c1_gpus.txt 2>&1
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c1_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c1_pytest.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c1_quick.json 2> gpurun_out/c1_quick.err
echo "quick rc=$?" >> gpurun_out/c1_quick.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c1_launches.csv \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c1_ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c1_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c1_ncu_records.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 87 -c 1 -f -o gpurun_out/c1_table_masks \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c1_ncu_masks.log 2>&1
ls -la gpurun_out
