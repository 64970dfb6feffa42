#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c13_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c13_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/c13_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c13_smoke.log
( time timeout 1200 python bench.py ) > gpurun_out/c13_bench.json 2> gpurun_out/c13_bench.err
echo "bench rc=$?" >> gpurun_out/c13_bench.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c13_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c13_ncu_records.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c13_launches.csv \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c13_ncu_launches.log 2>&1
