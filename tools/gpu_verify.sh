#!/bin/bash
# one-GPU verification under gpurun: GPU suite, smoke, full bench line, ncu --set full of the table kernel, launch list
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/v_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/v_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/v_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/v_smoke.log
( time timeout 1200 python bench.py ) > gpurun_out/v_bench.json 2> gpurun_out/v_bench.err
echo "bench rc=$?" >> gpurun_out/v_bench.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/v_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/v_ncu_records.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/v_launches.csv \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/v_ncu_launches.log 2>&1
