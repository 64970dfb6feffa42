#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c8_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c8_pytest.log
python -c "import __graft_entry__ as e; e.smoke()" > gpurun_out/c8_smoke.log 2>&1
echo "smoke rc=$?" >> gpurun_out/c8_smoke.log
timeout 600 python bench.py --quick --steps 20 --warmup 3 > gpurun_out/c8_quick.json 2> gpurun_out/c8_quick.err
timeout 900 ncu --set full --clock-control none --import-source on -k regex:besteffort_batch_table -s 3 -c 1 -f -o gpurun_out/c8_table_records \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c8_ncu_records.log 2>&1
