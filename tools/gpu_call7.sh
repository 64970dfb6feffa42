#!/bin/bash
mkdir -p gpurun_out
( time timeout 1200 python -m pytest tests -m gpu -x -q ) > gpurun_out/c7_pytest.log 2>&1
echo "pytest rc=$?" >> gpurun_out/c7_pytest.log
( time timeout 1200 python bench.py ) > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err
echo "bench rc=$?" >> gpurun_out/c7_bench.err
( time timeout 900 python bench.py --impl reference --steps 10 --warmup 1 ) > gpurun_out/c7_bench_ref.json 2> gpurun_out/c7_bench_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/c7_launches.csv \
  python bench.py --quick --steps 2 --warmup 3 > gpurun_out/c7_ncu_launches.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:besteffort_search -s 1 -c 1 -f -o gpurun_out/c7_search_n16k3 \
  python tools/search_one.py 16 3 enum 3 > gpurun_out/c7_ncu_search.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:subset_batch -s 2 -c 1 -f -o gpurun_out/c7_subset_batch \
  python tools/subset_bench.py > gpurun_out/c7_ncu_subset.log 2>&1
