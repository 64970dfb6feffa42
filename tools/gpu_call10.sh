#!/bin/bash
mkdir -p gpurun_out
for k in 5 6; do
  timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c10_dp_n16k${k}_warp.csv python tools/search_one.py 16 $k dp 2 > /dev/null 2>&1
  GPA_DP_SHARE=1 timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/c10_dp_n16k${k}_share.csv python tools/search_one.py 16 $k dp 2 > /dev/null 2>&1
done
