#!/bin/bash
mkdir -p gpurun_out
for z in 0 1 2; do
  GPA_ZERO_COPY=$z timeout 600 python bench.py --quick --steps 30 --warmup 3 > gpurun_out/c15_quick_z$z.json 2> gpurun_out/c15_quick_z$z.err
done
