#!/bin/bash
# Round 2, GPU session 18: ncu --set full (with source) of the Gumbel and EfficientZero step kernels at the headline shape
mkdir -p gpurun_out
for v in gmz ez; do
  CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --variant $v"
  $CMD > gpurun_out/plain_$v.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 2 -o gpurun_out/r02_step_$v -f $CMD > gpurun_out/ncu_full_$v.log 2>&1; echo ncu_$v=$?
done
ls -la gpurun_out/*.ncu-rep
