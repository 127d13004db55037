#!/bin/bash
# Round 2, GPU session 26: launch lists of the Sampled and Stochastic searches (per-kernel durations)
mkdir -p gpurun_out
for v in sez smz; do
  EX=""; [ "$v" = sez ] && EX="--actions 20"
  CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --variant $v $EX"
  $CMD > gpurun_out/plain_$v.log 2>&1 && \
  ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 300 -c 200 --csv --log-file gpurun_out/r02_launches_warm_$v.csv $CMD > gpurun_out/ncu_warm_$v.log 2>&1; echo list_$v=$?
done
