#!/bin/bash
# Warm launch lists (per-kernel durations under ncu) of one eager search of the Sampled, Stochastic, Gumbel and EfficientZero trees
mkdir -p gpurun_out
for v in sez smz gmz ez; do
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --variant $v"
$CMD > gpurun_out/plain_$v.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 400 -c 160 --csv --log-file gpurun_out/ll_$v.csv $CMD > gpurun_out/ncu_$v.log 2>&1; echo $v=$?
done
