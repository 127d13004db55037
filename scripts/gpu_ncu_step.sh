#!/bin/bash
# One ncu --set full capture of the fused step kernel (3 launches mid-search) + a warm launch list; each after the plain command exits 0.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e $EXTRA"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 3 -o gpurun_out/prof_step -f $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
$CMD > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --cache-control none -s 300 -c 200 --csv --log-file gpurun_out/launches_warm.csv $CMD > gpurun_out/ncu_warm.log 2>&1; echo ncu_list_warm=$?
