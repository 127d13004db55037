#!/bin/bash
# warm-cache launch list (per-kernel durations) of one eager search; EXTRA = extra bench flags
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e $EXTRA"
$CMD > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 300 -c 120 --csv --log-file gpurun_out/launches_warm_$TAG.csv $CMD > gpurun_out/ncu_warm.log 2>&1; echo ncu_list_warm=$?
