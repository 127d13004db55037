#!/bin/bash
# Round profile: headline bench, then (after the plain command exited 0) one ncu --set full capture of the step kernel
# and warm launch lists for L=1024 and L=0.  Outputs in gpurun_out/.
mkdir -p gpurun_out
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench=$?
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 3 -o gpurun_out/prof_step -f $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 300 -c 120 --csv --log-file gpurun_out/launches_warm_L1024.csv $CMD > gpurun_out/ncu_warm.log 2>&1; echo list_L1024=$?
$CMD --latent 0 > gpurun_out/plain3.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 300 -c 120 --csv --log-file gpurun_out/launches_warm_L0.csv $CMD --latent 0 > gpurun_out/ncu_warm0.log 2>&1; echo list_L0=$?
cat gpurun_out/bench.json
