#!/bin/bash
# Round 2, GPU session 8: ncu --set full (with source) of the step kernel and of the scalar-transform kernel, warm launch lists
# (identity and categorical heads), parity of the two experiment builds.
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 2 -o gpurun_out/r02_step -f $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_step=$?
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 300 -c 120 --csv --log-file gpurun_out/r02_launches_warm_L1024.csv $CMD > gpurun_out/ncu_warm.log 2>&1; echo ncu_list=$?
CMDC="$CMD --categorical"
$CMDC > gpurun_out/plain3.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_inverse_scalar -s 30 -c 2 -o gpurun_out/r02_scalar -f $CMDC > gpurun_out/ncu_full_c.log 2>&1; echo ncu_scalar=$?
$CMDC > gpurun_out/plain4.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum --clock-control none --cache-control none -s 400 -c 160 --csv --log-file gpurun_out/r02_launches_warm_categorical.csv $CMDC > gpurun_out/ncu_warm_c.log 2>&1; echo ncu_list_c=$?
for v in ep carry; do
  LZB200_LIB=lightzero_b200/liblzb200_$v.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py -m gpu -x -q > gpurun_out/pytest_$v.log 2>&1; echo pytest_$v=$?; tail -n 3 gpurun_out/pytest_$v.log
done
ls -la gpurun_out/*.ncu-rep
