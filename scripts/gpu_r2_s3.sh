#!/bin/bash
# Round 2, GPU session 3: the no-allocation entry point's test alone in a fresh process, then the whole GPU suite (which
# ends with it), then one ncu capture of the scalar-transform kernel (after the plain command exited 0).
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc.log 2>&1; echo noalloc_alone=$?; tail -n 4 gpurun_out/pytest_noalloc.log
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 6 gpurun_out/pytest.log
python -m pytest tests --collect-only -q -m gpu 2>/dev/null | tail -n 8
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --categorical"
$CMD > gpurun_out/plain_cat.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_inverse_scalar -s 60 -c 2 -o gpurun_out/prof_scalar_r2 -f $CMD > gpurun_out/ncu_scalar.log 2>&1; echo ncu_scalar=$?
ncu -i gpurun_out/prof_scalar_r2.ncu-rep --page raw --csv > gpurun_out/scalar_ncu_raw.csv 2>/dev/null
