#!/bin/bash
# Round 2, GPU session 22: leaf-row L2 prefetch on a certain walk end (LZB_X_ROW_PREFETCH = max visit count of the entered node)
mkdir -p gpurun_out
L() { echo "LZB200_LIB=lightzero_b200/liblzb200_$1.so"; }
echo "== mz"; bash scripts/gpu_ab.sh "-" "$(L pf1)" "$(L pf2)" "$(L pf3)" "-" "$(L pf1)"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$(L pf1)" "$(L pf2)"
LZB200_LIB=lightzero_b200/liblzb200_pf1.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q > gpurun_out/pytest_pf1.log 2>&1; echo pytest_pf1=$?; tail -n 3 gpurun_out/pytest_pf1.log
