#!/bin/bash
# Round 2, GPU session 47: trees that end together split between the bulk engine and a 32-lane gather (LZB_X_PAIR_SPLIT); the lone
# walker for Gumbel / stochastic / reuse again now that a 32-lane reduction is one CREDUX (LZB_X_LONE_WALKER_ALL)
mkdir -p gpurun_out
S="LZB200_LIB=lightzero_b200/liblzb200_split.so"
L="LZB200_LIB=lightzero_b200/liblzb200_loneall.so"
echo "== mz";  timeout 600 bash scripts/gpu_ab.sh "-" "$S" "-" "$S"
echo "== gmz"; AB_ARGS="--variant gmz" timeout 600 bash scripts/gpu_ab.sh "-" "$L" "-" "$L"
echo "== smz"; AB_ARGS="--variant smz" timeout 600 bash scripts/gpu_ab.sh "-" "$L"
LZB200_LIB=lightzero_b200/liblzb200_split.so timeout 900 python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
