#!/bin/bash
# Round 2, GPU session 44: lone walker + staged sibling row / 32-lane walker gather (LZB_X_LONE_STAGE), pb_c table on top of the
# lone walker (LZB_X_PBC2), both; parity of the staged variant
mkdir -p gpurun_out
S="LZB200_LIB=lightzero_b200/liblzb200_stage.so"
P="LZB200_LIB=lightzero_b200/liblzb200_pbc2.so"
SP="LZB200_LIB=lightzero_b200/liblzb200_stagepbc2.so"
echo "== mz";  bash scripts/gpu_ab.sh "-" "$S" "$P" "$SP" "-" "$S"
LZB200_LIB=lightzero_b200/liblzb200_stage.so timeout 900 python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
