#!/bin/bash
# Round 2, GPU session 35: Gumbel mixed-value sums over the visited lanes only (LZB_X_GMZ_SPARSE) A/B + parity
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_gsp.so"
echo "== gmz"; AB_ARGS="--variant gmz" bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
LZB200_LIB=lightzero_b200/liblzb200_gsp.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q -k "gumbel or gmz or Gumbel" 2>&1 | tail -n 3
