#!/bin/bash
# Round 2, GPU session 43: lone walker + L1 prefetch of the expanded children's blocks (LZB_X_LONE_PREFETCH) A/B; ez, sez too
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_lpf.so"
echo "== mz";  bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== sez"; AB_ARGS="--variant sez --actions 20" bash scripts/gpu_ab.sh "-" "$V"
LZB200_LIB=lightzero_b200/liblzb200_lpf.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py -m gpu -x -q 2>&1 | tail -n 2
