#!/bin/bash
# Round 2, GPU session 29: L2 prefetch of the later rounds of the leaf-row gather (LZB_X_GATHER_PREFETCH) A/B, mz / ez / sez
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_gpf.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== sez"; AB_ARGS="--variant sez --actions 20" bash scripts/gpu_ab.sh "-" "$V"
LZB200_LIB=lightzero_b200/liblzb200_gpf.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
