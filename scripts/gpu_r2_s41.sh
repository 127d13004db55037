#!/bin/bash
# Round 2, GPU session 41: "lone walker" - a tree whose warp sibling has finished walks on with all 32 lanes, one record per lane
# (LZB_X_LONE_WALKER).  First the refactored default (level body as a generic lambda) for parity and speed, then the variant.
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_lone.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$V"
echo "== sez"; AB_ARGS="--variant sez --actions 20" bash scripts/gpu_ab.sh "-" "$V"
python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
LZB200_LIB=lightzero_b200/liblzb200_lone.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 5
