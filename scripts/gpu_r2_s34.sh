#!/bin/bash
# Round 2, GPU session 34: libm tables in shared memory for the lane-group kernels (LZB_X_TABLES_IN_SHARED) A/B: mz, gmz, ez
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_shtab.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== gmz"; AB_ARGS="--variant gmz" bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$V"
LZB200_LIB=lightzero_b200/liblzb200_shtab.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
