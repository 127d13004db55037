#!/bin/bash
# Round 2, GPU session 25: Gumbel prior-softmax constants memoised per node (LZB_X_GMZ_STATS) A/B + parity
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_gst.so"
echo "== gmz"; AB_ARGS="--variant gmz" bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
LZB200_LIB=lightzero_b200/liblzb200_gst.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py tests/test_gumbel_helpers.py -m gpu -x -q > gpurun_out/pytest_gst.log 2>&1; echo pytest_gst=$?; tail -n 3 gpurun_out/pytest_gst.log
