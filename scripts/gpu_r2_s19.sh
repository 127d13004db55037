#!/bin/bash
# Round 2, GPU session 19: 8-deep gather loop (LZB_X_GATHER8) A/B on EfficientZero / Sampled; parity of the variant
mkdir -p gpurun_out
G8="LZB200_LIB=lightzero_b200/liblzb200_g8.so"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$G8" "-" "$G8"
echo "== sez"; AB_ARGS="--variant sez --actions 20" bash scripts/gpu_ab.sh "-" "$G8"
LZB200_LIB=lightzero_b200/liblzb200_g8.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q > gpurun_out/pytest_g8.log 2>&1; echo pytest_g8=$?; tail -n 3 gpurun_out/pytest_g8.log
