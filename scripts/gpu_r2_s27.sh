#!/bin/bash
# Round 2, GPU session 27: continuous sampler with both Box-Muller variates per uniform pair (default) vs one (LZB_X_SAMPLER_ONE_PER_PAIR)
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_smp1.so"
echo "== sez"; AB_ARGS="--variant sez --actions 20" bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
python -m pytest tests/test_gpu_sampler.py tests/test_gpu_api.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py -m gpu -x -q 2>&1 | tail -n 3
