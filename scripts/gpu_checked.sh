#!/bin/bash
# Round 2 hygiene: compute-sanitizer is closed on this GPU pool, so the fused parity cases run under a CHECKED build instead
# (-DLZB_CHECKED: every data-dependent index of the tree kernels is tested against its bound, the violation code is read at
# the end of every search and raises; lightzero_b200/forest.py, LZB_CHECKED=1).  Log -> gpurun_out/checked.log
mkdir -p gpurun_out
export LZB200_LIB=lightzero_b200/liblzb200_checked.so LZB_CHECKED=1
python -m pytest tests/test_gpu_checked_build.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_headline_parity.py \
  tests/test_gpu_api.py tests/test_gpu_callers.py tests/test_gpu_collect.py tests/test_gpu_tiebreak.py -m gpu -q -rs -x > gpurun_out/checked.log 2>&1
echo checked=$?; tail -n 8 gpurun_out/checked.log
