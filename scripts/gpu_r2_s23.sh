#!/bin/bash
# Round 2, GPU session 23: the row prefetch issued from inside the descending branch
mkdir -p gpurun_out
L() { echo "LZB200_LIB=lightzero_b200/liblzb200_$1.so"; }
echo "== mz"; bash scripts/gpu_ab.sh "-" "$(L pf1b)" "$(L pf1)" "-" "$(L pf1b)"
