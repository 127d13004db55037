#!/bin/bash
# Round 2, GPU session 20: gathered rows stored with the streaming policy (LZB_X_GATHER_STCS) A/B, mz and ez
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_stcs.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$V" "-" "$V"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$V"
