#!/bin/bash
# Round 2, GPU session 21: row-load cache policy A/B (streaming = default, plain, read-only path), mz and ez
mkdir -p gpurun_out
A="LZB200_LIB=lightzero_b200/liblzb200_ldg.so"; B="LZB200_LIB=lightzero_b200/liblzb200_ldnc.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$A" "$B" "-"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$A" "$B"
