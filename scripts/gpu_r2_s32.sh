#!/bin/bash
# Round 2, GPU session 32: two more cache policies for the row loads (ld.global.lu "last use", ld.global.cg "L2 only")
mkdir -p gpurun_out
A="LZB200_LIB=lightzero_b200/liblzb200_ldlu.so"; B="LZB200_LIB=lightzero_b200/liblzb200_ldcg.so"
echo "== mz"; bash scripts/gpu_ab.sh "-" "$A" "$B" "-"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "-" "$A" "$B"
