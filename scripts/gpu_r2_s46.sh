#!/bin/bash
# Round 2, GPU session 46: new defaults (redux reductions + bulk-staged lone-walker rows) against each one switched off and against
# both off, same box; the other flavours; the whole GPU suite on the new default; the checked build
mkdir -p gpurun_out
B="LZB200_LIB=lightzero_b200/liblzb200_before.so"
NR="LZB200_LIB=lightzero_b200/liblzb200_noredux.so"
NS="LZB200_LIB=lightzero_b200/liblzb200_nostage.so"
echo "== mz";  timeout 600 bash scripts/gpu_ab.sh "-" "$B" "$NR" "$NS" "-" "$B"
echo "== ez";  AB_ARGS="--variant ez" timeout 600 bash scripts/gpu_ab.sh "-" "$B"
echo "== gmz"; AB_ARGS="--variant gmz" timeout 600 bash scripts/gpu_ab.sh "-" "$B"
echo "== sez"; AB_ARGS="--variant sez --actions 20" timeout 600 bash scripts/gpu_ab.sh "-" "$B"
echo "== smz"; AB_ARGS="--variant smz" timeout 600 bash scripts/gpu_ab.sh "-" "$B"
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 3
timeout 900 bash scripts/gpu_checked.sh
