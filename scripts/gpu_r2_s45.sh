#!/bin/bash
# Round 2, GPU session 45: group reductions through redux.sync / CREDUX (LZB_X_REDUX) for mz, gmz, ez; the finished tree's row staged
# by the bulk-copy engine (LZB_X_LONE_STAGE + LZB_X_LONE_STAGE_BULK); parity of both variants
mkdir -p gpurun_out
R="LZB200_LIB=lightzero_b200/liblzb200_redux.so"
B="LZB200_LIB=lightzero_b200/liblzb200_stagebulk.so"
echo "== mz";  timeout 600 bash scripts/gpu_ab.sh "-" "$R" "$B" "-" "$R"
echo "== gmz"; AB_ARGS="--variant gmz" timeout 600 bash scripts/gpu_ab.sh "-" "$R" "-" "$R"
echo "== ez";  AB_ARGS="--variant ez" timeout 600 bash scripts/gpu_ab.sh "-" "$R"
for v in redux stagebulk; do
  LZB200_LIB=lightzero_b200/liblzb200_$v.so timeout 900 python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_api.py -m gpu -x -q 2>&1 | tail -n 3
done
