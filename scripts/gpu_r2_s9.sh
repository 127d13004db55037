#!/bin/bash
# Round 2, GPU session 9: per-level instruction diet A/B (pb_c table, cached child values, sparse ordered sum), parity of each.
mkdir -p gpurun_out
D="-"
L() { echo "LZB200_LIB=lightzero_b200/liblzb200_$1.so"; }
echo "== mz"; bash scripts/gpu_ab.sh "$D" "$(L pbc2)" "$(L cval)" "$(L sparse)" "$(L lean)" "$D" "$(L lean)"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "$D" "$(L lean)"
for v in pbc2 cval sparse lean; do
  LZB200_LIB=lightzero_b200/liblzb200_$v.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q > gpurun_out/pytest_$v.log 2>&1; echo pytest_$v=$?; tail -n 3 gpurun_out/pytest_$v.log
done
