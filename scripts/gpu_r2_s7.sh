#!/bin/bash
# Round 2, GPU session 7: step-kernel chain experiments A/B (early path-record loads; backup -> select register carry),
# parity of each variant at the headline shape.
mkdir -p gpurun_out
D="-"
EP="LZB200_LIB=lightzero_b200/liblzb200_ep.so"
CA="LZB200_LIB=lightzero_b200/liblzb200_carry.so"
echo "== mz"; bash scripts/gpu_ab.sh "$D" "$EP" "$CA" "$D" "$CA"
echo "== ez"; AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "$D" "$EP" "$CA"
echo "== two-player mz"; AB_ARGS="--two-player" bash scripts/gpu_ab.sh "$D" "$CA"
for v in ep carry; do
  LZB200_LIB=lightzero_b200/liblzb200_$v.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py -m gpu -x -q > gpurun_out/pytest_$v.log 2>&1; echo pytest_$v=$?; tail -n 3 gpurun_out/pytest_$v.log
done
