#!/bin/bash
# Round 2, GPU session 42: lone walker as the default for every lane-group tree (pUCT incl. reuse / stochastic, Gumbel):
# A/B against -DLZB_X_NO_LONE_WALKER, then the full GPU suite on the default build
mkdir -p gpurun_out
V="LZB200_LIB=lightzero_b200/liblzb200_nolone.so"
echo "== mz";  bash scripts/gpu_ab.sh "$V" "-"
echo "== gmz"; AB_ARGS="--variant gmz" bash scripts/gpu_ab.sh "$V" "-"
echo "== smz"; AB_ARGS="--variant smz" bash scripts/gpu_ab.sh "$V" "-"
echo "== ez";  AB_ARGS="--variant ez" bash scripts/gpu_ab.sh "$V" "-"
python -m pytest tests -m gpu -x -q 2>&1 | tail -n 4
