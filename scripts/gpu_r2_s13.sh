#!/bin/bash
# Round 2, GPU session 13: two-rows-per-warp packed scalar kernel vs the one-row kernel (graph-replayed launches), tests per build
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" onerow pair4 pair128; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python scripts/scalar_bench.py 2>&1 | tail -n 1
done
done
for v in "" pair4 pair128; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python -m pytest tests/test_scaling.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 3
done
