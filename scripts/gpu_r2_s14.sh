#!/bin/bash
# Round 2, GPU session 14: balanced grid (P full passes) for the scalar kernels: one-row (default) vs unbalanced vs the pair kernel
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" unbal pairb pairb4; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python scripts/scalar_bench.py 2>&1 | tail -n 1
done
done
for v in "" pairb; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python -m pytest tests/test_scaling.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 3
done
