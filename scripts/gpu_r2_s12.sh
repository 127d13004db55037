#!/bin/bash
# Round 2, GPU session 12: scalar-transform kernel A/B with the launches replayed from a CUDA graph (device-bound timing)
mkdir -p gpurun_out
for rep in 1 2; do
for v in "" sb256 s64 sbar128 sold; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python scripts/scalar_bench.py 2>&1 | tail -n 1
done
done
