#!/bin/bash
# A/B of library builds (tuning): headline bench, value path only, for each lightzero_b200/variants/lib_*.so, two rounds
mkdir -p gpurun_out
for round in 1 2; do
for f in lightzero_b200/variants/lib_*.so; do
  LZB200_LIB=$PWD/$f python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-e2e $EXTRA > gpurun_out/v.json 2> gpurun_out/v.err || tail -n 3 gpurun_out/v.err
  python - "$f" <<PY
import json, sys
d = json.load(open("gpurun_out/v.json"))
print(sys.argv[1].split("/")[-1], "value %.1fM" % (d["value"] / 1e6), "ms", d["ms_per_step"], "step_us", d["roofline"]["avg_launch_us"])
PY
done; done
