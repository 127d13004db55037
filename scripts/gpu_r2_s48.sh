#!/bin/bash
# Round 2, GPU session 48 (last seconds of the budget): programmatic launch trigger issued by a lone-walker warp as soon as its walk
# ends, before its row work (LZB_X_LONE_EARLY_RELEASE), headline shape only
mkdir -p gpurun_out
for rep in 1 2; do
for lib in "" "lightzero_b200/liblzb200_early.so"; do
  LZB200_LIB=$lib timeout 40 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab48.json 2> gpurun_out/ab48.err
  python - "$lib" <<'PY'
import json, sys
d = json.loads([l for l in open("gpurun_out/ab48.json") if l.startswith("{")][-1])
print("%-40s value %.1fM ms %.4f step_us %s" % (sys.argv[1] or "default", d["value"] / 1e6, d["ms_per_step"], d["roofline"]["avg_launch_us"]))
PY
done
done
