#!/bin/bash
# Round 2, GPU session 37: e2e leg against the number of chunks / streams of the driver (search_streams)
mkdir -p gpurun_out
for st in 2 3 4 5 6 8; do
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-streams $st > gpurun_out/e2e_$st.json 2> gpurun_out/e2e_$st.err
  python - $st gpurun_out/e2e_$st.json <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[2]) if l.startswith("{")][-1])
    print("e2e streams %s: e2e %.1fM ms %.4f lists %.3f ms" % (sys.argv[1], d["e2e"]["value"] / 1e6, d["e2e"]["ms_per_step"], d["e2e_lists"]["ms_per_step"]))
except Exception as e:
    print(sys.argv[1], "ERR", e)
PY
done
