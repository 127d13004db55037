#!/bin/bash
# Round 2, GPU session 30: chunks of the batch on concurrent streams (bench --streams) for EfficientZero, whose step is half
# row gather (HBM-bound) and half tree work (latency-bound) - do the two halves of the machine's time overlap across chunks?
mkdir -p gpurun_out
for v in ez sez; do
for st in 1 2 3 4; do
  EX=""; [ "$v" = sez ] && EX="--actions 20"
  python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --variant $v $EX --streams $st > gpurun_out/st_$v$st.json 2> gpurun_out/st_$v$st.err
  python - $v $st gpurun_out/st_$v$st.json <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[3]) if l.startswith("{")][-1])
    print("%s streams %s: value %.1fM ms %.4f" % (sys.argv[1], sys.argv[2], d["value"] / 1e6, d["ms_per_step"]))
except Exception as e:
    print(sys.argv[1], sys.argv[2], "ERR", e)
PY
done
done
