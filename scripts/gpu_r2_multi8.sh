#!/bin/bash
# Round 2, 8-GPU session on the final kernel (gpurun --gpus 8): weak and strong scaling at N = 8 only -> gpurun_out/multi8_*.json
mkdir -p gpurun_out
for name in weak8 strong8; do
  extra=""; [ $name = strong8 ] && extra="--scaling strong"
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline $extra > gpurun_out/multi8_$name.json 2> gpurun_out/multi8_$name.err
  echo "$name rc=$?"
  python - gpurun_out/multi8_$name.json <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    e = d.get("e2e", {})
    print("  n=%d %s value %.1fM ms %.4f e2e %.1fM exch_ms %s check %s" % (d["n_gpus"], d["scaling"], d["value"] / 1e6, d["ms_per_step"],
          e.get("value", 0) / 1e6, d["run"].get("exchange_ms_per_step"), d.get("exchange_check")))
except Exception as ex:
    print("  ERR", ex)
PY
done
