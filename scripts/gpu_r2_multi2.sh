#!/bin/bash
# Round 2, 2-GPU session on the final kernel (gpurun --gpus 2): weak and strong scaling at N = 2 with the gathered exchange check,
# the 1-GPU line of the same box for the efficiency.  One JSON line per run -> gpurun_out/multi2_*.json
mkdir -p gpurun_out
run() { # name N extra...
  name=$1; N=$2; shift 2
  if [ "$N" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/multi2_$name.json 2> gpurun_out/multi2_$name.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/multi2_$name.json 2> gpurun_out/multi2_$name.err
  fi
  echo "$name rc=$?"
  python - gpurun_out/multi2_$name.json <<'PY'
import json, sys
try:
    d = json.loads([l for l in open(sys.argv[1]) if l.startswith("{")][-1])
    e = d.get("e2e", {})
    print("  n=%d %s value %.1fM ms %.4f e2e %.1fM exch_ms %s check %s" % (d["n_gpus"], d["scaling"], d["value"] / 1e6, d["ms_per_step"],
          e.get("value", 0) / 1e6, d["run"].get("exchange_ms_per_step"), d.get("exchange_check")))
except Exception as ex:
    print("  ERR", ex)
PY
}
run weak1 1
run weak2 2
run strong2 2 --scaling strong
tail -n 5 gpurun_out/multi2_*.err | tail -n 20
