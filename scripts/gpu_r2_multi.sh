#!/bin/bash
# Round 2, multi-GPU session (gpurun --gpus 8): strong scaling (65,536 roots total over N = 1, 2, 4, 8), weak scaling at N = 2 and 8
# with the gathered exchange check, NCCL exchange for comparison.  One JSON line per run -> gpurun_out/multi_*.json
mkdir -p gpurun_out
run() { # name N extra...
  name=$1; N=$2; shift 2
  if [ "$N" = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/multi_$name.json 2> gpurun_out/multi_$name.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/multi_$name.json 2> gpurun_out/multi_$name.err
  fi
  echo "$name rc=$?"
  python - gpurun_out/multi_$name.json <<'PY'
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
nvidia-smi topo -m > gpurun_out/multi_topo.txt 2>&1
run strong1 1 --scaling strong
run strong2 2 --scaling strong
run strong4 4 --scaling strong
run strong8 8 --scaling strong
run weak8 8
run weak2 2
run weak8_nccl 8 --nccl-exchange --no-e2e
tail -n 5 gpurun_out/multi_*.err | tail -n 30
