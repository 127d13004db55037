#!/bin/bash
# A/B helper: headline bench (value only) under a list of env settings; prints value, step us
# usage: scripts/gpu_ab.sh "ENV1=a ENV2=b" "ENV1=c" ...   ("-" = no env)
mkdir -p gpurun_out
i=0
for envs in "$@"; do
  i=$((i+1))
  [ "$envs" = "-" ] && envs=""
  for extra in "" "--latent 0"; do
    env $envs python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e $extra $AB_ARGS > gpurun_out/ab_$i.json 2> gpurun_out/ab_$i.err || tail -n 3 gpurun_out/ab_$i.err
    python - "$envs $extra" gpurun_out/ab_$i.json <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[2]))
    print("%-40s value %.1fM ms %.4f step_us %s frac %s" % (sys.argv[1], d["value"] / 1e6, d["ms_per_step"], d["roofline"]["avg_launch_us"], d["roofline"]["frac"]))
except Exception as e:
    print(sys.argv[1], "ERR", e)
PY
  done
done
