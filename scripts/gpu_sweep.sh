#!/bin/bash
# headline shape variants and the C5 sweep (1 GPU).  One JSON line per run -> gpurun_out/sweep.jsonl
mkdir -p gpurun_out; : > gpurun_out/sweep.jsonl
run() { python bench.py --steps 5 --warmup 3 "$@" 2>>gpurun_out/sweep.err | tail -n 1 >> gpurun_out/sweep.jsonl; }
run --latent 1024
run --latent 256 --no-cpu-baseline
run --latent 0 --no-cpu-baseline
run --variant ez --latent 1024
run --two-player --latent 1024
run --variant gmz --latent 1024
for N in 1024 4096 16384 65536; do run --roots $N --latent 256 --no-cpu-baseline; done
for S in 100 200; do run --sims $S --latent 256 --no-cpu-baseline; done
run --roots 65536 --sims 200 --latent 0 --no-cpu-baseline
run --roots 8 --sims 25 --actions 2 --latent 128 --cpu-sample-roots 8
run --roots 8 --sims 50 --actions 6 --latent 1024 --variant ez --cpu-sample-roots 8
python - <<'PY'
import json
for l in open('gpurun_out/sweep.jsonl'):
    try: d=json.loads(l)
    except Exception: print('BAD', l[:200]); continue
    r=d['roofline']; c=d.get('cpu_baseline',{})
    print('%-75s value %.4g  ms %.3f  e2e %.4g  step_us %s frac %.3f depth %.2f cpu1 %s' % (d['config']['workload'], d['value'], d['ms_per_step'], d.get('e2e',{}).get('value',0), r['avg_launch_us_all'].get('step'), r['frac'], r['mean_search_depth'], c.get('value')))
PY
tail -n 5 gpurun_out/sweep.err
