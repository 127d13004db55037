#!/bin/bash
# Round 2 BASELINE.md table (1 GPU): every variant at the headline shape (incl. categorical heads, Sampled at K=20 and Stochastic),
# the C5 size sweep, the two reference-config shapes.  One JSON line per run -> gpurun_out/r02_sweep.jsonl (copied to profiles/).
mkdir -p gpurun_out; : > gpurun_out/r02_sweep.jsonl; : > gpurun_out/r02_sweep.err
run() { python bench.py --steps 5 --warmup 3 "$@" 2>>gpurun_out/r02_sweep.err | tail -n 1 >> gpurun_out/r02_sweep.jsonl; }
run --latent 1024
run --latent 1024 --categorical --no-cpu-baseline
run --latent 256 --no-cpu-baseline
run --latent 0 --no-cpu-baseline
run --variant ez --latent 1024 --no-cpu-baseline
run --variant ez --latent 1024 --categorical --no-cpu-baseline
run --two-player --latent 1024 --no-cpu-baseline
run --variant gmz --latent 1024 --no-cpu-baseline
run --variant sez --actions 20 --latent 1024 --no-cpu-baseline
run --variant smz --latent 1024 --no-cpu-baseline
for N in 1024 4096 16384 65536; do run --roots $N --latent 256 --no-cpu-baseline; done
for S in 100 200; do run --sims $S --latent 256 --no-cpu-baseline; done
run --roots 65536 --sims 200 --latent 0 --no-cpu-baseline
run --roots 8 --sims 25 --actions 2 --latent 128 --cpu-sample-roots 8
run --roots 8 --sims 50 --actions 6 --latent 1024 --variant ez --cpu-sample-roots 8
python - <<'PY'
import json
for l in open('gpurun_out/r02_sweep.jsonl'):
    try: d=json.loads(l)
    except Exception: print('BAD', l[:200]); continue
    r=d['roofline']; c=d.get('cpu_baseline',{}); t=r.get('scalar_transform') or {}
    print('%-70s %s value %.4g ms %.3f e2e %.4g lists %s step_us %s frac %.3f depth %.2f cpu %s xf_us %s par %s' % (d['config']['workload'][10:80], d['config']['scalar_transform'][:5], d['value'], d['ms_per_step'],
          d.get('e2e',{}).get('value',0), (d.get('e2e_lists') or {}).get('ms_per_step'), r['avg_launch_us_all'].get('step'), r['frac'], r['mean_search_depth'], c.get('value'), t.get('avg_launch_us'), (d.get('parity') or {}).get('visit_mismatches')))
PY
tail -n 5 gpurun_out/r02_sweep.err
python scripts/scalar_bench.py | tail -n 1
python scripts/reanalyze_bench.py 2>&1 | tail -n 4
