#!/bin/bash
# quick iteration: GPU parity suite + headline bench (graph) + tree-only bench; prints value / per-step / step-kernel us
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 3 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_q.json 2> gpurun_out/bench_q.err; echo bench=$?
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --latent 0 > gpurun_out/bench_q_l0.json 2>> gpurun_out/bench_q.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --variant ez > gpurun_out/bench_q_ez.json 2>> gpurun_out/bench_q.err
python - <<PY
import json
for f in ("bench_q", "bench_q_l0", "bench_q_ez"):
    try:
        d = json.load(open("gpurun_out/%s.json" % f))
        print(f, "value %.1fM" % (d["value"] / 1e6), "ms", d["ms_per_step"], "e2e %.1fM" % (d["e2e"]["value"] / 1e6), "step_us", d["roofline"]["avg_launch_us"], "frac", d["roofline"]["frac"])
    except Exception as e:
        print(f, "ERR", e)
PY
tail -n 3 gpurun_out/bench_q.err
