#!/bin/bash
# Round 2, GPU session 5: full GPU suite (callers, scaling, no-alloc ...), categorical + default bench (e2e_lists), reanalyze row.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 8 gpurun_out/pytest.log
python scripts/reanalyze_bench.py > gpurun_out/reanalyze.json 2> gpurun_out/reanalyze.err; echo reanalyze=$?; tail -n 3 gpurun_out/reanalyze.err; cat gpurun_out/reanalyze.json
python bench.py --steps 10 --warmup 3 --categorical --no-cpu-baseline > gpurun_out/bench_cat.json 2> gpurun_out/bench_cat.err; echo bench_cat=$?; tail -n 3 gpurun_out/bench_cat.err
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_s5.json 2> gpurun_out/bench_s5.err; echo bench=$?; tail -n 3 gpurun_out/bench_s5.err
python - <<'PY'
import json
for f in ("bench_cat", "bench_s5"):
    d = json.load(open("gpurun_out/%s.json" % f))
    print(f, "value %.1fM ms %.4f e2e %.1fM lists %s step_us %s transform %s parity %s" % (d["value"] / 1e6, d["ms_per_step"], d["e2e"]["value"] / 1e6,
          d.get("e2e_lists", {}).get("ms_per_step"), d["roofline"]["avg_launch_us"], (d["roofline"].get("scalar_transform") or {}).get("avg_launch_us"), d.get("parity")))
PY
