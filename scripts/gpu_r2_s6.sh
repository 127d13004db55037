#!/bin/bash
# Round 2, GPU session 6: full GPU suite after the sampler / callers / eval-path additions; default bench.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 12 gpurun_out/pytest.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_s6.json 2> gpurun_out/bench_s6.err; echo bench=$?; tail -n 3 gpurun_out/bench_s6.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_s6.json"))
print("value %.1fM ms %.4f e2e %.1fM lists %s step_us %s parity %s" % (d["value"] / 1e6, d["ms_per_step"], d["e2e"]["value"] / 1e6,
      d.get("e2e_lists"), d["roofline"]["avg_launch_us"], d.get("parity")))
PY
