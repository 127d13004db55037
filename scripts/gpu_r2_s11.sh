#!/bin/bash
# Round 2, GPU session 11: scalar-transform kernel A/B (barrier-free epilogue, block size, cheaper h^-1), its tests per build,
# and the order-dependent capture failure re-run with the cval experiment build (GC now off during capture).
mkdir -p gpurun_out
for v in "" sb256 s64 sbar128 sold; do
  if [ -z "$v" ]; then E=""; else E="LZB200_LIB=lightzero_b200/liblzb200_$v.so"; fi
  env $E python scripts/scalar_bench.py 2>&1 | tail -n 1
  env $E python -m pytest tests/test_scaling.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -n 1
done
LZB200_LIB=lightzero_b200/liblzb200_cval.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q > gpurun_out/pytest_cval.log 2>&1; echo pytest_cval=$?; tail -n 3 gpurun_out/pytest_cval.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --categorical > gpurun_out/bench_cat.json 2> gpurun_out/bench_cat.err; echo bench_cat=$?
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_cat.json"))
print("categorical: value %.1fM ms %.4f step_us %s transform %s" % (d["value"] / 1e6, d["ms_per_step"], d["roofline"]["avg_launch_us"], d["roofline"].get("scalar_transform")))
PY
