#!/bin/bash
# Round 2, GPU session 4: the vectorised scalar-transform kernel - tests, categorical bench, launch list with instruction counts.
mkdir -p gpurun_out
python -m pytest tests/test_scaling.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_scaling.log 2>&1; echo pytest_scaling=$?; tail -n 5 gpurun_out/pytest_scaling.log
python bench.py --steps 10 --warmup 3 --categorical --no-cpu-baseline > gpurun_out/bench_cat.json 2> gpurun_out/bench_cat.err; echo bench_cat=$?; tail -n 3 gpurun_out/bench_cat.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/bench_cat.json"))
print("value %.1fM ms %.4f e2e %.1fM step_us %s transform %s" % (d["value"] / 1e6, d["ms_per_step"], d["e2e"]["value"] / 1e6, d["roofline"]["avg_launch_us"], d["roofline"].get("scalar_transform")))
PY
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e --categorical"
$CMD > gpurun_out/plain_cat.log 2>&1 && \
ncu --metrics gpu__time_duration.sum,smsp__inst_executed.sum,dram__bytes_read.sum,sm__warps_active.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k regex:k_inverse_scalar -s 60 -c 3 --csv --log-file gpurun_out/scalar_launches.csv $CMD > gpurun_out/ncu_scalar2.log 2>&1; echo ncu=$?
tail -n 20 gpurun_out/scalar_launches.csv | cut -c1-260
