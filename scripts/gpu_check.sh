#!/bin/bash
# GPU-box check: parity tests, smoke, headline bench (+ variants), reference arm.  Outputs land in gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?
tail -n 15 gpurun_out/pytest.log
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo smoke=$?; tail -n 3 gpurun_out/smoke.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench=$?
python bench.py --steps 10 --warmup 3 --latent 0 --no-cpu-baseline > gpurun_out/bench_l0.json 2> gpurun_out/bench_l0.err; echo bench_l0=$?
python bench.py --steps 5 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo bench_ng=$?
if [ "$1" != "noref" ]; then python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref=$?; fi
cat gpurun_out/bench*.json; tail -n 5 gpurun_out/*.err
