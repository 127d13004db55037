#!/bin/bash
# Round 1 GPU session: GPU suite, smoke, headline bench (graph / tree-only / eager), reference arm, host CPU description
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo smoke=$?
python bench.py --steps 5 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo bench=$?
python bench.py --steps 5 --warmup 3 --latent 0 --no-cpu-baseline > gpurun_out/bench_l0.json 2> gpurun_out/bench_l0.err; echo bench_l0=$?
python bench.py --steps 5 --warmup 3 --no-graph --no-cpu-baseline > gpurun_out/bench_nograph.json 2> gpurun_out/bench_nograph.err; echo bench_ng=$?
python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref=$?
nproc; lscpu | head -20
tail -5 gpurun_out/pytest.log gpurun_out/smoke.log; cat gpurun_out/bench*.json; tail -5 gpurun_out/*.err
python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 500 -c 320 --csv --log-file gpurun_out/launches.csv python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e > gpurun_out/ncu.log 2>&1; echo ncu=$?
