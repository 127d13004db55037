#!/bin/bash
# Round 2, final GPU session: full suite (shipped build), checked-build suite, smoke, the BASELINE sweep, reference arm
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 4 gpurun_out/pytest.log
bash scripts/gpu_checked.sh
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo smoke=$?; tail -n 2 gpurun_out/smoke.log
bash scripts/gpu_r2_sweep.sh 2>&1 | tail -n 26
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref=$?; tail -n 1 gpurun_out/bench_ref.json | cut -c1-300
