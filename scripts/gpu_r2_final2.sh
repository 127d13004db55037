#!/bin/bash
# Round 2, last GPU session after the CREDUX reductions and the bulk-staged lone-walker rows became defaults: headline bench line,
# ncu --set full of the step kernel + warm launch lists, in-kernel timeline of the new kernel, smoke, the BASELINE sweep, reference arm
mkdir -p gpurun_out
bash scripts/gpu_profile_round.sh > gpurun_out/profile_round.log 2>&1; tail -n 6 gpurun_out/profile_round.log | cut -c1-400
timeout 600 python scripts/gpu_timeline.py --no-save > gpurun_out/timeline_L1024.txt 2>&1; echo timeline=$?
python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo smoke=$?; tail -n 2 gpurun_out/smoke.log
bash scripts/gpu_r2_sweep.sh 2>&1 | tail -n 26
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo ref=$?; tail -n 1 gpurun_out/bench_ref.json | cut -c1-300
