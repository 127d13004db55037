#!/bin/bash
# Round 2, GPU session 1: new parity tests, then A/B of the lane-group layouts / table placement / L2 window,
# the phase timeline of the new default and one ncu --set full capture (with source) of the new step kernel.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 6 gpurun_out/pytest.log
export AB_ARGS=""
scripts/gpu_ab.sh "-" "LZB_GROUP_LANES=16" "LZB_L2_PERSIST_MB=64" "LZB_L2_PERSIST_MB=100" "LZB200_LIB=lightzero_b200/liblzb200_tc.so" "LZB200_LIB=lightzero_b200/liblzb200_tc.so LZB_GROUP_LANES=16" 2>&1 | grep -v "LZB200_LIB is set"
AB_ARGS="--roots 65536" scripts/gpu_ab.sh "-" "LZB_GROUP_LANES=16" "LZB_GROUP_LANES=8" 2>&1 | grep -v "LZB200_LIB is set"
for v in ez gmz sez smz; do AB_ARGS="--variant $v" scripts/gpu_ab.sh "-" "LZB_GROUP_LANES=16" 2>&1 | grep -v "LZB200_LIB is set"; done
python scripts/gpu_timeline.py --no-save > gpurun_out/timeline_L1024.txt 2>&1; echo tl=$?
python scripts/gpu_timeline.py --no-save --latent 0 > gpurun_out/timeline_L0.txt 2>&1
python scripts/gpu_timeline.py --no-save --latent 0 --roots 96 > gpurun_out/timeline_L0_alone.txt 2>&1
CMD="python bench.py --steps 1 --warmup 1 --no-graph --no-cpu-baseline --no-e2e"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_step -s 130 -c 2 -o gpurun_out/prof_step_r2a -f $CMD > gpurun_out/ncu_full.log 2>&1; echo ncu_full=$?
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_s1.json 2> gpurun_out/bench_s1.err; echo bench=$?; cat gpurun_out/bench_s1.json
