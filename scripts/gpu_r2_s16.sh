#!/bin/bash
# Round 2, GPU session 16: new helper tests on the shipped build, then the checked-build run
mkdir -p gpurun_out
python -m pytest tests/test_gumbel_helpers.py tests/test_gpu_checked_build.py -m gpu -x -q -rs 2>&1 | tail -n 6
bash scripts/gpu_checked.sh
