#!/bin/bash
# Round 2, GPU session 10: full GPU suite on the default build; the sez graph test alone on default / pbc2 / cval builds.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo pytest=$?; tail -n 5 gpurun_out/pytest.log
for v in pbc2 cval; do
  LZB200_LIB=lightzero_b200/liblzb200_$v.so python -m pytest "tests/test_gpu_golden_and_scale.py::test_headline_size_properties_of_the_other_trees" -m gpu -x -q > gpurun_out/pytest_one_$v.log 2>&1; echo one_$v=$?; tail -n 3 gpurun_out/pytest_one_$v.log
  LZB200_LIB=lightzero_b200/liblzb200_$v.so python -m pytest tests/test_gpu_headline_parity.py tests/test_gpu_tree_parity.py tests/test_gpu_golden_and_scale.py tests/test_gpu_api.py -m gpu -x -q > gpurun_out/pytest_$v.log 2>&1; echo pytest_$v=$?; tail -n 3 gpurun_out/pytest_$v.log
done
