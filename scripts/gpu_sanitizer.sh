#!/bin/bash
# compute-sanitizer over the fused parity cases (VERDICT r1 hygiene item): memcheck on the golden / fused-step / scalar-transform /
# caller kernels, synccheck + racecheck-free kernels have no shared memory so synccheck covers the warp-level barriers.
# Logs -> gpurun_out/sanitizer_*.log (summaries are copied to profiles/r02_sanitizer.md by hand).
mkdir -p gpurun_out
SAN=/usr/local/cuda/bin/compute-sanitizer
MEM_TESTS="tests/test_gpu_tree_parity.py::test_fused_step_matches_oracle tests/test_gpu_tree_parity.py::test_fused_step_deep_paths tests/test_gpu_golden_and_scale.py::test_cuda_path_matches_reference_golden tests/test_scaling.py tests/test_gpu_callers.py::test_value_targets_kernel_matches_the_reference_golden tests/test_gpu_callers.py::test_target_policies_kernel_matches_the_reference_golden tests/test_gpu_collect.py tests/test_gpu_sampler.py::test_cuda_sampler_reproduces_the_reference_tables_and_search tests/test_zz_no_alloc.py"
timeout 1500 $SAN --tool memcheck --error-exitcode 9 --print-limit 20 --log-file gpurun_out/sanitizer_memcheck.log \
  python -m pytest $MEM_TESTS -m gpu -x -q -p no:cacheprovider > gpurun_out/sanitizer_memcheck_pytest.log 2>&1; echo memcheck=$?
tail -n 3 gpurun_out/sanitizer_memcheck_pytest.log; tail -n 4 gpurun_out/sanitizer_memcheck.log
SYNC_TESTS="tests/test_gpu_tree_parity.py::test_fused_step_matches_oracle tests/test_gpu_tree_parity.py::test_gumbel_m16_a18 tests/test_gpu_tree_parity.py::test_efficientzero_pong_shape tests/test_scaling.py::test_cuda_kernel_two_heads_one_launch_and_mixed_normalisation"
timeout 900 $SAN --tool synccheck --error-exitcode 9 --print-limit 20 --log-file gpurun_out/sanitizer_synccheck.log \
  python -m pytest $SYNC_TESTS -m gpu -x -q -p no:cacheprovider > gpurun_out/sanitizer_synccheck_pytest.log 2>&1; echo synccheck=$?
tail -n 3 gpurun_out/sanitizer_synccheck_pytest.log; tail -n 4 gpurun_out/sanitizer_synccheck.log
