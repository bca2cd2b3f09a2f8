# round 2: compute-sanitizer (memcheck, then racecheck / synccheck on the shared-memory kernels) over small GPU tests
set -x
mkdir -p gpurun_out
export PYTORCH_NO_CUDA_MEMORY_CACHING=1
SEL='test_large_map_50_steps_vs_oracle_and_pad or test_large_map_fused_multi_update_bitwise or test_batched_flag_variants_vs_oracle or test_in_kernel_generation_equals_perturb_then_step_bitwise or test_sharded_equals_unsharded_bitwise or test_batched_joseph_variants_bitwise_equal or test_association_overflow_keeps_slot_order or test_large_map_fused_joseph_bitwise'
timeout 1500 compute-sanitizer --tool memcheck --error-exitcode 9 --launch-timeout 0 python -m pytest tests -m gpu -x -q -k "$SEL" > gpurun_out/r2s_memcheck.log 2>&1; echo "memcheck rc=$?" >> gpurun_out/r2s_memcheck.log
SEL2='test_batched_flag_variants_vs_oracle or test_in_kernel_generation_equals_perturb_then_step_bitwise or test_large_map_fused_joseph_bitwise'
timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "$SEL2" > gpurun_out/r2s_racecheck.log 2>&1; echo "racecheck rc=$?" >> gpurun_out/r2s_racecheck.log
timeout 900 compute-sanitizer --tool synccheck --error-exitcode 9 python -m pytest tests -m gpu -x -q -k "$SEL2" > gpurun_out/r2s_synccheck.log 2>&1; echo "synccheck rc=$?" >> gpurun_out/r2s_synccheck.log
for f in memcheck racecheck synccheck; do echo == $f; grep -E "ERROR SUMMARY|passed|failed|rc=|RACECHECK SUMMARY|Hazard|Invalid" gpurun_out/r2s_$f.log | sort | uniq -c | head -12; done
