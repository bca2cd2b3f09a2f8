# round 1q: bench M + ncu captures of the final batched kernels (group kernel with the L2 prefetch)
set -x
mkdir -p gpurun_out
timeout 300 python bench.py --workload M --no-cpu-baseline --no-large-map > gpurun_out/q_benchM.json 2> gpurun_out/q_benchM.err
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/q_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/q_ncu_launch.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:batch_step_group -s 3 -c 1 -o gpurun_out/q_batchM python bench.py --workload M --steps 5 --warmup 3 --no-cpu-baseline --no-large-map > gpurun_out/q_ncu_fullM.log 2>&1
timeout 300 ncu --set full --clock-control none --import-source on -k regex:batch_step_regw -s 3 -c 1 -o gpurun_out/q_batch python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-large-map > gpurun_out/q_ncu_full.log 2>&1
cat gpurun_out/q_benchM.json; ls -la gpurun_out
