set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/t1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/t1_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/t1_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/t1_smoke.log
timeout 400 python bench.py > gpurun_out/t1_benchB.json 2> gpurun_out/t1_benchB.err
timeout 300 python bench.py --workload M --no-cpu-baseline --no-large-map > gpurun_out/t1_benchM.json 2> gpurun_out/t1_benchM.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/t1_ref.json 2> gpurun_out/t1_ref.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/t1_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/t1_ncu_launch.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_step_regw -s 3 -c 1 -o gpurun_out/t1_batch python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-large-map > gpurun_out/t1_ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_step_group -s 3 -c 1 -o gpurun_out/t1_batchM python bench.py --workload M --steps 5 --warmup 3 --no-cpu-baseline --no-large-map > gpurun_out/t1_ncu_fullM.log 2>&1
tail -3 gpurun_out/t1_pytest.log; cat gpurun_out/t1_benchB.json gpurun_out/t1_benchM.json
