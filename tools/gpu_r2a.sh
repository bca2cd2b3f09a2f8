# round 2, first GPU check (1 GPU): GPU tests, smoke, the default bench, the reference arm
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/r2a_gpu.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2a_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2a_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r2a_smoke.log
timeout 600 python bench.py > gpurun_out/r2a_benchB.json 2> gpurun_out/r2a_benchB.err; echo "bench rc=$?" >> gpurun_out/r2a_benchB.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2a_ref.json 2> gpurun_out/r2a_ref.err
tail -15 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_smoke.log; tail -3 gpurun_out/r2a_benchB.err; cat gpurun_out/r2a_benchB.json | head -c 6000
