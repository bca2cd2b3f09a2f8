# final check of the tree: GPU tests, smoke, both bench arms
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/f_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/f_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/f_smoke.log
timeout 400 python bench.py --impl reference > gpurun_out/f_ref.json 2> gpurun_out/f_ref.err
timeout 400 python bench.py > gpurun_out/f_benchB.json 2> gpurun_out/f_benchB.err
tail -n 5 gpurun_out/f_pytest.log; tail -n 2 gpurun_out/f_smoke.log
grep -o '"value": [0-9.]*' gpurun_out/f_*.json
