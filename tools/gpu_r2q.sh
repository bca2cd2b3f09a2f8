set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2q_pytest.log; tail -3 gpurun_out/r2q_pytest.log
python tools/lm_shard_probe.py 8; python tools/lm_shard_probe.py 4
for L in 256 1024 4096; do python bench_large_map.py --landmarks $L --updates 120 --warmup 10 --fused 3 | python -c "import sys,json;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('$L lm', d['value'], d['fused_step'])"; done
