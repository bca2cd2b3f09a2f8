# round 2: single-launch large-map update (gather CTAs + streaming tiles in one kernel): tests + A/B on 1 GPU
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r2h_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2h_pytest.log
for s in 0 1 0 1; do
  EKF_LM_SINGLE=$s timeout 300 python bench_large_map.py --landmarks 4096 --updates 120 --warmup 10 --fused 0 --check > gpurun_out/r2h_lm_single$s.json 2>> gpurun_out/r2h_lm.err
  python -c "
import json;d=json.loads([l for l in open('gpurun_out/r2h_lm_single$s.json') if l.startswith('{')][-1]);print('single=$s', d['value'], d['roofline']['frac'], d['sharded_equals_unsharded_bitwise'])"
done
EKF_LM_SINGLE=0 timeout 300 python bench_large_map.py --landmarks 1024 --updates 120 --warmup 10 --fused 0 | python -c "import sys,json;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('1024 two-launch', d['value'])"
EKF_LM_SINGLE=1 timeout 300 python bench_large_map.py --landmarks 1024 --updates 120 --warmup 10 --fused 0 | python -c "import sys,json;d=json.loads(sys.stdin.read().strip().splitlines()[-1]);print('1024 single-launch', d['value'])"
tail -5 gpurun_out/r2h_pytest.log; tail -3 gpurun_out/r2h_lm.err
