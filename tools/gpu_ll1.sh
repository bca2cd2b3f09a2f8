set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/ll1_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/ll1_suite.log
tail -8 gpurun_out/ll1_suite.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/ll1_bench.json 2> gpurun_out/ll1_bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/ll1_bench.json').read().splitlines() if l.startswith('{')][-1])
print('B',d['value'],'e2e',d['e2e']['value'],'frac',d['roofline']['frac'])
lm=d['large_map']; print(json.dumps({k:lm[k] for k in lm if k!='roofline'})[:1500])
PY
