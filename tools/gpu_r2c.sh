# round 2, third GPU check (2 GPUs): tests (p2p over two real devices), N=1 bench, N=2 bench (sharded objects)
set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2c_gpus.txt 2>&1
python -c "
import sys; sys.path.insert(0,'.')
from oracle import ref_loader as r
print('reference kind on the box:', r.reference_kind())" >> gpurun_out/r2c_gpus.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r2c_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2c_pytest.log
timeout 300 python tools/mc_probe.py > gpurun_out/r2c_mc_probe.json 2> gpurun_out/r2c_mc_probe.err
timeout 900 python bench.py > gpurun_out/r2c_bench1.json 2> gpurun_out/r2c_bench1.err; echo "bench1 rc=$?" >> gpurun_out/r2c_bench1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2c_bench2.json 2> gpurun_out/r2c_bench2.err; echo "bench2 rc=$?" >> gpurun_out/r2c_bench2.err
cat gpurun_out/r2c_gpus.txt; tail -12 gpurun_out/r2c_pytest.log; cat gpurun_out/r2c_mc_probe.json; tail -3 gpurun_out/r2c_bench1.err; tail -5 gpurun_out/r2c_bench2.err
python - <<'PY'
import json
def show(k,v,ind=0):
    if isinstance(v,dict):
        print(' '*ind+k+':')
        for a,b in v.items(): show(a,b,ind+2)
    else:
        print(' '*ind+f'{k}: {str(v)[:100]}')
try:
    d=json.load(open('gpurun_out/r2c_bench1.json'))
    for k in ('value','large_map','bag_ensemble','config1_replay','cpu_baseline'):
        show(k,d.get(k))
    print('M e2e', d['small_filters']['e2e']['value'], d['small_filters']['e2e']['frac_of_device_resident'])
except Exception as e: print('bench1 parse', e)
try:
    d=json.load(open('gpurun_out/r2c_bench2.json'))
    for k in ('value','e2e','small_filters','large_map_sharded'):
        show(k,d.get(k))
except Exception as e: print('bench2 parse', e)
PY
