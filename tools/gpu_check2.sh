# 2-GPU confidence check: p2p tests on distinct devices + bench at N = 2
set -x
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "p2p" > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c2_pytest.log
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29572 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/c2_bench2.json 2> gpurun_out/c2_bench2.err; echo "bench2 rc=$?" >> gpurun_out/c2_bench2.err
tail -3 gpurun_out/c2_pytest.log; tail -2 gpurun_out/c2_bench2.err
python - <<'PY'
import json
txt=open('gpurun_out/c2_bench2.json').read()
d=json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
sf=d['small_filters']; lm=d['large_map_sharded']
print('2 GPUs: B',d['value'],'e2e',d['e2e']['value'],'| M',sf['value'],'e2e',sf['e2e']['value'],sf['e2e']['frac_of_device_resident'])
for k in ('l4096','xl'):
    v=lm[k]
    print('  ',k,{a:v.get(a) for a in ('ms_per_update','sharded_equals_unsharded_bitwise','p2p_error','finite','error')}, 'fused',v.get('fused_step',{}).get('ms_per_step'),'parity',v.get('tracked_rows_parity',{}).get('worst'),'joseph',(v.get('joseph') or {}).get('ms_per_update'))
PY
