# round 2, 8-GPU check after the automatic 16-row tile selection: p2p tests on distinct devices + N=8 bench
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "p2p" > gpurun_out/r2s_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2s_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2s_bench8.json 2> gpurun_out/r2s_bench8.err; echo "bench8 rc=$?" >> gpurun_out/r2s_bench8.err
tail -4 gpurun_out/r2s_pytest.log; tail -3 gpurun_out/r2s_bench8.err
python - <<'PY'
import json
for n in (8,):
    try:
        txt=open(f'gpurun_out/r2s_bench{n}.json').read()
        d=json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
        sf=d['small_filters']; lm=d['large_map_sharded']
        print(n,'GPUs: B',d['value'],'e2e',d['e2e']['value'],'| M',sf['value'],'frac',sf['roofline']['frac'],'e2e',sf['e2e']['value'],sf['e2e']['frac_of_device_resident'])
        for k in ('l4096','xl'):
            v=lm[k]
            print('  ',k,{a:v.get(a) for a in ('ms_per_update','sharded_equals_unsharded_bitwise','p2p_error','finite','error')}, 'frac',v.get('roofline',{}).get('frac'),'fused',v.get('fused_step',{}).get('ms_per_step'),'parity',v.get('tracked_rows_parity',{}).get('worst'))
    except Exception as e: print(n,'parse',e)
PY
