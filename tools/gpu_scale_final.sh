# final multi-GPU record of the tree: bench.py at N = 8, 4, 2 on one box (as the driver's scaling run does)
set -x
mkdir -p gpurun_out
for N in 8 4 2; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2957$N bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/fs_bench$N.json 2> gpurun_out/fs_bench$N.err; echo "bench$N rc=$?" >> gpurun_out/fs_bench$N.err
tail -2 gpurun_out/fs_bench$N.err
done
python - <<'PY'
import json
for n in (8,4,2):
    try:
        txt=open(f'gpurun_out/fs_bench{n}.json').read()
        d=json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
        sf=d['small_filters']; lm=d['large_map_sharded']
        print(n,'GPUs: B',d['value'],'e2e',d['e2e']['value'],'| M',sf['value'],'frac',sf['roofline']['frac'],'e2e',sf['e2e']['value'],sf['e2e']['frac_of_device_resident'])
        for k in ('l4096','xl'):
            v=lm[k]
            print('  ',k,{a:v.get(a) for a in ('ms_per_update','sharded_equals_unsharded_bitwise','p2p_error','finite','error')}, 'frac',v.get('roofline',{}).get('frac'),'fused',v.get('fused_step',{}).get('ms_per_step'),'parity',v.get('tracked_rows_parity',{}).get('worst'),'joseph',(v.get('joseph') or {}).get('ms_per_update'))
    except Exception as e: print(n,'parse',e)
PY
