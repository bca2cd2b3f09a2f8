set -x
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29578 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/fz_bench8.json 2> gpurun_out/fz_bench8.err; echo "bench8 rc=$?" >> gpurun_out/fz_bench8.err
tail -2 gpurun_out/fz_bench8.err
python - <<'PY'
import json
txt=open('gpurun_out/fz_bench8.json').read()
d=json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
sf=d['small_filters']; lm=d['large_map_sharded']
print('8 GPUs: B',d['value'],'e2e',d['e2e']['value'],'gen',d['e2e'].get('generated_inputs',{}).get('value'),'| M',sf['value'],'frac',sf['roofline']['frac'],'e2e',sf['e2e']['value'],sf['e2e']['frac_of_device_resident'])
for k in ('l4096','xl'):
    v=lm[k]
    print('  ',k,{a:v.get(a) for a in ('ms_per_update','sharded_equals_unsharded_bitwise','p2p_error','finite','error')}, 'fused',v.get('fused_step',{}).get('ms_per_step'),'parity',v.get('tracked_rows_parity',{}).get('worst'),'joseph',(v.get('joseph') or {}).get('ms_per_update'))
PY
