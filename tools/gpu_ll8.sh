set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_sharding.py -m gpu -x -q -k "p2p or shard" > gpurun_out/ll8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ll8_pytest.log
run_tl() { # name L env...
  name=$1; L=$2; shift 2
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29563 tools/lm_p2p_timeline.py $L 60 > gpurun_out/tl_$name.txt 2> gpurun_out/tl_$name.err; echo "rc=$?" >> gpurun_out/tl_$name.err
}
run_tl 4096_auto 4096 EKF_LM_PREPUSH_SPREAD=1
run_tl 4096_equal 4096 EKF_LM_RANK0_FEWER=0
run_tl 4096_auto_nospread 4096 EKF_LM_PREPUSH_SPREAD=0
run_tl 4096_f150 4096 EKF_LM_RANK0_FEWER=150
run_tl xl_auto 32768 EKF_LM_PREPUSH_SPREAD=1
run_tl xl_equal 32768 EKF_LM_RANK0_FEWER=0
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus $N --steps 20 --warmup 3 > gpurun_out/ll8_bench.json 2> gpurun_out/ll8_bench.err; echo "bench rc=$?" >> gpurun_out/ll8_bench.err
tail -3 gpurun_out/ll8_pytest.log
for f in 4096_auto 4096_equal 4096_auto_nospread 4096_f150 xl_auto xl_equal; do echo "#### $f"; grep -A9 "== world" gpurun_out/tl_$f.txt; grep -A8 "^-- other" gpurun_out/tl_$f.txt | grep "period\|stream kernel\|arrived"; tail -1 gpurun_out/tl_$f.err; done
tail -3 gpurun_out/ll8_bench.err
python - <<'PY'
import json
try:
    txt=open('gpurun_out/ll8_bench.json').read()
    d=json.loads([l for l in txt.splitlines() if l.startswith('{')][-1])
    sf=d['small_filters']; lm=d['large_map_sharded']
    print(d['n_gpus'],'GPUs: B',d['value'],'e2e',d['e2e']['value'],'| M',sf['value'],'frac',sf['roofline']['frac'],'e2e',sf['e2e']['value'],sf['e2e']['frac_of_device_resident'])
    for k in ('l4096','xl'):
        v=lm[k]
        print('  ',k,{a:v.get(a) for a in ('ms_per_update','sharded_equals_unsharded_bitwise','p2p_error','finite','error')}, 'frac',v.get('roofline',{}).get('frac'),'fused',v.get('fused_step',{}).get('ms_per_step'),'parity',v.get('tracked_rows_parity',{}).get('worst'),'joseph',(v.get('joseph') or {}).get('ms_per_update'))
except Exception as e: print('parse',e)
PY
