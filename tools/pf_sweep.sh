# sweep of the group kernel's L2-prefetch distance (EKF_GROUP_PREFETCH, in CTAs) on config M
mkdir -p gpurun_out
for d in ${PF_LIST:-0 1036}; do
  EKF_BATCH_PREFETCH=$d timeout 300 python bench.py ${PF_WORKLOAD:---workload M} --no-cpu-baseline --no-large-map > gpurun_out/pf_$d.json 2> gpurun_out/pf_$d.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/pf_$d.json').read().strip().splitlines()[-1])
print('pf $d', round(d['value']/1e6,1), round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['finite_and_faultless'])
PY
done | tee gpurun_out/pf_sweep.txt
