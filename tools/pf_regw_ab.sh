# A/B of the warp-per-filter kernel's L2 prefetch restricted to the per-filter inputs + mu (EKF_REGW_PREFETCH < 0), config B
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r_pytest.log
tail -n 3 gpurun_out/r_pytest.log
i=0
for d in 0 -444 -888 -1776 0 -444; do
  i=$((i+1))
  EKF_REGW_PREFETCH=$d timeout 300 python bench.py --no-cpu-baseline --no-large-map > gpurun_out/rpf_$i.json 2> gpurun_out/rpf_$i.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/rpf_$i.json').read().strip().splitlines()[-1])
print('EKF_REGW_PREFETCH=$d', round(d['value']/1e6,1), round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'], d['finite_and_faultless'])
PY
done | tee gpurun_out/rpf_ab.txt
