# A/B of the warp-per-filter kernel's opt-in L2 prefetch on config B. EKF_REGW_PREFETCH: distance in CTAs (< 0: inputs +
# mu only, no P); EKF_REGW_PREFETCH_TAIL=1: issued after the filter's last store instead of at the head of its chain.
# (EKF_REGW_PREFETCH_TAIL needs the tail-placement build of commit 8a705c7; the shipped kernel ignores it.)
# AB_LIST: space-separated "distance:tail" pairs
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r_pytest.log
tail -n 3 gpurun_out/r_pytest.log
i=0
for dt in ${AB_LIST:-0:0 888:1 444:1 1332:1 -888:1 0:0 888:1}; do
  i=$((i+1)); d=${dt%%:*}; t=${dt##*:}
  EKF_REGW_PREFETCH=$d EKF_REGW_PREFETCH_TAIL=$t timeout 300 python bench.py --no-cpu-baseline --no-large-map > gpurun_out/rpf_$i.json 2> gpurun_out/rpf_$i.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/rpf_$i.json').read().strip().splitlines()[-1])
print('EKF_REGW_PREFETCH=$d TAIL=$t', round(d['value']/1e6,1), round(d['ms_per_step'],4), round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'], d['finite_and_faultless'])
PY
done | tee gpurun_out/rpf_ab.txt
