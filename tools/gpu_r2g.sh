# round 2, 1-GPU check after in-kernel generation of the Monte Carlo inputs
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/r2g_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2g_pytest.log
timeout 300 python tools/mc_probe.py > gpurun_out/r2g_mc_probe.json 2> gpurun_out/r2g_mc_probe.err
timeout 900 python bench.py > gpurun_out/r2g_bench1.json 2> gpurun_out/r2g_bench1.err; echo "bench1 rc=$?" >> gpurun_out/r2g_bench1.err
tail -6 gpurun_out/r2g_pytest.log; cat gpurun_out/r2g_mc_probe.json; tail -3 gpurun_out/r2g_bench1.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2g_bench1.json'))
print('B', d['value'], d['roofline']['frac'], 'e2e', d['e2e']['value'], 'gen', d['e2e']['generated_inputs']['value'])
print('sustained', d['sustained']['value'], d['sustained']['roofline']['frac_of_8TBs'], d['sustained']['clocks'])
j=d['joseph']; print('joseph', j['value'], j['roofline']['frac'], j['vs_plain'])
m=d['small_filters']; print('M', m['value'], m['roofline']['frac'], 'e2e', m['e2e']['value'], m['e2e']['frac_of_device_resident'])
b=d['bag_ensemble']; print('bag', b['value'], b['step_kernel'])
print('lm', d['large_map']['ms_per_update'], d['large_map']['predict_us'], d['large_map']['fused_step']['ms_per_step'])
print('cfg1', d['config1_replay'])
print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['kind'], d['cpu_baseline'].get('port_value'))
PY
