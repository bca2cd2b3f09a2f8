# round 2, second GPU check (1 GPU): diagnose oracle/_ref on the box, GPU tests, bench with the new objects
set -x
mkdir -p gpurun_out
python -V > gpurun_out/r2b_diag.txt 2>&1
ls -la oracle/_ref >> gpurun_out/r2b_diag.txt 2>&1
python -c "
import sys; sys.path.insert(0,'.')
from oracle import ref_loader as r
import importlib.util
print('kind', r.reference_kind(), 'magic', importlib.util.MAGIC_NUMBER.hex())
print(open('oracle/_ref/BUILD_INFO.txt').read())
" >> gpurun_out/r2b_diag.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/r2b_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2b_pytest.log
timeout 900 python bench.py > gpurun_out/r2b_benchB.json 2> gpurun_out/r2b_benchB.err; echo "bench rc=$?" >> gpurun_out/r2b_benchB.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2b_ref.json 2> gpurun_out/r2b_ref.err
cat gpurun_out/r2b_diag.txt; tail -15 gpurun_out/r2b_pytest.log; tail -5 gpurun_out/r2b_benchB.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r2b_benchB.json'))
def show(k,v,ind=0):
    if isinstance(v,dict):
        print(' '*ind+k+':')
        for a,b in v.items(): show(a,b,ind+2)
    else:
        t=str(v)
        print(' '*ind+f'{k}: {t[:110]}')
for k in ('value','e2e','sustained','cpu_baseline','vs_cpu_baseline','large_map','small_filters','bag_ensemble','config1_replay'):
    show(k,d.get(k))
PY
