# A/B of two builds of the library (ab/libekf_old.so vs ab/libekf_new.so) on workloads B and M, alternating
set -x
mkdir -p gpurun_out
L=slam-duckietown_b200/csrc/libekf_b200.so
cp ab/libekf_new.so $L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ab_pytest.log
for i in 1 2; do
  for v in old new; do
    cp ab/libekf_$v.so $L
    timeout 300 python bench.py --workload M --no-cpu-baseline --no-large-map > gpurun_out/ab_M_${v}_$i.json 2> gpurun_out/ab_M_${v}_$i.err
    timeout 300 python bench.py --no-cpu-baseline --no-large-map > gpurun_out/ab_B_${v}_$i.json 2> gpurun_out/ab_B_${v}_$i.err
  done
done
cp ab/libekf_new.so $L
tail -n 3 gpurun_out/ab_pytest.log
grep -o '"value": [0-9.]*' gpurun_out/ab_[MB]_*.json
