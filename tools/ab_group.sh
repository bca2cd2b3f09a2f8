# A/B of two builds of the library on config M (and B): ab/libekf_old.so vs ab/libekf_new.so
set -x
mkdir -p gpurun_out
L=slam-duckietown_b200/csrc/libekf_b200.so
cp ab/libekf_new.so $L
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/ab_pytest.log
for i in 1 2; do
  for v in old new; do
    cp ab/libekf_$v.so $L
    timeout 300 python bench.py --workload M --no-cpu-baseline --no-large-map > gpurun_out/ab_M_${v}_$i.json 2> gpurun_out/ab_M_${v}_$i.err
  done
done
cp ab/libekf_new.so $L
timeout 600 ncu --set full --clock-control none --import-source on -k regex:batch_step_group -s 3 -c 1 -o gpurun_out/ab_batchM_new python bench.py --workload M --steps 5 --warmup 3 --no-cpu-baseline --no-large-map > gpurun_out/ab_ncu_fullM.log 2>&1
tail -n 3 gpurun_out/ab_pytest.log
grep -o '"value": [0-9.]*' gpurun_out/ab_M_*.json
