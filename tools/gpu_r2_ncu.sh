# round 2: ncu evidence of the final build (1 GPU). Every profiled command first runs once WITHOUT ncu.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err; echo "rc=$?" >> gpurun_out/r2n_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2n_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r2n_ncu_launch.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batch_step_regw_kernel -s 12 -c 1 -o gpurun_out/r2n_batch python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-large-map --no-sustained > gpurun_out/r2n_ncu_full.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batch_step_group_kernel -s 8 -c 1 -o gpurun_out/r2n_batchM python bench.py --workload M --steps 5 --warmup 3 --no-cpu-baseline --no-large-map --no-sustained > gpurun_out/r2n_ncu_fullM.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:lm_rank_update_kernel -s 4 -c 1 -o gpurun_out/r2n_lm python bench_large_map.py --landmarks 4096 --updates 12 --warmup 3 --fused 0 > gpurun_out/r2n_ncu_lm.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:lm_gather_kernel -s 4 -c 1 -o gpurun_out/r2n_lm_gather python bench_large_map.py --landmarks 4096 --updates 12 --warmup 3 --fused 0 > gpurun_out/r2n_ncu_lmg.log 2>&1
timeout 400 ncu --set full --clock-control none --import-source on -k regex:batch_step_regw_kernel -s 80 -c 1 -o gpurun_out/r2n_bag python tools/mc_bag_probe.py > gpurun_out/r2n_ncu_bag.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/r2n_bench.err
for n in batch batchM lm lm_gather bag; do
  f=gpurun_out/r2n_$n.ncu-rep
  if [ -f $f ]; then
    python tools/ncu_key.py $f > gpurun_out/r2n_$n.key.txt 2>&1
    python tools/ncu_lines.py $f > gpurun_out/r2n_$n.lines.txt 2>&1
    ncu -i $f --page details > gpurun_out/r2n_$n.details.txt 2>&1
    rm -f $f
  fi
done
python tools/ncu_launch_summary.py gpurun_out/r2n_launches.csv > gpurun_out/r2n_launches.summary.txt 2>&1
head -20 gpurun_out/r2n_launches.summary.txt; cat gpurun_out/r2n_batch.key.txt | head -12
