# round 2, final tree: ncu of the large-map kernel the final code launches on one GPU (single launch per update) and the
# launch list of the default bench run. Every profiled command first runs once WITHOUT ncu.
set -x
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "rc=$?" >> gpurun_out/r2x_bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/r2x_launches.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r2x_ncu_launch.log 2>&1
timeout 300 python bench_large_map.py --landmarks 4096 --updates 12 --warmup 3 --fused 0 > gpurun_out/r2x_lm.json 2> gpurun_out/r2x_lm.err; echo "rc=$?" >> gpurun_out/r2x_lm.err
timeout 400 ncu --set full --clock-control none --import-source on -k regex:lm_update_single_launch_kernel -s 4 -c 1 -o gpurun_out/r2x_lm_single python bench_large_map.py --landmarks 4096 --updates 12 --warmup 3 --fused 0 > gpurun_out/r2x_ncu_lm.log 2>&1
ls -la gpurun_out/*.ncu-rep; tail -2 gpurun_out/r2x_bench.err; tail -2 gpurun_out/r2x_lm.err
f=gpurun_out/r2x_lm_single.ncu-rep
if [ -f $f ]; then
  python tools/ncu_key.py $f > gpurun_out/r2x_lm_single.key.txt 2>&1
  python tools/ncu_lines.py $f > gpurun_out/r2x_lm_single.lines.txt 2>&1
  ncu -i $f --page details > gpurun_out/r2x_lm_single.details.txt 2>&1
  rm -f $f
fi
python tools/ncu_launch_summary.py gpurun_out/r2x_launches.csv > gpurun_out/r2x_launches.summary.txt 2>&1
head -30 gpurun_out/r2x_launches.summary.txt; head -12 gpurun_out/r2x_lm_single.key.txt
