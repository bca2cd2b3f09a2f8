# ncu --set full of the large-map kernels of the final build (one launch each) + launch list of bench_large_map.py
set -x
mkdir -p gpurun_out
timeout 300 python bench_large_map.py --landmarks 4096 > gpurun_out/n_lm4096.json 2> gpurun_out/n_lm4096.err
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_rank_update_kernel -s 10 -c 1 -o gpurun_out/n_lm python bench_large_map.py --landmarks 4096 --updates 20 --fused 0 > gpurun_out/n_ncu_lm.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_rank_update_multi -s 3 -c 1 -o gpurun_out/n_lm_multi python bench_large_map.py --landmarks 4096 --updates 12 > gpurun_out/n_ncu_lm_multi.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:lm_gather -s 10 -c 2 -o gpurun_out/n_lm_gather python bench_large_map.py --landmarks 4096 --updates 12 > gpurun_out/n_ncu_lm_gather.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/n_lm_launches.csv python bench_large_map.py --landmarks 4096 --updates 30 > gpurun_out/n_ncu_lm_launch.log 2>&1
ls -la gpurun_out | tail -n 12
