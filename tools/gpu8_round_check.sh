# 8-GPU records of the final kernels: batched B (weak), batched M (strong), large map 32768 / 4096 lm (peer-memory exchange)
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531"
timeout 400 $TR bench.py --gpus 8 --no-cpu-baseline --no-large-map > gpurun_out/g8_benchB.json 2> gpurun_out/g8_benchB.err
timeout 400 $TR bench.py --gpus 8 --workload M --no-cpu-baseline --no-large-map > gpurun_out/g8_benchM.json 2> gpurun_out/g8_benchM.err
timeout 500 $TR bench_large_map.py --landmarks 32768 --updates 40 > gpurun_out/g8_lm32768.json 2> gpurun_out/g8_lm32768.err
timeout 300 $TR bench_large_map.py --landmarks 4096 --check > gpurun_out/g8_lm4096.json 2> gpurun_out/g8_lm4096.err
grep -o '"value": [0-9.]*' gpurun_out/g8_*.json
