# N-GPU records of the final kernels (N = $1): batched B (weak), batched M (strong), large map 4096 lm sharded
set -x
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531"
timeout 400 $TR bench.py --gpus $N --no-cpu-baseline --no-large-map > gpurun_out/g${N}_benchB.json 2> gpurun_out/g${N}_benchB.err
timeout 400 $TR bench.py --gpus $N --workload M --no-cpu-baseline --no-large-map > gpurun_out/g${N}_benchM.json 2> gpurun_out/g${N}_benchM.err
timeout 300 $TR bench_large_map.py --landmarks 4096 --check > gpurun_out/g${N}_lm4096.json 2> gpurun_out/g${N}_lm4096.err
grep -o '"value": [0-9.]*' gpurun_out/g${N}_*.json
