# round 2: which spin of the peer-memory exchange times out at 32,768 landmarks (4 GPUs, per-phase error words)
set -x
mkdir -p gpurun_out
EKF_BENCH_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2j_bench4.json 2> gpurun_out/r2j_bench4.err
grep debug gpurun_out/r2j_bench4.err
EKF_LM_SINGLE=0 EKF_BENCH_DEBUG=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus 4 --steps 5 --warmup 3 > gpurun_out/r2j_bench4_two.json 2> gpurun_out/r2j_bench4_two.err
grep debug gpurun_out/r2j_bench4_two.err
