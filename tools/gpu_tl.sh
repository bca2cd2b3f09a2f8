set -x
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 tools/lm_p2p_timeline.py 4096 60 > gpurun_out/tl_${N}.txt 2> gpurun_out/tl_${N}.err; echo "rc=$?" >> gpurun_out/tl_${N}.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29562 tools/lm_p2p_timeline.py 2048 60 > gpurun_out/tl_2.txt 2> gpurun_out/tl_2.err; echo "rc=$?" >> gpurun_out/tl_2.err
cat gpurun_out/tl_${N}.txt; tail -3 gpurun_out/tl_${N}.err; cat gpurun_out/tl_2.txt; tail -3 gpurun_out/tl_2.err
