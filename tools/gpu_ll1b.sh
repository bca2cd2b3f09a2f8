set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/ll1b_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/ll1b_suite.log
tail -8 gpurun_out/ll1b_suite.log
