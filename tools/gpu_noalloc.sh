# the no-alloc entry point: alone in a fresh process, then after the whole GPU suite
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/noalloc_suite.log 2>&1; echo "suite rc=$?" >> gpurun_out/noalloc_suite.log
tail -5 gpurun_out/noalloc_alone.log; tail -5 gpurun_out/noalloc_suite.log
