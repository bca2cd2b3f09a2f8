# round 2, fourth GPU check (1 GPU): tests after the p2p sequence fix, launch list of the bag-shaped Monte Carlo
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/r2d_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2d_pytest.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/r2d_mc_launches.csv python tools/mc_bag_probe.py > gpurun_out/r2d_mc_ncu.log 2>&1
tail -8 gpurun_out/r2d_pytest.log
python tools/ncu_launch_summary.py gpurun_out/r2d_mc_launches.csv 2>&1 | head -30
