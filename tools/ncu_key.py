#!/usr/bin/env python
"""Print the key metrics of the LAST kernel of an .ncu-rep (raw page). Usage: python tools/ncu_key.py file.ncu-rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, v = rows[0], rows[-1]
want = """gpu__time_duration.sum dram__bytes_read.sum dram__bytes_write.sum dram__throughput.avg.pct_of_peak_sustained_elapsed
smsp__inst_executed.sum l1tex__data_pipe_lsu_wavefronts_mem_shared.sum l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum
l1tex__data_pipe_lsu_wavefronts.sum l1tex__lsu_writeback_active.avg.pct_of_peak_sustained_elapsed l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed
sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active smsp__issue_active.avg.pct_of_peak_sustained_active
sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active
launch__occupancy_limit_shared_mem launch__occupancy_limit_registers sm__warps_active.avg.per_cycle_active launch__registers_per_thread
smsp__average_warp_latency_per_inst_issued.ratio
smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio
smsp__average_warps_issue_stalled_wait_per_issue_active.ratio smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio
smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio
smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio
smsp__average_warps_issue_stalled_drain_per_issue_active.ratio smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio""".split()
for w in want:
    if w in h:
        print(f"{w:90s} {v[h.index(w)]}")
if len(sys.argv) > 2:
    for i, name in enumerate(h):
        if sys.argv[2] in name:
            print(f"{name:90s} {v[i]}")
