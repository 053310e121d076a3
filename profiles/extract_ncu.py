#!/usr/bin/env python
"""ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python profiles/extract_ncu.py X_raw.csv "title" > profiles/NAME.txt
Keeps the metrics the roofline discussion in DESIGN.md uses."""
import csv
import sys

KEEP = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors.sum", "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_tensor_subpipe_hmma.avg.pct_of_peak_sustained_active",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio"]

rows = list(csv.reader(open(sys.argv[1])))
H, U, R = rows[0], rows[1], rows[2:]
print("#", sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
print("# ncu --set full --clock-control none; one column per captured launch")
ki = H.index("Kernel Name")
print(f"{'kernel':100s}", [r[ki].split('(')[0][-40:] for r in R])
for k in KEEP:
    if k in H:
        i = H.index(k)
        print(f"{k:100s} {U[i]:16s}", [r[i] for r in R])
