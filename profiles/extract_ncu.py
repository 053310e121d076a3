#!/usr/bin/env python
"""ncu -i X.ncu-rep --page raw --csv > X_raw.csv ; python profiles/extract_ncu.py X_raw.csv "title" > profiles/NAME.txt
Keeps the metrics the roofline discussion in DESIGN.md uses. With a third argument (a .json path) it also writes the record
bench.py reads for `roofline.traffic` / `roofline.ncu` (profiles/ncu_eval_tc_current.json): the LARGEST captured launch."""
import csv
import json
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
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio", "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio", "smsp__inst_executed.sum", "l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed",
        "l1tex__m_xbar2l1tex_read_bytes.sum", "lts__t_sectors_srcunit_tex.sum", "sm__inst_executed_pipe_tex.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"]

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

if len(sys.argv) > 3:
    ti = H.index("gpu__time_duration.sum")
    big = max(R, key=lambda r: float(r[ti]))

    def val(name):
        if name not in H:
            return None
        i = H.index(name)
        v = float(big[i])
        u = U[i]
        return v * {"Mbyte": 1e6, "Kbyte": 1e3, "Gbyte": 1e9, "byte": 1.0}.get(u, 1.0)

    rec = {"source": sys.argv[1], "kernel": big[ki], "launch_us": float(big[ti]),
           "dram_bytes": val("dram__bytes_read.sum") + val("dram__bytes_write.sum"),
           "dram_pct": val("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"), "l2_pct": val("lts__throughput.avg.pct_of_peak_sustained_elapsed"),
           "l2_hit_pct": val("lts__t_sector_hit_rate.pct"), "l1tex_pct": val("l1tex__throughput.avg.pct_of_peak_sustained_elapsed"),
           "tex_data_pipe_pct": val("l1tex__data_pipe_tex_wavefronts.avg.pct_of_peak_sustained_elapsed"), "l1_hit_pct": val("l1tex__t_sector_hit_rate.pct"),
           "issue_active_pct": val("smsp__issue_active.avg.pct_of_peak_sustained_active"), "warps_active_pct": val("sm__warps_active.avg.pct_of_peak_sustained_active"),
           "tensor_pipe_pct": val("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active"), "warp_instructions": val("smsp__inst_executed.sum"),
           "stall_long_scoreboard": val("smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio"),
           "stall_tex_throttle": val("smsp__average_warps_issue_stalled_tex_throttle_per_issue_active.ratio"),
           "registers_per_thread": val("launch__registers_per_thread")}
    json.dump(rec, open(sys.argv[3], "w"), indent=1)
