#!/usr/bin/env python
"""Print the handful of ncu metrics the roofline discussion needs from a .ncu-rep (run where ncu is installed)."""
import csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
hdr, units = rows[0], rows[1]
KEYS = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "sass__inst_executed_shared_loads", "sass__inst_executed_shared_stores", "sass__inst_executed_global_loads",
        "sass__inst_executed_global_stores", "l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled", "smsp__average_warps_issue_stalled", "sm__cycles_elapsed.max",
        "smsp__cycles_active.avg", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__lsu_writeback_active", "smsp__inst_executed_pipe_lsu", "smsp__inst_executed_pipe_alu", "smsp__inst_executed_pipe_fma",
        "smsp__inst_executed_pipe_fp64", "smsp__inst_executed_pipe_xu", "smsp__inst_executed_pipe_cbu", "smsp__inst_executed_pipe_adu",
        "smsp__inst_executed_pipe_uniform", "smsp__pcsamp_warps_issue_stalled"]
for r in rows[2:]:
    for h, u, v in zip(hdr, units, r):
        if any(h.startswith(k) for k in KEYS) and v not in ("", "0"):
            print(f"{h} [{u}] = {v}")
    print("-" * 60)
