#!/usr/bin/env python
"""tools/ncu_pick.py REPORT.ncu-rep [substr ...]: print selected raw metrics of the first kernel in a report."""
import csv, subprocess, sys
rep = sys.argv[1]
want = sys.argv[2:] or [
    "gpu__time_duration.sum", "dram__bytes_read.sum ", "dram__bytes_write.sum ", "launch__registers_per_thread",
    "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum ", "sm__inst_executed_pipe_tex.sum ",
    "l1tex__data_pipe_tex_wavefronts.sum ", "l1tex__data_pipe_lsu_wavefronts.sum ", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum ",
    "l1tex__data_pipe_tex_wavefronts.sum.pct", "l1tex__data_pipe_lsu_wavefronts.sum.pct", "l1tex__f_wavefronts.sum.pct",
    "l1tex__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_tex", "l1tex__t_requests_pipe_tex", "lts__t_sector_hit_rate.pct",
    "lts__t_sectors_srcunit_tex.sum ", "l1tex__throughput.avg.pct", "lts__throughput.avg.pct", "dram__throughput.avg.pct",
    "sm__throughput.avg.pct", "smsp__average_warps_issue_stalled", "smsp__average_warp_latency_issue_stalled", "l1tex__texin",
    "sm__cycles_elapsed.max", "l1tex__m_xbar2l1tex_read_bytes.sum ", "smsp__thread_inst_executed_per_inst_executed",
    "sm__inst_executed_pipe_", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum "]
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u, v = rows[0], rows[1], rows[2]
for i, n in enumerate(h):
    key = n + " "
    if any(w in key for w in want):
        try:
            if float(v[i].replace(",", "")) == 0.0 and "stalled" in n: continue
        except ValueError: pass
        print(f"{n:90s} {v[i]:>16s} {u[i]}")
