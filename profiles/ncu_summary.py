#!/usr/bin/env python3
"""Key metrics of every kernel in an .ncu-rep (ncu --set full) as JSON.  usage: ncu_summary.py REPORT.ncu-rep [NAME]"""
import csv, io, json, subprocess, sys
WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__shared_mem_per_block_dynamic", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed.avg.per_cycle_active",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "sass__inst_executed_local_loads", "sass__inst_executed_local_stores",
        "memory_l2_theoretical_sectors_local", "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio", "sm__icc_request_hit_rate.pct", "sm__icc_requests.sum",
        "l1tex__t_sector_pipe_lsu_mem_local_op_ld_hit_rate.pct", "l1tex__t_sector_pipe_lsu_mem_local_op_st_hit_rate.pct", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
h, units = rows[0], rows[1]
out = {}
for n, r in enumerate(rows[2:]):
    d = {"kernel": r[h.index("Kernel Name")]}
    for w in WANT:
        if w in h:
            i = h.index(w)
            try: d[w] = {"value": float(r[i].replace(",", "")), "unit": units[i]}
            except ValueError: pass
    out[(sys.argv[2] if len(sys.argv) > 2 else "launch") + (f"_{n}" if len(rows) > 3 else "")] = d
print(json.dumps(out, indent=1))
