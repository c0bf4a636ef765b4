#!/usr/bin/env python
"""Prints the metrics DESIGN.md quotes from an `ncu -i X.ncu-rep --page raw --csv` dump.  usage: ncu_metrics.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h, units, v = rows[0], rows[1], rows[2]
keep = ("gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__", "sm__inst_executed_pipe_",
        "sm__pipe_fma_cycles_active", "sm__pipe_alu_cycles_active", "sm__throughput.avg.pct", "sm__warps_active.avg.pct",
        "smsp__issue_active.avg.pct", "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__warps_eligible.avg.per_cycle_active",
        "smsp__inst_executed.sum", "l1tex__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__average_warps_issue_stalled_")
for k, u, x in sorted(zip(h, units, v)):
    if any(k.startswith(p) for p in keep) and not k.endswith(".max_rate") and "pct_of_peak_sustained_elapsed" not in k.replace("sm__throughput", ""):
        print(f"{k:100s} {u:16s} {x}")
