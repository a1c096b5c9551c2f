#!/usr/bin/env python
"""ncu --page raw --csv dump of ONE captured launch -> the short metric,unit,value summary kept under profiles/.
usage: ncu -i x.ncu-rep --page raw --csv > raw.csv ; ncu_summary.py raw.csv > summary.csv"""
import csv, sys
KEEP = ("dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__time_duration.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "l1tex__t_sector_hit_rate.pct", "launch__block_size", "launch__grid_size", "launch__registers_per_thread",
        "lts__t_sector_hit_rate.pct", "lts__t_sectors_srcunit_tex_lookup_hit.sum", "lts__t_sectors_srcunit_tex_op_read.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active")
rows = list(csv.reader(open(sys.argv[1])))
names, units, vals = rows[0], rows[1], rows[2]
print("metric,unit,value")
print("kernel,,%s" % vals[names.index("Kernel Name")].replace(",", ";"))
for n, u, v in sorted(zip(names, units, vals)):
    if n in KEEP or (n.startswith("smsp__average_warps_issue_stalled") and n.endswith("per_issue_active.ratio")):
        print("%s,%s,%s" % (n, u, v))
