#!/usr/bin/env python
"""Cut an `ncu --page raw --csv` export down to the columns the profiles/ summaries keep.
Usage: python tools/ncu_summary.py raw.csv "header comment" > profiles/xxx_summary.csv"""
import csv
import sys

COLS = ["ID", "Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum",
        "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__waves_per_multiprocessor", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max"]

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
idx = [hdr.index(c) for c in COLS if c in hdr]
w = csv.writer(sys.stdout)
if len(sys.argv) > 2:
    print('"# %s"' % sys.argv[2].replace('"', '""'))
for r in rows:
    if len(r) == len(hdr):
        w.writerow([r[i] for i in idx])
