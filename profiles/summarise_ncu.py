#!/usr/bin/env python
"""Summarise `ncu --page raw --csv` output: one block of key metrics per distinct kernel launch shape.
usage: ncu -i X.ncu-rep --page raw --csv > raw.csv ; python profiles/summarise_ncu.py raw.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__occupancy_limit_registers", "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor"]
idx = {w: hdr.index(w) for w in want if w in hdr}
kn = hdr.index("Kernel Name")
seen = set()
for r in rows[2:]:
    name = r[kn].split("(")[0]
    key = (name, r[idx["launch__grid_size"]])
    if key in seen:
        continue
    seen.add(key)
    print("\n" + name)
    for w in want:
        if w in idx:
            print(f"   {w:68s} {r[idx[w]]:>16s} {units[idx[w]]}")
