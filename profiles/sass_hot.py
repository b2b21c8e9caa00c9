#!/usr/bin/env python
"""Summarise `ncu --page source --csv --print-source sass` of ONE kernel launch: stall mix and executed opcode mix.
usage: ncu -i X.ncu-rep --page source --csv -k KERNEL --print-source sass > k.csv ; python profiles/sass_hot.py k.csv"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]; n = len(hdr)
data = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break                      # first launch only
    if len(r) >= n and r[0] != "Address":
        data.append(r)
si = hdr.index("# Samples"); ie = hdr.index("Instructions Executed")
stalls = [c for c in hdr if c.startswith("stall_") and "Not" not in c]
tot = {c: sum(int(r[hdr.index(c)] or 0) for r in data) for c in stalls}
T = max(1, sum(tot.values()))
print("stalls %:", {k: round(100 * v / T, 1) for k, v in sorted(tot.items(), key=lambda x: -x[1])[:8]})
ops = {}
for r in data:
    t = r[1].split()
    op = t[1] if t and t[0].startswith("@") else (t[0] if t else "")
    ops[op] = ops.get(op, 0) + int(r[ie] or 0)
te = sum(ops.values())
print("warp instructions:", te)
print("opcode %:", [(k, round(100 * v / te, 1)) for k, v in sorted(ops.items(), key=lambda x: -x[1])[:18]])
top = sorted(data, key=lambda r: -int(r[si] or 0))[:12]
for r in top:
    print(f"{r[si]:>7s} {r[ie]:>10s}  {r[1][:90]}")
