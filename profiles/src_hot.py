#!/usr/bin/env python
"""Per-source-line instruction / stall-sample shares of ONE kernel from an ncu report captured with --import-source on.
usage: ncu -i X.ncu-rep --page source --csv -k regex:KERNEL --print-source sass,cuda > k.csv ; python profiles/src_hot.py k.csv [N]"""
import csv, sys
num = lambda v: int(v) if v.strip().isdigit() else 0
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
seen, cur, hdr = {}, None, None
for r in rows:
    if r and r[0] == "File Path": cur = r[1].split("/")[-1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr and len(r) > 8 and r[0].isdigit() and (cur, int(r[0])) not in seen:
        seen[(cur, int(r[0]))] = (r[1].strip()[:100], num(r[hdr.index("Instructions Executed")]), num(r[hdr.index("# Samples")]))
tot = sum(v[1] for v in seen.values()) or 1; ts = sum(v[2] for v in seen.values()) or 1
print("warp instructions", tot, "samples", ts)
for (f, l), (s, ie, sm) in sorted(seen.items(), key=lambda x: -x[1][2])[:top]:
    print(f"{f[:16]:16s} {l:4d} {100*ie/tot:5.1f}% inst {100*sm/ts:5.1f}% smp  {s}")
