#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --set full` capture of `bench.py --batch 32` (tools/gpu_trip.sh ncu):
usage: ncu -i full.ncu-rep --page raw --csv > raw.csv ; python profiles/make_traffic.py raw.csv [images]"""
import csv, json, os, sys
rows = list(csv.reader(open(sys.argv[1])))
n_img = int(sys.argv[2]) if len(sys.argv) > 2 else 64
hdr = rows[0]
kn, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
assert rows[1][ir] == "Mbyte" and rows[1][iw] == "Mbyte", (rows[1][ir], rows[1][iw])
per = {}
for r in rows[2:]:
    per.setdefault(r[kn].split("(")[0].replace("void ", ""), []).append((float(r[ir]) + float(r[iw])) * 1e6)
stage = {"fast": "k_fast", "blur": "k_blur", "describe": "k_describe", "select": "k_select"}
out = {"_source": f"profiles/r02_ncu_full_b32_summary.txt: ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch "
                  f"over {n_img} images ({n_img // 2} pairs); bench.py scales by images per step"}
for s, k in stage.items():
    out[s] = {"dram_bytes_per_image": per[k][0] / n_img}
rs = [v for k, v in per.items() if k.startswith("k_resize")]
out["pyramid"] = {"dram_bytes_per_image": sum(sum(v) for v in rs) / n_img,
                  "note": f"{sum(len(v) for v in rs)} launches summed; produced levels stay in the 126 MB L2, so DRAM traffic is below the algorithmic bytes"}
out["match"] = {"dram_bytes_per_pair": [v for k, v in per.items() if k.startswith("k_hamming")][0][0] / (n_img // 2)}
json.dump(out, open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "traffic.json"), "w"), indent=1)
print(json.dumps(out, indent=1))
