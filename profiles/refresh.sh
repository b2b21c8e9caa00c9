#!/bin/bash
# usage: profiles/refresh.sh vN   -- copy the artefacts of the last `tools/gpu_round.sh ncu` run from gpurun_out/ into profiles/ as round-1 vN
V=$1
cp gpurun_out/bench.json profiles/r01_bench_b256_$V.json
cp gpurun_out/launches.csv profiles/r01_launches_b256_$V.csv
ncu -i gpurun_out/full.ncu-rep --page raw --csv > /tmp/raw_$V.csv 2>/dev/null
python profiles/summarise_ncu.py /tmp/raw_$V.csv > profiles/r01_ncu_full_b32_summary_$V.txt
python - $V <<'PY'
import re, json, sys
V = sys.argv[1]
txt = open(f'profiles/r01_ncu_full_b32_summary_{V}.txt').read()
agg = {}
for b in re.split(r'\n(?=\S)', txt):
    lines = b.strip().splitlines()
    if not lines: continue
    def val(key):
        for l in lines:
            if key in l:
                p = l.split(); return float(p[1]) * {'Mbyte': 1e6, 'Kbyte': 1e3, 'Gbyte': 1e9, 'byte': 1}.get(p[2], 1)
        return 0
    agg.setdefault(lines[0].strip(), []).append(val('dram__bytes_read.sum') + val('dram__bytes_write.sum'))
t = json.load(open('profiles/traffic.json'))
t['_source'] = re.sub(r'_v\d+', '_' + V, t['_source'])
for stage, k in (('fast', 'k_fast'), ('blur', 'k_blur'), ('describe', 'k_describe'), ('select', 'k_select')):
    t[stage]['dram_bytes_per_image'] = sum(agg[k]) / 64
pyr = [k for k in agg if 'k_resize' in k]
t['pyramid']['dram_bytes_per_image'] = sum(sum(agg[k]) for k in pyr) / 64
hm = [k for k in agg if 'k_hamming' in k][0]
t['match']['dram_bytes_per_pair'] = sum(agg[hm]) / 32
json.dump(t, open('profiles/traffic.json', 'w'), indent=1)
print({k: (round(v.get('dram_bytes_per_image', v.get('dram_bytes_per_pair', 0)) / 1e6, 2) if isinstance(v, dict) else '') for k, v in t.items()})
PY
