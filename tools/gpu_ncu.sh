#!/bin/bash
# usage: gpu_ncu.sh <kernel-regex> <out-name> [extra bench args]   -- one ncu --set full capture of the named kernels
K="$1"; O="$2"; shift 2
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$K" --launch-skip 4 -c 6 -o gpurun_out/$O -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --batch 32 --inflight 1 "$@" > gpurun_out/ncu_$O.log 2>&1; echo "ncu rc=$?"
