#!/bin/bash
# GPU-box helper: pipeline-shape sweep (engine handles, chunk pairs, chunk streams) of the bench line.
mkdir -p gpurun_out
for cs in "1 32 1" "1 128 2" "1 64 2" "1 64 3" "1 32 4" "2 32 1" "2 128 2"; do set -- $cs
  timeout 300 python bench.py --no-cpu-baseline --inflight $1 --chunk $2 --streams $3 > gpurun_out/b_$1_$2_$3.json 2>gpurun_out/b.err || tail -3 gpurun_out/b.err
  python - gpurun_out/b_$1_$2_$3.json "$cs" <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("handles/chunk/streams", sys.argv[2], "value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "stream", round(d["stream"]["value"],1))
PY
done
