#!/bin/bash
# GPU-box helper: parity tests, then a short bench with the per-stage breakdown.
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -8
timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench2.log 2>gpurun_out/bench2.err || tail -5 gpurun_out/bench2.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench2.log").read().strip().splitlines()[-1])
print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/step", round(d["ms_per_step"],3))
print({k:round(v["ms_per_step"],3) for k,v in d["stages"].items()})
PY
