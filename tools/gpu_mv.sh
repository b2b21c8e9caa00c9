# usage: gpu_mv.sh v1 v2 ...  -- bench the matcher variants (RPE_MATCH_VARIANT), then matcher parity with the last one
for v in "$@"; do RPE_MATCH_VARIANT=$v python bench.py --no-cpu-baseline --no-stream --steps 10 > gpurun_out/mv_$v.json 2>gpurun_out/mv.err || tail -3 gpurun_out/mv.err; python - $v <<PY
import json,sys
d=json.loads(open("gpurun_out/mv_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print("variant",sys.argv[1],"value",round(d["value"],1),"match ms",round(d["stages"]["match"]["ms_per_step"],3))
PY
done
RPE_MATCH_VARIANT=$v python -m pytest tests -m gpu -x -q -k "match or cfg3 or smoke or cfg1" 2>&1 | tail -2
