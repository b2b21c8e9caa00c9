for v in 0 2 3 4 1; do RPE_MATCH_VARIANT=$v python bench.py --no-cpu-baseline --no-stream --steps 10 > gpurun_out/mv_$v.json 2>gpurun_out/mv.err || tail -3 gpurun_out/mv.err; python - $v <<PY
import json,sys
d=json.loads(open("gpurun_out/mv_%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print("variant",sys.argv[1],"value",round(d["value"],1),"match ms",round(d["stages"]["match"]["ms_per_step"],3))
PY
done
RPE_MATCH_VARIANT=4 python -m pytest tests -m gpu -x -q -k "match or cfg3" 2>&1 | tail -2
