# usage: gpu_ab.sh ENVVAR v1 v2 ...   -- bench A/B over values of one environment knob
VAR=$1; shift
for v in "$@"; do env $VAR=$v python bench.py --no-cpu-baseline --steps 20 > gpurun_out/ab_$v.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err; python - $VAR $v <<PY
import json,sys
d=json.loads(open("gpurun_out/ab_%s.json"%sys.argv[2]).read().strip().splitlines()[-1])
print(sys.argv[1],sys.argv[2],"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"stream",round(d["stream"]["value"],1),{k:round(x["ms_per_step"],3) for k,x in d["stages"].items() if k in ("pyramid","fast","blur","describe","select","match","ransac")})
PY
done
