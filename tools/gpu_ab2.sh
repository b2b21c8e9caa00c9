# usage: gpu_ab2.sh "ENV1=a ENV2=b" "ENV1=c" ...   -- bench A/B over whole environment settings
i=0
for setting in "$@"; do i=$((i+1)); env $setting python bench.py --no-cpu-baseline --steps 20 > gpurun_out/ab2_$i.json 2>gpurun_out/ab2.err || tail -3 gpurun_out/ab2.err; python - "$setting" $i <<PY
import json,sys
d=json.loads(open("gpurun_out/ab2_%s.json"%sys.argv[2]).read().strip().splitlines()[-1])
print(sys.argv[1],"| value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"stream",round(d["stream"]["value"],1),{k:round(x["ms_per_step"],3) for k,x in d["stages"].items() if k in ("pyramid","fast","blur","describe","select","match","ransac")})
PY
done
