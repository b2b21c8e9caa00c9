#!/bin/bash
# usage: gpu_scale.sh N  -- bench.py on N GPUs of one box exactly as the driver launches it; prints the headline figures
N=$1
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --no-cpu-baseline > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err || tail -5 gpurun_out/bench_n$N.err
python - $N <<PY
import json,sys
d=json.loads(open("gpurun_out/bench_n%s.json"%sys.argv[1]).read().strip().splitlines()[-1])
print("n_gpus",d["n_gpus"],"value",round(d["value"],1),"e2e",round(d["e2e"]["value"],1),"ms/step",round(d["ms_per_step"],3),"stream",round(d["stream"]["value"],1),round(d["stream"]["e2e"],1),d["clocks"])
PY
