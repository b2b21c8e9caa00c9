#!/bin/bash
# 8-GPU box visit: upload ceiling at N = 1, 2, 4, 8 (pinned and write-combined), then bench.py at N ranks exactly as the
# driver launches it.  usage: bash tools/gpu_scale.sh N [bench args]
N=${1:-8}; shift
mkdir -p gpurun_out
: > gpurun_out/h2d_ceiling.jsonl
for n in 1 2 4 8; do
  [ $n -le $N ] || continue
  for wc in "" "--write-combined"; do
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29517 tools/h2d_ceiling.py $wc 2>/dev/null | tail -1 >> gpurun_out/h2d_ceiling.jsonl
  done
done
cat gpurun_out/h2d_ceiling.jsonl
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1; head -14 gpurun_out/topo.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; tail -3 gpurun_out/bench_n$N.err
python tools/show_bench.py gpurun_out/bench_n$N.json
