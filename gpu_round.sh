#!/bin/bash
# GPU-box helper: parity tests, bench (both arms), ncu launch list + full capture of the top kernels.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench.json").read().strip().splitlines()[-1])
print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/step", round(d["ms_per_step"],3), "cpu", d["cpu_baseline"])
print({k:round(v["ms_per_step"],3) for k,v in d["stages"].items()})
print(d["roofline"]); print(d["clocks"])
PY
if [ "$1" = "ncu" ]; then
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_blur|k_hamming|k_resize|k_ransac|k_describe|k_pose|k_select|k_build' --launch-skip 20 -c 20 -o gpurun_out/full -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --batch 32 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
