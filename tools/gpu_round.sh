#!/bin/bash
# GPU-box helper: parity tests, smoke, bench; optional ncu launch list + full capture.
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" ; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
show() { python - "$1" <<PY
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print("value", round(d["value"],1), "e2e", round(d["e2e"]["value"],1), "ms/step", round(d["ms_per_step"],3), "serial", round(d["roofline"]["serial_ms_per_step"],3), "cpu", (d["cpu_baseline"] or {}).get("value"))
print({k:round(v["ms_per_step"],3) for k,v in d["stages"].items()})
print("stream", d.get("stream"))
print("roofline", d["roofline"]["kernel"], round(d["roofline"]["frac"],3), "xu", round(d["roofline"]["xu_pipe"]["frac"],3), d["clocks"])
PY
}
timeout 600 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err; show gpurun_out/bench.json
for extra in "$@"; do
  case "$extra" in
    ncu)
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --inflight 1 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
      timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_blur|k_hamming|k_resize|k_ransac|k_describe|k_pose|k_select|k_build' --launch-skip 20 -c 20 -o gpurun_out/full -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --batch 32 --inflight 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?" ;;
    sweep)
      for cs in "1 128 1" "2 128 1" "3 128 1" "2 64 3" "2 32 4" "1 32 4" "3 64 3"; do set -- $cs
        timeout 300 python bench.py --no-cpu-baseline --inflight $1 --chunk $2 --streams $3 > gpurun_out/b_$1_$2_$3.json 2>gpurun_out/b.err || tail -3 gpurun_out/b.err; echo "inflight $1 chunk $2 streams $3"; show gpurun_out/b_$1_$2_$3.json 2>/dev/null | head -1; done ;;
  esac
done
