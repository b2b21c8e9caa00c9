#!/bin/bash
# After `gpurun -- 'bash tools/gpu_trip.sh tests_all checked smoke bench ncu'`: turn gpurun_out/ into the artefacts committed
# under profiles/ (bench line, launch list, ncu summary, per-line profile of k_fast, DRAM traffic per kernel, test logs).
set -e
cd "$(dirname "$0")/.."
ncu -i gpurun_out/full.ncu-rep --page raw --csv > /tmp/rpe_raw.csv 2>/dev/null
python profiles/summarise_ncu.py /tmp/rpe_raw.csv > profiles/r02_ncu_full_b32_summary.txt
python profiles/make_traffic.py /tmp/rpe_raw.csv > /dev/null
ncu -i gpurun_out/full.ncu-rep --page source --csv -k regex:k_fast --print-source sass,cuda > /tmp/rpe_src_fast.csv 2>/dev/null
python profiles/src_hot.py /tmp/rpe_src_fast.csv 45 > profiles/r02_k_fast_lines.txt
cp gpurun_out/bench.json profiles/r02_bench_b256.json
cp gpurun_out/launches.csv profiles/r02_launches_b256.csv
cp gpurun_out/r02_cfg3_parity_rank0.json gpurun_out/r02_cfg3_parity_rank7.json gpurun_out/r02_cfg4_parity.json profiles/
[ -f gpurun_out/cfg1_latency.json ] && tail -1 gpurun_out/cfg1_latency.json > profiles/r02_cfg1_latency.json
{
  echo "# RPE_LIB_VARIANT=checked RPE_EXTRA_DEFS=-DRPE_CHECKED=1 python -m pytest tests -m gpu -q -x (device-side assertions on; B200)"
  cat gpurun_out/pytest_gpu_checked.log
  echo "# tools/sanitize_target.py against the checked build"; cat gpurun_out/checked_target.log
  echo "# plain build, same trip: python -m pytest tests -m gpu -q"; cat gpurun_out/pytest_gpu.log
  [ -f gpurun_out/pytest_noalloc_alone.log ] && { echo "# tests/test_zz_no_alloc.py alone in a fresh process"; cat gpurun_out/pytest_noalloc_alone.log; }
} > profiles/r02_pytest_gpu_checked_build.log
