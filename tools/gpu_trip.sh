#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench (+ A/B legs), pipe rates, optional sanitizer / ncu captures.
# usage: bash tools/gpu_trip.sh [tests|tests_all] [checked] [smoke] [bench] [rates] [sanitize] [ncu] [cfg2]
mkdir -p gpurun_out
for what in "$@"; do
  case "$what" in
    tests)
      timeout 1500 python -m pytest tests -m gpu -q -x --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log ;;
    tests_all)
      timeout 1800 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -40 gpurun_out/pytest_gpu.log ;;
    checked)
      # the GPU suite against the checked build (device-side assertions on every hand-computed index; stands in for
      # compute-sanitizer, which is closed on this pool)
      RPE_LIB_VARIANT=checked RPE_EXTRA_DEFS=-DRPE_CHECKED=1 timeout 1800 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu_checked.log 2>&1; echo "checked pytest rc=$?"; tail -6 gpurun_out/pytest_gpu_checked.log
      RPE_LIB_VARIANT=checked RPE_EXTRA_DEFS=-DRPE_CHECKED=1 timeout 600 python tools/sanitize_target.py > gpurun_out/checked_target.log 2>&1; echo "checked target rc=$?"; tail -2 gpurun_out/checked_target.log ;;
    ab:*)
      # A/B over variant libraries (RPE_LIB_VARIANT=x RPE_EXTRA_DEFS=-DX=1 python relative-pose-estimation_b200/build.py): ab:base,x,y
      for v in $(echo "${what#ab:}" | tr ',' ' '); do
        RPE_LIB_VARIANT=$v timeout 300 python bench.py --lean --no-cpu-baseline --steps 10 > gpurun_out/bench_ab_$v.json 2> gpurun_out/bench_ab.err || tail -3 gpurun_out/bench_ab.err
        echo "A/B $v"; python tools/show_bench.py gpurun_out/bench_ab_$v.json | head -2; done ;;
    match_ab)
      # cross-check matcher: this round's one-pass kernel against round 1's two-direction launch (a variant library
      # linked here from round 1's kernels_match.cu: librpe_b200_r01match.so) on the simulator sequence (config 2)
      for v in "" r01match; do
        RPE_LIB_VARIANT=$v timeout 300 python tools/gpu_cfg2_time.py > gpurun_out/cfg2_time_${v:-r02}.log 2>&1; echo "cfg2 ${v:-r02} rc=$?"; sed -n 2p gpurun_out/cfg2_time_${v:-r02}.log; done ;;
    smoke)
      timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log ;;
    bench)
      timeout 900 python bench.py > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; tail -3 gpurun_out/bench.err
      python tools/show_bench.py gpurun_out/bench.json ;;
    rates)
      timeout 120 tools/build/pipe_rates > gpurun_out/pipe_rates.json; cat gpurun_out/pipe_rates.json ;;
    sanitize)
      timeout 1200 compute-sanitizer --tool memcheck --error-exitcode 9 python tools/sanitize_target.py > gpurun_out/sanitizer_memcheck.log 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/sanitizer_memcheck.log
      timeout 1500 compute-sanitizer --tool racecheck --error-exitcode 9 python tools/sanitize_target.py > gpurun_out/sanitizer_racecheck.log 2>&1; echo "racecheck rc=$?"; tail -4 gpurun_out/sanitizer_racecheck.log ;;
    ncu)
      timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --lean --inflight 1 > gpurun_out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
      timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'k_fast|k_blur|k_hamming|k_resize|k_ransac_solve|k_describe|k_pose|k_select|k_build' --launch-skip 20 -c 24 -o gpurun_out/full -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline --lean --batch 32 --inflight 1 > gpurun_out/ncu_full.log 2>&1; echo "ncu full rc=$?" ;;
    cfg2)
      timeout 300 python tools/gpu_cfg2_time.py > gpurun_out/cfg2_time.log 2>&1; echo "cfg2 rc=$?"; tail -5 gpurun_out/cfg2_time.log
      timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_ransac_solve|k_ransac_score|k_hamming' --launch-skip 2 -c 9 -o gpurun_out/cfg2 -f python tools/gpu_cfg2_time.py > gpurun_out/ncu_cfg2.log 2>&1; echo "ncu cfg2 rc=$?" ;;
  esac
done
