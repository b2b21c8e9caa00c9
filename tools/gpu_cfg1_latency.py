"""Config 1 (the reference's default single pair, 640x480, ORB 4000, cross-check, 500 matches, RANSAC 1000): latency of ONE
PoseEstimator.estimate() call on the GPU (host frames in, pose out, blocking) with the per-stage CUDA-event times, beside
the reference's cv2 call sequence on the same pair with all host threads."""
import json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rpe_b200
from conftest import default_K, load_gray
from oracle.reference_path import PathConfig, ReferencePath
a, b = load_gray("sim_000000.png"), load_gray("sim_000015.png")
K = default_K(640, 480)
est = rpe_b200.PoseEstimator(K, nfeatures=4000, max_matches=500)
for _ in range(5):
    est.estimate(a, b)
n = 200
t0 = time.perf_counter()
for _ in range(n):
    R, t = est.estimate(a, b)
gpu_ms = 1e3 * (time.perf_counter() - t0) / n
eng = est.engine
eng.set_pipeline(32, 1)
for _ in range(5):
    est.estimate(a, b)
t0 = time.perf_counter()
for _ in range(n):
    est.estimate(a, b)
gpu_ms_1stream = 1e3 * (time.perf_counter() - t0) / n
eng.set_profiling(True)
est.estimate(a, b)
ms, launches = eng.stage_times()
eng.set_profiling(False)
ref = ReferencePath(K, PathConfig())
ref.estimate(a, b)
t0 = time.perf_counter()
for _ in range(10):
    ref.estimate(a, b)
cpu_ms = 1e2 * (time.perf_counter() - t0)
print(json.dumps({"workload": "cfg1: one 640x480 pair, ORB 4000, cross-check, 500 matches, RANSAC 1000; PoseEstimator.estimate(), blocking",
                  "gpu_ms_per_call": gpu_ms, "gpu_ms_per_call_unchunked": gpu_ms_1stream, "cpu_ms_per_call_cv2_all_threads": cpu_ms,
                  "stages_ms": {k: round(v, 4) for k, v in ms.items()}, "stage_sum_ms": round(sum(ms.values()), 4),
                  "launches": int(sum(launches.values())) if hasattr(launches, "values") else launches}))
