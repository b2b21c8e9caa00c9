"""Config 2 line item: the reference's simulator sequence (59 frames at step 15 -> 58 pairs), one sequence call per pass,
host frames in / poses out, against the reference's cv2 call sequence on the same frames (all host threads)."""
import os, sys, time, json
import numpy as np, cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rpe_b200
from conftest import default_K
from oracle.reference_path import PathConfig, ReferencePath
D = os.path.join(ROOT, "baseline/_ref/simulator-data/images")
idx = sorted(int(n[:6]) for n in os.listdir(D))
frames = [cv2.cvtColor(cv2.imread(os.path.join(D, f"{f:06d}.png")), cv2.COLOR_BGR2GRAY) for f in idx]
K = default_K(640, 480)
est = rpe_b200.PoseEstimator(K, nfeatures=4000, max_matches=500)
for _ in range(3):
    est.estimate_sequence(frames)
n = 20
t0 = time.perf_counter()
for _ in range(n):
    R, t, status = est.estimate_sequence(frames)
dt = (time.perf_counter() - t0) / n
ref = ReferencePath(K, PathConfig())
ref.estimate(frames[0], frames[1])
t0 = time.perf_counter()
for p in range(len(frames) - 1):
    ref.estimate(frames[p], frames[p + 1])
dc = time.perf_counter() - t0
print(json.dumps({"workload": "cfg2: simulator sequence, 59 frames 640x480 -> 58 pairs, ORB 4000, cross-check, 500 matches, RANSAC 1000",
                  "gpu_ms_per_sequence": 1e3 * dt, "gpu_pairs_per_s": 58 / dt, "ok": int((status == 0).sum()),
                  "cpu_s_per_sequence": dc, "cpu_pairs_per_s": 58 / dc, "cpu_threads": cv2.getNumThreads()}))
eng = est.engine
eng.set_profiling(True)
est.estimate_sequence(frames)
ms, launches = eng.stage_times()
print({k: round(v, 3) for k, v in ms.items()}, "sum", round(sum(ms.values()), 3))
eng.set_profiling(False)
import torch
pin = torch.from_numpy(np.stack(frames)).pin_memory().numpy()
for _ in range(3): eng.estimate_sequence(pin, K)
t0 = time.perf_counter()
for _ in range(n): eng.estimate_sequence(pin, K)
print("pinned contiguous frames: ms per sequence", round(1e3 * (time.perf_counter() - t0) / n, 3))
