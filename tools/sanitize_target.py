"""Workload for compute-sanitizer (memcheck / racecheck): smoke() + a 4-pair 1280x720 cfg3 batch (ratio mode) + a
cross-check sequence of 3 frames, all through the C ABI.  Small on purpose: the tools slow kernels 10-100x."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as g   # noqa: E402
import rpe_b200               # noqa: E402
from bench_support.synth import synth_K, synth_batch   # noqa: E402

g.smoke()
W, H, B = 1280, 720, 4
frames = synth_batch(3, 0, B, W, H)
K = synth_K(W, H)
eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=B, nfeatures=5000, match_mode="ratio",
                                             max_matches=500, ransac_max_iters=2000))
out = eng.estimate_batch_interleaved(frames, K)
assert (out["status"] == 0).all(), out["status"]
eng.submit_interleaved(frames, K)
eng.submit_interleaved(frames, K)
a, b = eng.wait(), eng.wait()
assert np.array_equal(a["R"], out["R"]) and np.array_equal(b["R"], out["R"])
eng.close()
eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=2, nfeatures=4000))
o = eng.estimate_sequence(np.ascontiguousarray(frames[[0, 1, 0]]), K)
assert (o["status"] == 0).all()
noise = np.random.default_rng(0).integers(0, 255, (2, 480, 640), dtype=np.uint8)      # dense-tile path of k_fast
eng.estimate_batch_interleaved(np.ascontiguousarray(np.pad(noise, ((0, 0), (0, 240), (0, 640)))), K)
eng.close()
print("sanitize target ok")
