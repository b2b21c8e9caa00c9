"""Stage-by-stage diagnosis of config 2 (simulator sequence) against the reference's cv2 call sequence."""
import os, sys
import numpy as np, cv2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rpe_b200
from rpe_b200 import _lib as L
from conftest import default_K
from oracle import geometry as og
from oracle.reference_path import PathConfig, ReferencePath
D = os.path.join(ROOT, "baseline/_ref/simulator-data/images")
idx = sorted(int(n[:6]) for n in os.listdir(D))
frames = [cv2.cvtColor(cv2.imread(os.path.join(D, f"{f:06d}.png")), cv2.COLOR_BGR2GRAY) for f in idx]
K = default_K(640, 480)
est = rpe_b200.PoseEstimator(K, nfeatures=4000, max_matches=500)
R, t, status = est.estimate_sequence(frames)
eng = est.engine
info = est.last["info"]
ref = ReferencePath(K, PathConfig(), canonical=True)
bad = {}
for p in range(len(frames) - 1):
    d = ref.estimate_with_debug(frames[p], frames[p + 1])
    dr = og.rotation_angle_deg(R[p], d["R"]); dt = og.direction_angle_deg(t[p], d["t"])
    m = eng.get_stage(p, 0, L.STAGE_MATCHES, 0, np.int32).reshape(-1, 3)
    E = eng.get_stage(p, 0, L.STAGE_E, 0, np.float64)
    rm = eng.get_stage(p, 0, L.STAGE_RANSAC_MASK, 0, np.uint8)[:len(m)]
    same_m = m.shape == d["matches"].shape and np.array_equal(m, d["matches"])
    Er = d["E"].ravel()
    dE = min(np.abs(E - Er).max(), np.abs(E + Er).max())
    mask_eq = rm.shape == d["ransac_mask"].shape and int((rm != d["ransac_mask"]).sum())
    if dr > 0.01 or dt > 0.05 or not same_m:
        print(f"pair {p} frames {idx[p]}->{idx[p+1]}: dR {dr:.4f} dt {dt:.4f} matches_equal {same_m} dE {dE:.2e} mask_diff {mask_eq} "
              f"gpu(inl {info[p,2]} it {info[p,3]} pose {info[p,1]}) ref(inl {int(d['ransac_mask'].sum())} pose {d['inliers']})")
        bad[p] = d
print("pairs checked", len(frames) - 1, "flagged", len(bad))
if bad:
    p = sorted(bad)[0]; d = bad[p]
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "cfg2_bad.npz"), pair=p, pts1=d["pts1"], pts2=d["pts2"], E_ref=d["E"], mask_ref=d["ransac_mask"],
                        E_gpu=eng.get_stage(p, 0, L.STAGE_E, 0, np.float64), R_gpu=R[p], t_gpu=t[p], R_ref=d["R"], t_ref=d["t"], K=K)
