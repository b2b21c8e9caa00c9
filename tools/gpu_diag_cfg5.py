"""Stage-by-stage diagnosis of a config-5 stream (1280x720, cfg3 parameters) against the reference's cv2 call sequence."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rpe_b200
from rpe_b200 import _lib as L
from bench_support.synth import synth_K, synth_stream, stream_relative_pose
from oracle import geometry as og
from oracle.reference_path import PathConfig, ReferencePath
W, H, F = 1280, 720, int(os.environ.get("F", "25"))
first = int(os.environ.get("FIRST", "1000")); step = int(os.environ.get("STEP", "1"))
K = synth_K(W, H)
frames = synth_stream(5, first, F, W, H, step=step)
eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=F - 1, nfeatures=5000, nlevels=12,
                                            match_mode="ratio", ratio=0.75, max_matches=500, ransac_max_iters=2000))
out = eng.estimate_sequence(frames, K)
info = out["info"]
ref = ReferencePath(K, PathConfig(nfeatures=5000, nlevels=12, match_mode="ratio", ratio=0.75, max_matches=500,
                                  ransac_max_iters=2000), canonical=True)
bad = {}
for p in range(F - 1):
    d = ref.estimate_with_debug(frames[p], frames[p + 1])
    dr = og.rotation_angle_deg(out["R"][p], d["R"]); dt = og.direction_angle_deg(out["t"][p], d["t"])
    m = eng.get_stage(p, 0, L.STAGE_MATCHES, 0, np.int32).reshape(-1, 3)
    E = eng.get_stage(p, 0, L.STAGE_E, 0, np.float64)
    rm = eng.get_stage(p, 0, L.STAGE_RANSAC_MASK, 0, np.uint8)[:len(m)]
    same_m = m.shape == d["matches"].shape and np.array_equal(m, d["matches"])
    Er = d["E"].ravel()
    dE = min(np.abs(E - Er).max(), np.abs(E + Er).max())
    mask_diff = int((rm != d["ransac_mask"].ravel()[:len(rm)]).sum()) if len(rm) == d["ransac_mask"].size else -1
    Rg, tg = stream_relative_pose(5, 0, first + p * step, W, H, step)
    flag = dr > 0.01 or dt > 0.05 or not same_m
    print(f"{'BAD ' if flag else 'ok  '}pair {p}: dR {dr:.4f} dt {dt:.4f} matches_equal {same_m} dE {dE:.2e} mask_diff {mask_diff} "
          f"gpu(inl {info[p,2]} it {info[p,3]} pose {info[p,1]}) ref(inl {int(d['ransac_mask'].sum())} pose {d['inliers']}) "
          f"gt: gpu dR {og.rotation_angle_deg(out['R'][p], Rg):.3f} dt {og.direction_angle_deg(out['t'][p], Rg @ np.zeros(3) + tg):.2f} "
          f"ref dR {og.rotation_angle_deg(d['R'], Rg):.3f} dt {og.direction_angle_deg(d['t'], tg):.2f}")
    if flag:
        bad[p] = d
print("pairs checked", F - 1, "flagged", len(bad))
if bad:
    p = sorted(bad)[0]; d = bad[p]
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", "cfg5_bad.npz"), pair=p, pts1=d["pts1"], pts2=d["pts2"], E_ref=d["E"],
                        mask_ref=d["ransac_mask"], E_gpu=eng.get_stage(p, 0, L.STAGE_E, 0, np.float64),
                        mask_gpu=eng.get_stage(p, 0, L.STAGE_RANSAC_MASK, 0, np.uint8),
                        R_gpu=out["R"][p], t_gpu=out["t"][p], R_ref=d["R"], t_ref=d["t"], K=K, info=info[p])
