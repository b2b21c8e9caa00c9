"""SURVEY 8 f3 settled with data: how often does the reference's VP refinement change a result?

Runs the UNMODIFIED reference PoseEstimator (imported from /root/reference, plotly stubbed) with
use_vp_refinement=True over the reference's own three datasets at the step its runners use
(simulator 15, phone 5, salah 5), exactly as BatchProcessor.process_sequence does (R_prev from the
ground truth), and counts per dataset: pairs, pairs whose two VP gates pass
(src/core/pose_estimator.py:545-550), pairs whose R_rel changes (:565-567).

CPU only, build container only (needs /root/reference).  Output: profiles/r02_vp_gate_count.json
"""
from __future__ import annotations

import json
import os
import sys
import time
import types

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "profiles", "r02_vp_gate_count.json")


def main():
    for name in ("plotly", "plotly.graph_objects", "plotly.subplots", "matplotlib", "matplotlib.pyplot",
                 "matplotlib.animation"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, REF)
    from src.core.ground_truth_loader import GroundTruthLoader
    from src.core.pose_estimator import PoseEstimator
    from src.utils.geometry import euler_to_rotation
    from src.utils.image_loader import load_image

    sets = [
        ("simulator-data", "camera_poses.txt", 15, "yup", None),
        ("phone-data", "camera_poses_zyx.txt", 5, "zyx", "calibration_scaled.npz"),
        ("vo_dataset_salah", "camera_poses_zyx.txt", 5, "zyx", "calibration.npz"),
    ]
    only = sys.argv[1:]
    report = {}
    for name, gtf, step, conv, calib in sets:
        if only and name not in only:
            continue
        d = os.path.join(REF, "evaluation-runs", name, "data")
        gt = GroundTruthLoader(os.path.join(d, gtf))
        gt.load()
        frames = gt.get_frame_indices(step=step)
        img0 = load_image(os.path.join(d, "images", f"{frames[0]:06d}.png"), to_gray=True)
        if calib and os.path.exists(os.path.join(d, calib)):
            z = np.load(os.path.join(d, calib))
            K = np.asarray(z["K"] if "K" in z.files else z["intrinsic_matrix"], dtype=np.float64).reshape(3, 3)
        else:
            from src.core.camera_calibration import CameraCalibration
            K = CameraCalibration().get_matrix_from_image(img0)
        est = PoseEstimator(K, use_vp_refinement=True)
        # instrument the gates without touching the reference: wrap _estimate_manhattan_dirs / _optimize
        calls = []
        orig = est._estimate_manhattan_dirs

        def wrapped(gray, n_lat=90, n_lon=360, rng_seed=0, _o=orig):
            D, ok, dbg = _o(gray, n_lat=n_lat, n_lon=n_lon, rng_seed=rng_seed)
            calls.append((ok, dbg.get("acc_max", 0.0), dbg.get("vp2_score", 0.0), dbg.get("num_lines", 0)))
            return D, ok, dbg
        est._estimate_manhattan_dirs = wrapped
        opt_calls = []
        orig_opt = est._optimize_rotation_from_vps

        def wrapped_opt(R_init, Dc, Dw, _o=orig_opt):
            opt_calls.append(1)
            return _o(R_init, Dc, Dw)
        est._optimize_rotation_from_vps = wrapped_opt
        est_off = PoseEstimator(K, use_vp_refinement=False)
        n = gates = changed = failed = 0
        per_img_pass = 0
        t0 = time.time()
        acc_vals, vp2_vals = [], []
        for a, b in zip(frames[:-1], frames[1:]):
            i1 = load_image(os.path.join(d, "images", f"{a:06d}.png"), to_gray=True)
            i2 = load_image(os.path.join(d, "images", f"{b:06d}.png"), to_gray=True)
            p = gt.get_pose(a)
            R_prev = euler_to_rotation(p["yaw"], p["pitch"], p["roll"], convention=conv)
            n += 1
            k0, o0 = len(calls), len(opt_calls)
            try:
                R_on, _ = est.estimate(i1, i2, R_prev=R_prev)
                R_off, _ = est_off.estimate(i1, i2)
            except Exception:
                failed += 1
                continue
            for ok, am, v2, _nl in calls[k0:]:
                acc_vals.append(am); vp2_vals.append(v2)
                per_img_pass += bool(ok and am >= est.vp_acc_min and v2 >= est.vp_vp2_min)
            gates += len(opt_calls) > o0
            changed += not np.array_equal(R_on, R_off)
        report[name] = {"step": step, "pairs": n, "failed_pairs": failed, "gate_pass_pairs": gates,
                        "R_changed_pairs": changed, "images_passing_both_gates": per_img_pass,
                        "acc_max_max": float(max(acc_vals, default=0.0)), "acc_max_median": float(np.median(acc_vals)) if acc_vals else 0.0,
                        "vp2_score_max": float(max(vp2_vals, default=0.0)),
                        "gate_thresholds": {"vp_acc_min": est.vp_acc_min, "vp_vp2_min": est.vp_vp2_min},
                        "image_size": list(img0.shape[::-1]), "seconds": round(time.time() - t0, 1)}
        print(name, report[name], flush=True)
    prev = json.load(open(OUT)) if os.path.exists(OUT) else {}
    prev.update(report)
    json.dump(prev, open(OUT, "w"), indent=1)


if __name__ == "__main__":
    main()
