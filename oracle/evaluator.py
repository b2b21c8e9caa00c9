"""Numpy restatement of the reference's sequence evaluation.

TEST INFRASTRUCTURE ONLY (see oracle/orb.py header for the import rule): the checker for
``rpe_evaluate_sequence`` / ``rpe_b200.PoseEvaluator``; never imported by the product.

Follows, line by line:
  * ``PoseEvaluator.evaluate_sequence``            reference src/core/pose_evaluator.py:32-134
  * ``PoseEvaluator._wrap_angle_error``            :187-198
  * ``PoseEvaluator.compute_summary_statistics``   :136-157
  * ``rotation_error``                             reference src/utils/geometry.py:128-149
  * ``translation_direction_error``                :152-174
  * ``euler_to_rotation_yup`` / ``_zyx``           :82-125 / :208-236
Pinned: ``tests/golden/eval_golden.npz`` holds inputs and the outputs of the UNMODIFIED reference class run in the
build container (``tests/golden/make_golden_eval.py``); ``tests/test_oracle_evaluator.py`` checks this file against it.
"""
from __future__ import annotations

import numpy as np

METRICS = ("roll_error", "pitch_error", "yaw_error", "rotation_error", "translation_dir_error")
STATS = ("mean", "std", "median", "max", "min")


def euler_to_rotation(yaw_deg, pitch_deg, roll_deg, convention="yup"):
    y, p, r = np.deg2rad(yaw_deg), np.deg2rad(pitch_deg), np.deg2rad(roll_deg)
    cy, sy, cp, sp, cr, sr = np.cos(y), np.sin(y), np.cos(p), np.sin(p), np.cos(r), np.sin(r)
    if convention == "zyx":                                   # geometry.py:208-236
        return np.array([[cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
                         [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
                         [-sp, cp * sr, cp * cr]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])     # geometry.py:82-125
    Rx = np.array([[1, 0, 0], [0, cp, -sp], [0, sp, cp]])
    Rz = np.array([[cr, -sr, 0], [sr, cr, 0], [0, 0, 1]])
    return Ry @ Rx @ Rz


def rotation_error(R_est, R_gt):
    c = np.clip((np.trace(R_est @ R_gt.T) - 1) / 2, -1.0, 1.0)
    return np.rad2deg(np.arccos(c))


def translation_direction_error(t_est, t_gt):
    with np.errstate(invalid="ignore", divide="ignore"):
        a = t_est.flatten() / np.linalg.norm(t_est)
        b = t_gt.flatten() / np.linalg.norm(t_gt)
    return np.rad2deg(np.arccos(np.clip(np.dot(a, b), -1.0, 1.0)))


def wrap_angle_error(e):
    return abs(((e + 180) % 360) - 180)


def evaluate_sequence(est_R, est_t, est_euler, gt_euler, gt_pos, convention="yup"):
    """est_euler / gt_euler rows are (yaw, pitch, roll) degrees.  Returns errors (n, 5) in METRICS order."""
    n = len(est_R)
    out = np.zeros((n, 5))
    prev = None
    for i in range(n):
        gy, gp, gr = gt_euler[i]
        out[i, 0] = wrap_angle_error(abs(est_euler[i][2] - gr))
        out[i, 1] = wrap_angle_error(abs(est_euler[i][1] - gp))
        out[i, 2] = wrap_angle_error(abs(est_euler[i][0] - gy))
        out[i, 3] = rotation_error(np.asarray(est_R[i]).reshape(3, 3), euler_to_rotation(gy, gp, gr, convention))
        pos = np.asarray(gt_pos[i], dtype=np.float64)
        out[i, 4] = translation_direction_error(np.asarray(est_t[i], dtype=np.float64), pos - prev) if prev is not None else 0.0
        prev = pos
    return out


def summary_statistics(errors):
    """(5, 5): per metric mean, std, median, max, min (pose_evaluator.py:148-155)."""
    e = np.asarray(errors, dtype=np.float64)
    return np.array([[np.mean(e[:, m]), np.std(e[:, m]), np.median(e[:, m]), np.max(e[:, m]), np.min(e[:, m])]
                     for m in range(5)])
