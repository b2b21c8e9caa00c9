"""Device twin of the reference's ``PoseEvaluator`` (reference: src/core/pose_evaluator.py:14-198), SURVEY 8 f4.

Same constructor, ``evaluate_sequence(estimated_results)`` and ``compute_summary_statistics(evaluation_results)``
with the same result dicts; the per-frame Python loop (:75-120: ``euler_to_rotation``, ``rotation_error``,
``translation_direction_error``, angle wrapping) and the 25 numpy reductions (:148-155) run as two CUDA kernels
behind ``rpe_evaluate_sequence`` (kernels_eval.cu).  Only the ground-truth lookup (``gt_loader.get_pose``, a
pandas row per frame) stays on the host.  There is no CPU fallback: an engine (GPU) is required.
"""
from __future__ import annotations

import numpy as np

from . import _lib as L
from .geometry import CONVENTION_YUP, CONVENTION_ZYX

_METRICS = ("roll_error", "pitch_error", "yaw_error", "rotation_error", "translation_dir_error")
_STATS = ("mean", "std", "median", "max", "min")


class PoseEvaluator:
    def __init__(self, ground_truth_loader, euler_convention=CONVENTION_YUP, engine=None):
        self.gt_loader = ground_truth_loader
        self.euler_convention = euler_convention
        self._engine = engine
        self._summary = None

    def _get_engine(self):
        if self._engine is None:
            from .engine import Engine, EngineConfig
            # a minimal handle: the evaluation kernels use none of the image workspaces
            self._engine = Engine(EngineConfig(max_width=64, max_height=64, max_pairs=1, nfeatures=64, nlevels=1))
        return self._engine

    def evaluate_sequence(self, estimated_results):
        frames = estimated_results["frames"]
        n = len(frames)
        gt = [self.gt_loader.get_pose(f) for f in frames]
        gt_euler = np.array([[g["yaw"], g["pitch"], g["roll"]] for g in gt], dtype=np.float64).reshape(n, 3)
        gt_pos = np.array([[g["x"], g["y"], g["z"]] for g in gt], dtype=np.float64).reshape(n, 3)
        est_roll = estimated_results["roll"]; est_pitch = estimated_results["pitch"]; est_yaw = estimated_results["yaw"]
        est_euler = np.stack([np.asarray(est_yaw, np.float64), np.asarray(est_pitch, np.float64),
                              np.asarray(est_roll, np.float64)], axis=1)
        est_R = np.asarray(estimated_results["R"], dtype=np.float64).reshape(n, 9)
        est_t = np.asarray(estimated_results["t"], dtype=np.float64).reshape(n, 3)
        conv = L.EULER_ZYX if self.euler_convention == CONVENTION_ZYX else L.EULER_YUP
        errors, summary = self._get_engine().evaluate_sequence(est_R, est_t, est_euler, gt_euler, gt_pos, conv)
        out = {"frames": frames}
        for k, name in enumerate(_METRICS):
            out[name] = errors[:, k].copy()
        out.update({"gt_roll": gt_euler[:, 2].copy(), "gt_pitch": gt_euler[:, 1].copy(), "gt_yaw": gt_euler[:, 0].copy(),
                    "est_roll": est_roll, "est_pitch": est_pitch, "est_yaw": est_yaw})
        self._summary = (out, summary)
        return out

    def compute_summary_statistics(self, evaluation_results):
        """reference: pose_evaluator.py:136-157.  The statistics of the dict returned by the last
        ``evaluate_sequence`` were reduced on the device in the same call; any other dict is re-reduced there."""
        if self._summary is not None and self._summary[0] is evaluation_results:
            summary = self._summary[1]
        else:
            n = len(evaluation_results["rotation_error"])
            errors = np.stack([np.asarray(evaluation_results[m], np.float64) for m in _METRICS], axis=1)
            summary = self._get_engine().summarize_errors(errors) if n else np.full((5, 5), np.nan)
        return {f"{m}_{s}": summary[i, j] for i, m in enumerate(_METRICS) for j, s in enumerate(_STATS)}

    # ------------------------------------------------------------------ host-side reporting (no arithmetic of the path)
    # The shipped pipeline calls these two right after evaluate_sequence (reference: src/pipeline.py:160, :163), so a
    # drop-in evaluator needs them; they only format what the device kernels returned.
    _FRAME_COLUMNS = ("gt_roll", "gt_pitch", "gt_yaw", "est_roll", "est_pitch", "est_yaw") + _METRICS

    def create_comparison_dataframe(self, evaluation_results):
        """reference: pose_evaluator.py:159-184 — one row per frame: frame, GT / estimated angles, the five errors."""
        import pandas as pd
        cols = {"frame": evaluation_results["frames"]}
        cols.update({c: evaluation_results[c] for c in self._FRAME_COLUMNS})
        return pd.DataFrame(cols)

    @staticmethod
    def _wrap_angle_error(error_deg):
        """reference: pose_evaluator.py:186-198 — |error| after wrapping to [-180, 180)."""
        return abs(((error_deg + 180) % 360) - 180)

    # (title, metric, statistics shown) in the order the reference prints them (pose_evaluator.py:214-238)
    _SUMMARY_LAYOUT = (("Rotation Errors (degrees):", "rotation_error", ("mean", "std", "median", "max", "min")),
                       ("Roll Errors (degrees):", "roll_error", ("mean", "std")),
                       ("Pitch Errors (degrees):", "pitch_error", ("mean", "std")),
                       ("Yaw Errors (degrees):", "yaw_error", ("mean", "std")),
                       ("Translation Direction Errors (degrees):", "translation_dir_error", ("mean", "std")))

    def print_summary(self, evaluation_results):
        """reference: pose_evaluator.py:200-240 — same console text, from the device-reduced statistics."""
        stats = self.compute_summary_statistics(evaluation_results)
        bar = "=" * 60
        lines = ["\n" + bar, "POSE ESTIMATION EVALUATION SUMMARY", bar,
                 f"\nNumber of frames evaluated: {len(evaluation_results['frames'])}"]
        for title, metric, shown in self._SUMMARY_LAYOUT:
            lines.append("\n" + title)
            for s in shown:
                lines.append(f"  {(s.capitalize() + ':'):<7} {stats[f'{metric}_{s}']:.2f}")
        lines.append("\n" + bar + "\n")
        print("\n".join(lines))
