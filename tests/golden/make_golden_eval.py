#!/usr/bin/env python
"""Generate tests/golden/eval_golden.npz from the UNMODIFIED reference ``PoseEvaluator``
(src/core/pose_evaluator.py, imported from /root/reference, which exists only in the build container).

    python tests/golden/make_golden_eval.py

Per case (yup / zyx, odd and even lengths): seeded inputs in the shape ``BatchProcessor`` produces
({frames, roll, pitch, yaw, R, t}) plus a ground-truth table served through the ``get_pose`` interface of the
reference's GroundTruthLoader (src/core/ground_truth_loader.py), and the reference's outputs:
evaluate_sequence -> the five error arrays, compute_summary_statistics -> the 25 statistics.
"""
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


class TableGT:
    """get_pose(frame) -> {x, y, z, roll, pitch, yaw}: the only GroundTruthLoader method evaluate_sequence uses."""
    def __init__(self, frames, pos, euler):
        self.rows = {int(f): (pos[i], euler[i]) for i, f in enumerate(frames)}

    def get_pose(self, frame_idx):
        p, e = self.rows[int(frame_idx)]
        return {"x": p[0], "y": p[1], "z": p[2], "yaw": e[0], "pitch": e[1], "roll": e[2]}


def make_case(seed, n, convention, geometry):
    rng = np.random.default_rng(seed)
    frames = [15 * (i + 1) for i in range(n)]
    gt_euler = np.c_[rng.uniform(-179, 179, n), rng.uniform(-80, 80, n), rng.uniform(-179, 179, n)]   # yaw, pitch, roll
    gt_pos = np.cumsum(rng.normal(0, 0.3, (n, 3)), axis=0)
    est_euler = gt_euler + rng.normal(0, 2.0, (n, 3))
    est_euler[::7, 0] += 360.0          # exercise the wrap
    est_euler[3::11, 2] -= 355.0
    est_R = np.stack([geometry.euler_to_rotation(*est_euler[i], convention=convention) for i in range(n)])
    if n > 9:
        est_R[5] = geometry.euler_to_rotation(*gt_euler[5], convention=convention)      # exact match: acos near 1
    t = np.diff(gt_pos, axis=0, prepend=gt_pos[:1]) + rng.normal(0, 0.05, (n, 3))
    t[0] = rng.normal(0, 1, 3)
    est_t = (t / np.linalg.norm(t, axis=1, keepdims=True)).reshape(n, 3, 1)
    if n > 9:
        est_t[9] *= -1.0                    # opposite direction: acos near -1
    return frames, gt_euler, gt_pos, est_euler, est_R, est_t


def main():
    for name in ("plotly", "plotly.graph_objects", "plotly.subplots"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, REF)
    from src.core.pose_evaluator import PoseEvaluator
    from src.utils import geometry
    out = {}
    for ci, (seed, n, conv) in enumerate([(11, 58, "yup"), (12, 201, "zyx"), (13, 2, "yup"), (14, 1, "zyx")]):
        frames, gt_euler, gt_pos, est_euler, est_R, est_t = make_case(seed, n, conv, geometry)
        ev = PoseEvaluator(TableGT(frames, gt_pos, gt_euler), euler_convention=conv)
        res = ev.evaluate_sequence({"frames": frames, "roll": est_euler[:, 2], "pitch": est_euler[:, 1],
                                    "yaw": est_euler[:, 0], "R": list(est_R), "t": list(est_t)})
        stats = ev.compute_summary_statistics(res)
        metrics = ("roll_error", "pitch_error", "yaw_error", "rotation_error", "translation_dir_error")
        p = f"c{ci}_"
        out[p + "convention"] = np.array(conv)
        out[p + "frames"] = np.array(frames)
        out[p + "gt_euler"] = gt_euler; out[p + "gt_pos"] = gt_pos; out[p + "est_euler"] = est_euler
        out[p + "est_R"] = est_R; out[p + "est_t"] = est_t
        out[p + "errors"] = np.stack([res[m] for m in metrics], axis=1)
        out[p + "summary"] = np.array([[stats[f"{m}_{s}"] for s in ("mean", "std", "median", "max", "min")] for m in metrics])
        if ci == 0:
            # what the shipped pipeline does next (src/pipeline.py:160, :163): the console summary and the comparison table
            import contextlib
            import io
            buf = io.StringIO()
            with contextlib.redirect_stdout(buf):
                ev.print_summary(res)
            df = ev.create_comparison_dataframe(res)
            with open(os.path.join(HERE, "eval_report_c0.txt"), "w") as f:
                f.write(buf.getvalue())
            with open(os.path.join(HERE, "eval_report_c0_columns.txt"), "w") as f:
                f.write(",".join(df.columns) + "\n")
    np.savez_compressed(os.path.join(HERE, "eval_golden.npz"), **out)
    print("wrote eval_golden.npz:", {k: v.shape for k, v in out.items() if k.endswith("errors")})


if __name__ == "__main__":
    main()
