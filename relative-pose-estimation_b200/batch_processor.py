"""Batched twin of the reference's ``BatchProcessor`` (reference: src/core/batch_processor.py:14-141).

Same constructor, same ``process_sequence`` / ``process_at_interval`` / ``get_image_path``
and the same result dict ``{frames, roll, pitch, yaw, R, t}``; the per-pair Python loop
(:71-108) becomes ONE ``estimate_sequence`` call per chunk of pairs: every frame is decoded once
instead of twice (:79) and goes through ORB once instead of twice (pose_estimator.py:505-506).  World-frame composition ``R_prev @ R_rel`` and the Euler
conversion (:88-101) stay on the host (9 doubles per pair).  A pair whose status is not OK
raises the reference's RuntimeError, so a failing pair aborts the sequence as documented at
:55-56.  PNG decode stays on the CPU (out of the hot path).
"""
from __future__ import annotations

from pathlib import Path

import numpy as np

from .geometry import CONVENTION_YUP, euler_to_rotation, rotation_to_euler


def load_image(path, to_gray=True):
    """reference: src/utils/image_loader.py:9-30 (imread + BGR2GRAY; decode is host work)."""
    import cv2
    img = cv2.imread(str(path))
    if img is None:
        raise FileNotFoundError(f"Could not read image from: {path}")
    return cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) if to_gray else img


class BatchProcessor:
    def __init__(self, images_dir, pose_estimator, ground_truth_loader, euler_convention=CONVENTION_YUP,
                 chunk_pairs=64, gpu_gray=True):
        self.images_dir = Path(images_dir)
        self.pose_estimator = pose_estimator
        self.gt_loader = ground_truth_loader
        self.euler_convention = euler_convention
        self.chunk_pairs = int(chunk_pairs)
        # True: frames go to the GPU as decoded (BGR) and load_image's cvtColor (image_loader.py:28) runs there,
        # bit-identical (k_bgr2gray); False: cv2.cvtColor on the host
        self.gpu_gray = bool(gpu_gray)

    def get_image_path(self, frame_idx):
        return self.images_dir / f"{frame_idx:06d}.png"

    def process_sequence(self, frame_indices):
        if len(frame_indices) < 2:
            raise ValueError("Need at least 2 frames to process")
        frame_indices = list(frame_indices)
        results = {"frames": [], "roll": [], "pitch": [], "yaw": [], "R": [], "t": []}
        n_pairs = len(frame_indices) - 1
        for c0 in range(0, n_pairs, self.chunk_pairs):
            c1 = min(c0 + self.chunk_pairs, n_pairs)
            frames = [load_image(self.get_image_path(f), to_gray=not self.gpu_gray) for f in frame_indices[c0:c1 + 1]]
            # the reference hands R_prev to estimate() on every pair (:82-92), which arms its VP branch
            note = getattr(self.pose_estimator, "_note_vp", None)
            if note is not None:
                note(True)
            R_rel, t_rel, status = self.pose_estimator.estimate_sequence(frames)
            for k in range(c1 - c0):
                self.pose_estimator._raise_for(status[k], self.pose_estimator.last["info"][k, 0])
                gt = self.gt_loader.get_pose(frame_indices[c0 + k])
                R_prev = euler_to_rotation(gt["yaw"], gt["pitch"], gt["roll"], convention=self.euler_convention)
                R_new = R_prev @ R_rel[k]
                yaw, pitch, roll = rotation_to_euler(R_new, convention=self.euler_convention)
                results["frames"].append(frame_indices[c0 + k + 1])
                results["roll"].append(roll)
                results["pitch"].append(pitch)
                results["yaw"].append(yaw)
                results["R"].append(R_new)
                results["t"].append(t_rel[k])
        for k in ("roll", "pitch", "yaw"):
            results[k] = np.array(results[k])
        return results

    def process_at_interval(self, step=15):
        return self.process_sequence(self.gt_loader.get_frame_indices(step=step))
