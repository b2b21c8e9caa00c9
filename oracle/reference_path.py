"""The reference's hot path, restated call-for-call on its CPU library (cv2).

TEST INFRASTRUCTURE ONLY (see oracle/orb.py header for the import rule).  This is the
``cpu_baseline`` / ``bench.py --impl reference`` arm and the end-to-end checker.

The reference (``PoseEstimator.estimate``, src/core/pose_estimator.py:487-569) is twenty
lines of Python around four OpenCV calls; /root/reference cannot travel to the GPU box, so
this module repeats exactly that call sequence with the same arguments:

* ORB_create(nfeatures, 1.1, 12, fastThreshold=15, HARRIS)       (pose_estimator.py:85-91)
* detectAndCompute(image, None) on both images                   (:108, :505-506)
* BFMatcher(NORM_HAMMING, crossCheck=True).match -> stable sort -> [:max_matches]  (:131-151)
  (or knnMatch(k=2) + ratio for BASELINE configs 3-5)
* float32 point gather                                           (:518-519)
* findEssentialMat(pts1, pts2, K, RANSAC, 0.999, 1.0[, maxIters]) (:522-527)
* recoverPose(E, pts1, pts2, K)                                  (:533)
and raises the same RuntimeErrors (:509, :515, :530).  VP refinement (:535-567) is out of
scope and off.  ``tests/test_oracle_reference.py`` checks this port against
the unmodified reference class imported from /root/reference (in the build container) and
against the committed golden vectors made by ``tests/golden/make_golden.py``.

``canonical=True`` feeds the matcher with keypoints in the canonical
(octave, y_level, x_level) order the CUDA engine emits: the same cv2
calls on the same features, only reordered, which is what stage-wise parity is judged on.
"""
from __future__ import annotations

from dataclasses import dataclass

import cv2
import numpy as np

from .orb import OrbParams, canonicalize_cv2, level_geometry


@dataclass
class PathConfig:
    nfeatures: int = 4000
    nlevels: int = 12
    scale_factor: float = 1.1
    fast_threshold: int = 15
    match_mode: str = "crosscheck"      # or "ratio"
    ratio: float = 0.75
    max_matches: int | None = 500
    ransac_prob: float = 0.999
    ransac_threshold: float = 1.0
    ransac_max_iters: int = 1000


class ReferencePath:
    def __init__(self, K, cfg: PathConfig | None = None, canonical: bool = False):
        self.K = np.asarray(K, dtype=np.float64)
        self.cfg = cfg or PathConfig()
        self.canonical = canonical
        c = self.cfg
        self.extractor = cv2.ORB_create(nfeatures=c.nfeatures, scaleFactor=c.scale_factor,
                                        nlevels=c.nlevels, fastThreshold=c.fast_threshold,
                                        scoreType=cv2.ORB_HARRIS_SCORE)
        self.matcher = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=(c.match_mode == "crosscheck"))

    # -- stages -------------------------------------------------------------------------
    def detect(self, img):
        kps, desc = self.extractor.detectAndCompute(img, None)
        if desc is None:
            return np.zeros((0, 2), np.float32), None
        if self.canonical:
            c = self.cfg
            lv = level_geometry(img.shape[1], img.shape[0],
                                OrbParams(c.nfeatures, c.scale_factor, c.nlevels, c.fast_threshold))
            r = canonicalize_cv2(kps, desc, lv)
            return np.stack([r.x, r.y], axis=1), r.desc
        return np.array([k.pt for k in kps], dtype=np.float32).reshape(-1, 2), desc

    def match(self, d1, d2):
        """-> (M, 3) int32 (queryIdx, trainIdx, distance) after the stable sort and cut."""
        c = self.cfg
        if c.match_mode == "crosscheck":
            ms = self.matcher.match(d1, d2)
        else:
            ms = []
            for pair in self.matcher.knnMatch(d1, d2, k=2):
                if len(pair) == 2 and pair[0].distance < c.ratio * pair[1].distance:
                    ms.append(pair[0])
        ms = sorted(ms, key=lambda m: m.distance)
        if c.max_matches is not None:
            ms = ms[:c.max_matches]
        return np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], dtype=np.int32).reshape(-1, 3)

    def essential(self, pts1, pts2):
        c = self.cfg
        return cv2.findEssentialMat(pts1, pts2, self.K, method=cv2.RANSAC, prob=c.ransac_prob,
                                    threshold=c.ransac_threshold, maxIters=c.ransac_max_iters)

    # -- the path -----------------------------------------------------------------------
    def estimate_with_debug(self, img1, img2):
        p1, d1 = self.detect(img1)
        p2, d2 = self.detect(img2)
        if d1 is None or d2 is None:
            raise RuntimeError("Could not compute descriptors for one of the images.")
        m = self.match(d1, d2)
        if len(m) < 5:
            raise RuntimeError(f"Insufficient matches: {len(m)} (minimum 5 required)")
        pts1 = p1[m[:, 0]].astype(np.float32)
        pts2 = p2[m[:, 1]].astype(np.float32)
        E, mask = self.essential(pts1, pts2)
        if E is None:
            raise RuntimeError("Could not estimate Essential matrix.")
        n, R, t, _ = cv2.recoverPose(E, pts1, pts2, self.K)
        return {"R": R, "t": t, "num_matches": len(m), "pts1": pts1, "pts2": pts2, "inliers": n,
                "E": E, "ransac_mask": mask.ravel(), "matches": m}

    def estimate(self, img1, img2):
        d = self.estimate_with_debug(img1, img2)
        return d["R"], d["t"]
