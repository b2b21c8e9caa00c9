"""Drop-in for the reference's ``PoseEstimator`` (reference: src/core/pose_estimator.py:11-153,
487-688): same constructor signature, same ``estimate`` / ``estimate_with_debug`` results and
RuntimeErrors, computed by hand-written sm_100a kernels through librpe_b200.so instead of cv2 on
the CPU.  Adds ``estimate_batch`` (one launch sequence for many pairs) and stage accessors.

Not implemented (documented scope): SIFT / L2 raise NotImplementedError.
VP refinement (reference :155-481, applied at :535-567) is not built: run with the unmodified reference over
its three datasets at the runners' steps (tools/vp_gate_count.py -> profiles/r02_vp_gate_count.json) its two
reliability gates (:545-550) pass on 0 of 58 (simulator), 0 of 9 (phone) and 0 of 80 (salah) pairs and R_rel is
bit-identical with the refinement on and off on all 147.  The constructor accepts the ``vp_*`` arguments; the
first ``estimate*`` call that would have entered the branch (``use_vp_refinement`` and ``R_prev`` given) warns
once (``VPRefinementNotApplied``) instead of dropping it silently.
"""
from __future__ import annotations

import warnings

import numpy as np

from . import _lib as L
from .engine import Engine, EngineConfig


class VPRefinementNotApplied(UserWarning):
    """use_vp_refinement=True and R_prev were given: the reference would run its VP branch here."""

_MESSAGES = {
    L.STATUS_NO_DESCRIPTORS: "Could not compute descriptors for one of the images.",
    L.STATUS_E_FAILED: "Could not estimate Essential matrix.",
}


class PoseEstimator:
    def __init__(self, camera_matrix, feature_method="ORB", norm_type="Hamming", max_matches=500,
                 nfeatures=4000, use_vp_refinement=False, vp_max_lines=120, vp_max_pairs=3000,
                 vp_acc_min=8e5, vp_vp2_min=8000.0, vp_iters=12, vp_lm_lambda=1e-2,
                 vp_cost_improve_eps=1e-3, *, device=0, max_pairs=1, match_mode="crosscheck", ratio=0.75,
                 nlevels=12, ransac_max_iters=1000, euler_convention="yup", max_image_size=None):
        self.K = np.asarray(camera_matrix, dtype=np.float64).reshape(3, 3)
        self.feature_method = feature_method
        self.norm_type = norm_type
        self.max_matches = max_matches
        self.nfeatures = nfeatures
        self.use_vp_refinement = use_vp_refinement
        method = feature_method.upper()
        if method == "SIFT":
            raise NotImplementedError("SIFT is outside the B200 hot path (ORB only)")
        if method != "ORB":
            raise ValueError(f"Unknown feature extraction method: {method}")
        norm = norm_type.upper()
        if norm == "L2":
            raise NotImplementedError("L2 matching is outside the B200 hot path (Hamming only)")
        if norm != "HAMMING":
            raise ValueError(f"Unknown norm type: {norm}")
        self._device = device
        self._kw = dict(max_pairs=max_pairs, nfeatures=nfeatures, nlevels=nlevels, match_mode=match_mode,
                        ratio=ratio, max_matches=max_matches, ransac_max_iters=ransac_max_iters,
                        euler_convention=euler_convention)
        self._engine = None
        self._size = (0, 0)
        self._vp_warned = False
        if max_image_size is not None:
            self._ensure_engine(max_image_size[0], max_image_size[1], max_pairs)

    # engines are sized lazily from the first image seen (and re-created if a larger one arrives)
    def _ensure_engine(self, W, H, B):
        if self._engine is not None and W <= self._size[0] and H <= self._size[1] and B <= self._kw["max_pairs"]:
            return self._engine
        if self._engine is not None:
            self._engine.close()
        self._size = (max(W, self._size[0]), max(H, self._size[1]))
        self._kw["max_pairs"] = max(B, self._kw["max_pairs"])
        self._engine = Engine(EngineConfig(max_width=self._size[0], max_height=self._size[1], **self._kw),
                              device=self._device)
        return self._engine

    @property
    def engine(self) -> Engine:
        if self._engine is None:
            raise RuntimeError("engine not created yet (call estimate or pass max_image_size)")
        return self._engine

    @staticmethod
    def _gray(img):
        """HxW uint8 gray (the reference's input), or HxWx3 uint8 BGR as cv2.imread returns it: the batch and
        sequence calls then convert on the GPU with cv2.cvtColor's arithmetic (load_image, image_loader.py:28)."""
        a = np.asarray(img)
        if a.dtype != np.uint8 or not (a.ndim == 2 or (a.ndim == 3 and a.shape[2] == 3)):
            raise ValueError("images must be uint8 HxW grayscale (or HxWx3 BGR) arrays")
        return a

    # ------------------------------------------------------------------ reference API
    def estimate_batch(self, imgs1, imgs2):
        """-> (R (B,3,3) f64, t (B,3,1) f64, status (B,) i32).  Never raises per pair."""
        imgs1 = [self._gray(i) for i in imgs1]
        imgs2 = [self._gray(i) for i in imgs2]
        H, W = imgs1[0].shape[:2]
        out = self._ensure_engine(W, H, len(imgs1)).estimate_batch(imgs1, imgs2, self.K)
        self.last = out
        return out["R"], out["t"], out["status"]

    def estimate_sequence(self, frames):
        """Consecutive pairs (k, k+1) of a frame sequence, ORB once per frame (the reference runs it on
        both frames of every pair, src/core/pose_estimator.py:505-506, so interior frames twice);
        results are bit-identical to ``estimate_batch(frames[:-1], frames[1:])``.
        -> (R (F-1,3,3), t (F-1,3,1), status (F-1,))."""
        frames = [self._gray(f) for f in frames]
        H, W = frames[0].shape[:2]
        out = self._ensure_engine(W, H, len(frames) - 1).estimate_sequence(frames, self.K)
        self.last = out
        return out["R"], out["t"], out["status"]

    def _raise_for(self, status, num_matches):
        code = int(status) & 0xFF
        if code == L.STATUS_INSUFFICIENT_MATCHES:
            raise RuntimeError(f"Insufficient matches: {num_matches} (minimum 5 required)")
        if code != L.STATUS_OK:
            raise RuntimeError(_MESSAGES[code])

    def _note_vp(self, R_prev):
        """The reference enters its VP branch when ``use_vp_refinement and R_prev is not None`` (:536).  That branch
        is not built here (see the module docstring: its gates pass on none of the reference's 147 dataset pairs)."""
        if self.use_vp_refinement and R_prev is not None and not self._vp_warned:
            self._vp_warned = True
            warnings.warn("use_vp_refinement=True: VP refinement is not implemented in the B200 engine and is NOT "
                          "applied; results equal the reference's whenever its VP reliability gates fail (all 147 "
                          "pairs of the reference's own datasets, profiles/r02_vp_gate_count.json)",
                          VPRefinementNotApplied, stacklevel=3)

    def estimate(self, img1, img2, R_prev=None):
        """reference: src/core/pose_estimator.py:487-569 -> (R 3x3, t 3x1)."""
        self._note_vp(R_prev)
        R, t, st = self.estimate_batch([img1], [img2])
        self._raise_for(st[0], self.last["info"][0, 0])
        return R[0], t[0]

    def estimate_with_debug(self, img1, img2, R_prev=None):
        """reference: src/core/pose_estimator.py:571-688 (same dict keys)."""
        self._note_vp(R_prev)
        R, t, st = self.estimate_batch([img1], [img2])
        info = self.last["info"][0]
        self._raise_for(st[0], info[0])
        e = self.engine
        return {"R": R[0], "t": t[0], "num_matches": int(info[0]),
                "pts1": e.get_stage(0, 0, L.STAGE_PTS1, 0, np.float32).reshape(-1, 2).copy(),
                "pts2": e.get_stage(0, 0, L.STAGE_PTS2, 0, np.float32).reshape(-1, 2).copy(),
                "inliers": int(info[1]), "vp_used": False, "vp_debug": {}}

    # seams the reference keeps private, exposed as arrays for the parity harness
    def _detect_and_compute(self, image):
        image = self._gray(image)
        return self._ensure_engine(image.shape[1], image.shape[0], 1).orb_detect(image)

    def _match_descriptors(self, desc1, desc2):
        return self.engine.match(desc1, desc2)
