"""Deterministic synthetic image pairs for BASELINE configs 3-5.

BENCH / TEST INPUT GENERATOR (host-side; not on the hot path; uses cv2 on the host for the texture upsampling
and the homography warp).  The reference has no synthetic data; this is the generator the
survey defines so that the CPU baseline and the GPU engine see byte-identical inputs.

Per pair: seed = 1000*cfg + pair_idx.  View 1 = multi-octave noise texture scaled to [40, 200]
plus floor(W*H/1500) filled axis-aligned rectangles; view 2 = view 1 warped by the homography
induced by a small random motion (rvec ~ U(-0.03,0.03)^3, t ~ U(-0.1,0.1)^3) of a camera
looking at the plane z = 5; K = [[0.9W,0,W/2],[0,0.9W,H/2],[0,0,1]].

Stream (config 5): ONE base texture per stream (seed = 1000*cfg + stream_idx) seen by a camera whose pose
is a smooth bounded function of the frame index (one sinusoid per axis, random phase); frame k =
the base texture warped by the plane homography of pose k, so any frame can be generated independently
(shards need no state).  The motion between CONSECUTIVE frames is sized like a config-3 pair's
(|dt| ~ 0.03-0.2 per axis at scene depth 5, |dr| ~ 0.01-0.04 rad per axis: periods of 8-20 frames), so that
stride-1 pairs have a determinate pose.  (Round 1 used periods of 40-120 frames: a per-frame baseline of ~1/400 of
the depth, for which recoverPose's answer - cv2's too - is rounding noise)
"""
from __future__ import annotations

import cv2
import numpy as np


def synth_K(W: int, H: int) -> np.ndarray:
    return np.array([[0.9 * W, 0, W / 2.0], [0, 0.9 * W, H / 2.0], [0, 0, 1.0]])


def synth_texture(rng: np.random.Generator, W: int, H: int) -> np.ndarray:
    acc = np.zeros((H, W), np.float32)
    for o in range(6):
        h, w = H // (1 << o) + 2, W // (1 << o) + 2
        layer = rng.standard_normal((h, w)).astype(np.float32)
        acc += (2.0 ** (0.7 * o)) * cv2.resize(layer, (W, H), interpolation=cv2.INTER_CUBIC)
    lo, hi = float(acc.min()), float(acc.max())
    img = (40.0 + (acc - lo) * (160.0 / max(hi - lo, 1e-9))).astype(np.uint8)
    for _ in range((W * H) // 1500):
        x, y = int(rng.integers(0, W)), int(rng.integers(0, H))
        s1, s2 = int(rng.integers(6, 60)), int(rng.integers(6, 60))
        img[y:y + s2, x:x + s1] = int(rng.integers(0, 256))
    return img


def synth_motion(rng: np.random.Generator):
    rvec = rng.uniform(-0.03, 0.03, 3)
    t = rng.uniform(-0.1, 0.1, 3)
    R, _ = cv2.Rodrigues(rvec)
    return R, t


def synth_pair(cfg: int, pair_idx: int, W: int, H: int):
    """-> (view1 u8 HxW, view2 u8 HxW, K, R_gt, t_gt_unit)."""
    rng = np.random.default_rng(1000 * cfg + pair_idx)
    img1 = synth_texture(rng, W, H)
    R, t = synth_motion(rng)
    K = synth_K(W, H)
    n = np.array([0.0, 0.0, 1.0])
    Hm = K @ (R + np.outer(t, n) / 5.0) @ np.linalg.inv(K)
    img2 = cv2.warpPerspective(img1, Hm, (W, H), flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    return img1, img2, K, R, t / np.linalg.norm(t)


def synth_batch(cfg: int, first_pair: int, n_pairs: int, W: int, H: int, out: np.ndarray | None = None):
    """(2*n_pairs, H, W) uint8, view 1 of pair i at 2i, view 2 at 2i+1."""
    if out is None:
        out = np.empty((2 * n_pairs, H, W), np.uint8)
    for i in range(n_pairs):
        a, b, _, _, _ = synth_pair(cfg, first_pair + i, W, H)
        out[2 * i] = a
        out[2 * i + 1] = b
    return out


_STREAM_CACHE: dict = {}


def _stream_state(cfg: int, stream_idx: int, W: int, H: int):
    key = (cfg, stream_idx, W, H)
    if key not in _STREAM_CACHE:
        rng = np.random.default_rng(1000 * cfg + stream_idx)
        base = synth_texture(rng, W, H)
        # per axis: amplitude, angular frequency (periods of 8 .. 20 frames), phase
        par = {"r_amp": rng.uniform(0.02, 0.05, 3), "r_w": 2 * np.pi / rng.uniform(8, 20, 3),
               "r_ph": rng.uniform(0, 2 * np.pi, 3), "t_amp": rng.uniform(0.5, 0.9, 3),
               "t_w": 2 * np.pi / rng.uniform(8, 20, 3), "t_ph": rng.uniform(0, 2 * np.pi, 3)}
        _STREAM_CACHE.clear()
        _STREAM_CACHE[key] = (base, par)
    return _STREAM_CACHE[key]


def stream_pose(cfg: int, stream_idx: int, k: int, W: int, H: int):
    """Camera pose (R_k, t_k) of frame k relative to the base view."""
    _, p = _stream_state(cfg, stream_idx, W, H)
    rvec = p["r_amp"] * np.sin(p["r_w"] * k + p["r_ph"])
    t = p["t_amp"] * np.sin(p["t_w"] * k + p["t_ph"])
    R, _ = cv2.Rodrigues(rvec)
    return R, t


def stream_relative_pose(cfg: int, stream_idx: int, k: int, W: int, H: int, step: int = 1):
    """Ground-truth relative pose of the pair (frame k, frame k+step): R_rel, unit t_rel."""
    R0, t0 = stream_pose(cfg, stream_idx, k, W, H)
    R1, t1 = stream_pose(cfg, stream_idx, k + step, W, H)
    Rr = R1 @ R0.T
    tr = t1 - Rr @ t0
    return Rr, tr / max(np.linalg.norm(tr), 1e-300)


def synth_stream(cfg: int, first_frame: int, n_frames: int, W: int, H: int, stream_idx: int = 0,
                 out: np.ndarray | None = None, step: int = 1):
    """(n_frames, H, W) uint8: frames first_frame, first_frame + step, ... of stream ``stream_idx``
    (``step`` > 1 = the stream sampled at a lower frame rate: proportionally larger motion per pair)."""
    base, _ = _stream_state(cfg, stream_idx, W, H)
    K = synth_K(W, H)
    Ki = np.linalg.inv(K)
    n = np.array([0.0, 0.0, 1.0])
    if out is None:
        out = np.empty((n_frames, H, W), np.uint8)
    for i in range(n_frames):
        R, t = stream_pose(cfg, stream_idx, first_frame + i * step, W, H)
        Hm = K @ (R + np.outer(t, n) / 5.0) @ Ki
        cv2.warpPerspective(base, Hm, (W, H), dst=out[i], flags=cv2.INTER_LINEAR, borderMode=cv2.BORDER_REFLECT_101)
    return out
