"""CPU restatement of the brute-force Hamming matcher stage (numpy).

TEST INFRASTRUCTURE ONLY (see oracle/orb.py header for the import rule).

Restates ``cv2.BFMatcher(NORM_HAMMING, crossCheck=True).match(d1, d2)`` followed by the
reference's stable sort and ``max_matches`` cut (reference: src/core/pose_estimator.py:131,
:144-151), plus ``knnMatch(k=2)`` + ratio test for BASELINE configs 3-5.
Tie rules (verified against cv2 4.13): lowest train index among equal minima in the forward
direction, lowest query index in the backward direction; knn = first two of a stable
ascending sort.
"""
from __future__ import annotations

import numpy as np


def hamming_matrix(d1: np.ndarray, d2: np.ndarray) -> np.ndarray:
    """(N1, N2) uint16 Hamming distances between 32-byte descriptors."""
    a = np.ascontiguousarray(d1).view(np.uint64).reshape(-1, 4)
    b = np.ascontiguousarray(d2).view(np.uint64).reshape(-1, 4)
    out = np.zeros((a.shape[0], b.shape[0]), dtype=np.uint16)
    step = max(1, (1 << 24) // max(1, b.shape[0]))
    for s in range(0, a.shape[0], step):
        x = a[s:s + step, None, :] ^ b[None, :, :]
        out[s:s + step] = np.bitwise_count(x).sum(axis=2).astype(np.uint16)
    return out


def match_crosscheck(d1: np.ndarray, d2: np.ndarray) -> np.ndarray:
    """(K, 3) int32 rows (queryIdx, trainIdx, distance), query-ascending, cross-checked."""
    if d1.shape[0] == 0 or d2.shape[0] == 0:
        return np.zeros((0, 3), np.int32)
    D = hamming_matrix(d1, d2)
    fwd = D.argmin(axis=1)                 # first minimum = lowest train index
    bwd = D.argmin(axis=0)                 # lowest query index
    q = np.arange(D.shape[0])
    keep = bwd[fwd] == q
    return np.stack([q[keep], fwd[keep], D[q[keep], fwd[keep]]], axis=1).astype(np.int32)


def knn2(d1: np.ndarray, d2: np.ndarray):
    """Top-2 per query: (idx (N1,2) int32, dist (N1,2) int32); needs N2 >= 2."""
    D = hamming_matrix(d1, d2).astype(np.int64)
    key = D * (1 << 32) + np.arange(D.shape[1])[None, :]      # stable (dist, index) order
    part = np.sort(key, axis=1)[:, :2]
    return (part & 0xFFFFFFFF).astype(np.int32), (part >> 32).astype(np.int32)


def match_ratio(d1: np.ndarray, d2: np.ndarray, ratio: float = 0.75) -> np.ndarray:
    """knnMatch(k=2) + Lowe ratio test d0 < ratio*d1 (strict); rows (q, t, dist), q ascending."""
    if d1.shape[0] == 0 or d2.shape[0] < 2:
        return np.zeros((0, 3), np.int32)
    idx, dist = knn2(d1, d2)
    keep = dist[:, 0].astype(np.float64) < ratio * dist[:, 1].astype(np.float64)
    q = np.arange(d1.shape[0])[keep]
    return np.stack([q, idx[keep, 0], dist[keep, 0]], axis=1).astype(np.int32)


def sort_and_cut(matches: np.ndarray, max_matches: int | None) -> np.ndarray:
    """The reference's ``sorted(matches, key=distance)[:max_matches]`` (stable)."""
    order = np.argsort(matches[:, 2], kind="stable")
    m = matches[order]
    if max_matches is not None:
        m = m[:max_matches]
    return m
