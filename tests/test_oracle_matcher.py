"""Pins the matcher oracle (oracle/matcher.py) against live cv2.BFMatcher and the golden match lists."""
import hashlib
import json
import os

import cv2
import numpy as np
from conftest import GOLDEN, load_gray

from oracle import matcher as om
from oracle import orb as oo

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


def _cv_crosscheck(d1, d2, cut):
    ms = sorted(cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(d1, d2), key=lambda m: m.distance)[:cut]
    return np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], np.int32).reshape(-1, 3)


def _cv_ratio(d1, d2, ratio):
    out = []
    for p in cv2.BFMatcher(cv2.NORM_HAMMING).knnMatch(d1, d2, k=2):
        if len(p) == 2 and p[0].distance < ratio * p[1].distance:
            out.append([p[0].queryIdx, p[0].trainIdx, int(p[0].distance)])
    return np.array(out, np.int32).reshape(-1, 3)


def test_real_descriptors_vs_cv2_and_golden():
    r1 = oo.orb_detect_and_compute(load_gray("sim_000000.png"))
    r2 = oo.orb_detect_and_compute(load_gray("sim_000015.png"))
    m = om.sort_and_cut(om.match_crosscheck(r1.desc, r2.desc), 500)
    assert np.array_equal(m, _cv_crosscheck(r1.desc, r2.desc, 500))
    g = G["pairs"][0]["canonical"]
    assert hashlib.sha256(m.tobytes()).hexdigest() == g["matches"]
    assert m[:8].tolist() == g["first_matches"]
    full = om.match_crosscheck(r1.desc, r2.desc)
    # (SURVEY 3.1 quotes 1300 in cv2's own keypoint order; ties make the count order-dependent)
    assert len(full) == len(_cv_crosscheck(r1.desc, r2.desc, None)) == 1299
    ratio = om.sort_and_cut(om.match_ratio(r1.desc, r2.desc, 0.75), None)
    want = _cv_ratio(r1.desc, r2.desc, 0.75)
    assert np.array_equal(ratio, want[np.argsort(want[:, 2], kind="stable")])


def test_ties_lowest_index_both_ways():
    rng = np.random.default_rng(3)
    d = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    d2 = np.concatenate([d[:60], d[:60], d[60:]])              # exact duplicates -> distance-0 ties
    got = om.match_crosscheck(d, d2)
    assert np.array_equal(om.sort_and_cut(got, None), _cv_crosscheck(d, d2, None))
    assert (got[:60, 1] == np.arange(60)).all()
    # low-entropy descriptors: many equal minima at non-zero distance
    lo = (rng.integers(0, 2, (300, 32), dtype=np.uint8) * 0x0F).astype(np.uint8)
    lo2 = (rng.integers(0, 2, (280, 32), dtype=np.uint8) * 0x0F).astype(np.uint8)
    assert np.array_equal(om.sort_and_cut(om.match_crosscheck(lo, lo2), None), _cv_crosscheck(lo, lo2, None))
    r = om.match_ratio(lo, lo2, 0.75)
    w = _cv_ratio(lo, lo2, 0.75)
    assert np.array_equal(r, w)


def test_edge_cases():
    d = np.random.default_rng(0).integers(0, 256, (10, 32), dtype=np.uint8)
    assert om.match_crosscheck(d[:0], d).shape == (0, 3) and om.match_crosscheck(d, d[:0]).shape == (0, 3)
    assert om.match_ratio(d, d[:1]).shape == (0, 3)            # knn needs two train descriptors
    one = om.match_crosscheck(d[:1], d[4:5])
    assert one.tolist() == [[0, 0, int(np.unpackbits(d[0] ^ d[4]).sum())]]
    m = np.array([[0, 5, 9], [1, 4, 3], [2, 3, 9], [3, 2, 3]], np.int32)
    assert om.sort_and_cut(m, 3)[:, 0].tolist() == [1, 3, 0]   # stable: query order inside a distance
    assert om.hamming_matrix(d, d).diagonal().sum() == 0
