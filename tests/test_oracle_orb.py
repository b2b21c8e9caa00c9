"""Pins the ORB oracle (oracle/orb.py): against the committed golden vectors (made from cv2 as the
reference constructs it, tests/golden/make_golden.py) and against live cv2 primitives."""
import hashlib
import json
import os

import cv2
import numpy as np
import pytest
from conftest import GOLDEN, load_gray

from oracle import orb as oo
from oracle.orb_pattern import PATTERN

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_known_answers():
    k = G["known"]
    p = oo.OrbParams()
    assert [lv.quota for lv in oo.level_geometry(640, 480, p)] == k["quotas_4000_12"] == \
        [534, 485, 441, 401, 365, 331, 301, 274, 249, 226, 206, 187]
    assert [lv.quota for lv in oo.level_geometry(1280, 720, oo.OrbParams(nfeatures=5000))] == k["quotas_5000_12"]
    assert [lv.quota for lv in oo.level_geometry(3840, 2160, oo.OrbParams(20000, 1.1, 8))] == k["quotas_20000_8"] == \
        [3408, 3098, 2817, 2561, 2328, 2116, 1924, 1748]
    assert [[lv.w, lv.h] for lv in oo.level_geometry(640, 480, p)] == k["sizes_640x480"]
    assert k["sizes_640x480"][1] == [582, 436] and k["sizes_640x480"][-1] == [224, 168]
    assert oo.UMAX == k["umax"] == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    assert sha(PATTERN.astype(np.int8)) == k["pattern_sha256"] == \
        "2164181aea6ff9ac426ca512d5130d15e1f6e3cd47b1cbdd568bbe1e55d49023"
    assert PATTERN[:3].tolist() == [[8, -3, 9, 5], [4, 2, 7, -12], [-11, 9, -8, 2]]
    assert oo.BLUR_KERNEL.view(np.uint32).tolist()[:4] == [0x3d8fafb1, 0x3e06387e, 0x3e434a39, 0x3e5d4ae0]


# the 1080p fixtures take ~10 s each in numpy: two small + one wide image keep the CPU suite short
@pytest.mark.parametrize("name", ["sim_000000.png", "sim_000015.png", "phone_000005.png"])
def test_oracle_orb_equals_golden(name):
    g = G["images"][name]
    r = oo.orb_detect_and_compute(load_gray(name), oo.OrbParams())
    assert len(r) == g["n_keypoints"]
    assert np.bincount(r.octave, minlength=12).tolist() == g["per_level"]
    assert sha(np.stack([r.x, r.y])) == g["xy"]
    assert sha(r.size) == g["size"] and sha(r.octave) == g["octave"]
    assert sha(r.angle) == g["angle"] and sha(r.response) == g["response"]
    assert sha(r.desc) == g["desc"]


def test_oracle_orb_equals_live_cv2():
    img = load_gray("sim_000030.png")
    p = oo.OrbParams()
    kps, desc = cv2.ORB_create(nfeatures=4000, scaleFactor=1.1, nlevels=12, fastThreshold=15,
                               scoreType=cv2.ORB_HARRIS_SCORE).detectAndCompute(img, None)
    ref = oo.canonicalize_cv2(kps, desc, oo.level_geometry(img.shape[1], img.shape[0], p))
    r = oo.orb_detect_and_compute(img, p)
    assert len(r) == len(ref) == G["images"]["sim_000030.png"]["n_keypoints"]
    for f in ("x", "y", "size", "angle", "response", "octave", "xl", "yl"):
        assert np.array_equal(getattr(r, f), getattr(ref, f)), f
    assert np.array_equal(r.desc, ref.desc)


def test_stage_primitives_against_cv2():
    img = load_gray("phone_000000.png")
    # A.1 resize (chained twice)
    a = oo.resize_linear_exact(img, 771, 435)
    assert np.array_equal(a, cv2.resize(img, (771, 435), interpolation=cv2.INTER_LINEAR_EXACT))
    b = oo.resize_linear_exact(a, 701, 395)
    assert np.array_equal(b, cv2.resize(a, (701, 395), interpolation=cv2.INTER_LINEAR_EXACT))
    # A.2 FAST + NMS: positions, scores, raster order
    kps = cv2.FastFeatureDetector_create(threshold=15, nonmaxSuppression=True).detect(img, None)
    score = oo.fast_score_map(img, 15)
    ys, xs = np.nonzero(oo.nms_3x3(score))
    assert [(int(k.pt[0]), int(k.pt[1])) for k in kps] == list(zip(xs.tolist(), ys.tolist()))
    assert [int(k.response) for k in kps] == score[ys, xs].tolist()
    # A.8 blur == sepFilter2D with the f32 kernel
    k = oo.BLUR_KERNEL.reshape(-1, 1)
    assert np.array_equal(oo.blur7(img), cv2.sepFilter2D(img, -1, k, k, borderType=cv2.BORDER_REFLECT_101))
    # A.6 fastAtan2
    rng = np.random.default_rng(0)
    y = rng.integers(-40000, 40000, 20000).astype(np.float32)
    x = rng.integers(-40000, 40000, 20000).astype(np.float32)
    want = np.array([cv2.fastAtan2(float(b), float(a)) for a, b in zip(x[:4000], y[:4000])], np.float32)
    assert np.array_equal(oo.fast_atan2(y[:4000], x[:4000]), want)


def test_retain_best_keeps_ties():
    r = np.array([5, 3, 3, 3, 1, 9], np.float32)
    assert oo.retain_best(r, 3).tolist() == [True, True, True, True, False, True]   # ties at the cut survive
    assert oo.retain_best(r, 10).all() and not oo.retain_best(r, 0).any()


def test_degenerate_images():
    blank = np.full((200, 200), 77, np.uint8)
    assert len(oo.orb_detect_and_compute(blank, oo.OrbParams())) == 0
    tiny = np.random.default_rng(0).integers(0, 255, (40, 50), dtype=np.uint8)    # narrower than 2*31+1
    assert len(oo.orb_detect_and_compute(tiny, oo.OrbParams())) == 0
