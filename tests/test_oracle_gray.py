"""The gray-conversion oracle (oracle/gray.py) against live cv2, and the kernel's host-compiled arithmetic
(gray_from_bgr in orb_math.cuh) against the oracle.  CPU only."""
import numpy as np

import cv2
import _hostcheck as hc

from oracle.gray import bgr_to_gray


def _lattice():
    # every (b, g, r) with components on a 37-step lattice plus the extremes: catches any rounding drift
    v = np.unique(np.concatenate([np.arange(0, 256, 37), [1, 2, 127, 128, 254, 255]])).astype(np.uint8)
    b, g, r = np.meshgrid(v, v, v, indexing="ij")
    return np.stack([b.ravel(), g.ravel(), r.ravel()], axis=1).reshape(1, -1, 3)


def test_cvtcolor_mode_bit_exact():
    rng = np.random.default_rng(7)
    for img in (rng.integers(0, 256, (97, 131, 3), dtype=np.uint8), _lattice()):
        assert np.array_equal(bgr_to_gray(img, 0), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY))


def test_imread_grayscale_png_mode_bit_exact():
    rng = np.random.default_rng(8)
    for img in (rng.integers(0, 256, (97, 131, 3), dtype=np.uint8), _lattice()):
        ok, buf = cv2.imencode(".png", img)
        assert ok and np.array_equal(cv2.imdecode(buf, cv2.IMREAD_COLOR), img)
        assert np.array_equal(bgr_to_gray(img, 1), cv2.imdecode(buf, cv2.IMREAD_GRAYSCALE))


def test_the_two_modes_differ_as_the_survey_says():
    img = np.random.default_rng(9).integers(0, 256, (200, 200, 3), dtype=np.uint8)
    d = bgr_to_gray(img, 0).astype(int) - bgr_to_gray(img, 1).astype(int)
    assert set(np.unique(d)) <= {0, 1} and 0.2 < (d != 0).mean() < 0.7


def test_kernel_arithmetic_equals_oracle():
    img = np.random.default_rng(10).integers(0, 256, (64, 80, 3), dtype=np.uint8)
    for mode in (0, 1):
        for src in (img, _lattice()):
            assert np.array_equal(hc.gray(src, mode), bgr_to_gray(src, mode))
