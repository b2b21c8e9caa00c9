"""Pins oracle/reference_path.py (the call-for-call cv2 restatement of PoseEstimator.estimate) against
the golden vectors produced by the UNMODIFIED reference (tests/golden/make_golden.py) and, where
/root/reference exists (build container only), against the reference class imported live."""
import json
import os
import sys
import types

import numpy as np
import pytest
from conftest import GOLDEN, load_gray

from oracle.reference_path import PathConfig, ReferencePath

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


@pytest.mark.parametrize("i", range(len(G["pairs"])))
def test_port_equals_golden_reference(i):
    g = G["pairs"][i]
    a, b, K = load_gray(g["img1"]), load_gray(g["img2"]), np.array(g["K"])
    d = ReferencePath(K, PathConfig()).estimate_with_debug(a, b)
    assert d["num_matches"] == g["reference"]["num_matches"] and d["inliers"] == g["reference"]["inliers"]
    assert np.abs(d["R"] - np.array(g["reference"]["R"])).max() < 1e-12
    assert np.abs(d["t"].ravel() - np.array(g["reference"]["t"])).max() < 1e-12
    c = ReferencePath(K, PathConfig(), canonical=True).estimate_with_debug(a, b)
    assert c["num_matches"] == g["canonical"]["num_matches"] and c["inliers"] == g["canonical"]["inliers"]
    assert np.abs(c["R"] - np.array(g["canonical"]["R"])).max() < 1e-12
    assert int(c["ransac_mask"].sum()) == g["canonical"]["ransac_inliers"]


def test_cfg1_golden_of_survey():
    """default single pair under cv2 4.13."""
    r = G["pairs"][0]["reference"]
    assert (r["num_matches"], r["inliers"]) == (500, 497)
    assert np.allclose(r["t"], [0.0623287603, 0.2878066596, -0.9556581252], atol=1e-9)
    assert np.allclose(r["euler_yup"], [-0.30540, 1.38952, -0.00392], atol=1e-4)


def test_failure_contract():
    K = np.array(G["pairs"][0]["K"])
    blank = np.full((480, 640), 128, np.uint8)
    with pytest.raises(RuntimeError, match="Could not compute descriptors"):
        ReferencePath(K).estimate(blank, load_gray("sim_000000.png"))


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference tree only exists in the build container")
def test_port_equals_live_reference():
    for name in ("plotly", "plotly.graph_objects"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, "/root/reference")
    try:
        from src.core.pose_estimator import PoseEstimator
    finally:
        sys.path.remove("/root/reference")
    g = G["pairs"][1]
    a, b, K = load_gray(g["img1"]), load_gray(g["img2"]), np.array(g["K"])
    R, t = PoseEstimator(K, "ORB", nfeatures=4000, use_vp_refinement=False).estimate(a, b)
    R2, t2 = ReferencePath(K, PathConfig()).estimate(a, b)
    assert np.array_equal(R, R2) and np.array_equal(t, t2)
