"""CPU checks of the drop-in boundary: librpe_b200.so builds / loads, exports every symbol
include/rpe.h declares, fails loudly without a B200 (no CPU fallback), and the host-side mirror of
the reference interface (constructor contract, Euler helpers) behaves like the reference."""
import ctypes as C
import json
import os
import re
import subprocess
import sys

import numpy as np
import pytest
from conftest import GOLDEN, ROOT

import rpe_b200
from rpe_b200 import _lib as L

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "rpe.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rpe_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = L.load()
    declared = _header_symbols()
    assert len(declared) >= 17
    assert sorted(L.EXPORTED_SYMBOLS) == declared
    for name in declared:
        assert getattr(lib, name) is not None
    nm = subprocess.run(["nm", "-D", "--defined-only", L.library_path()], capture_output=True, text=True).stdout
    for name in declared:
        assert re.search(rf"\bT {name}\b", nm), name
    assert lib.rpe_abi_version() == 1


def test_default_config_is_the_reference_configuration():
    c = L.RpeConfig()
    L.load().rpe_default_config(C.byref(c))
    # reference: src/core/pose_estimator.py:85-91 (ORB), :131 (crossCheck), :150 (500), :525-526 (RANSAC)
    assert (c.nfeatures, c.nlevels, c.fast_threshold, c.max_matches) == (4000, 12, 15, 500)
    assert np.float32(c.scale_factor) == np.float32(1.1)
    assert c.match_mode == L.MATCH_CROSSCHECK and c.ransac_max_iters == 1000
    assert (c.ransac_prob, c.ransac_threshold, c.ratio) == (0.999, 1.0, 0.75)


def test_sass_is_sm100a_only():
    out = subprocess.run(["cuobjdump", "-lelf", L.library_path()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_no_cpu_fallback(gpu_available):
    if gpu_available:
        pytest.skip("GPU present")
    lib = L.load()
    c = L.RpeConfig()
    lib.rpe_default_config(C.byref(c))
    h = C.c_void_p()
    assert lib.rpe_create(C.byref(c), 0, C.byref(h)) == -2          # RPE_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.rpe_last_error()
    with pytest.raises(L.RpeError, match="rpe_create"):
        rpe_b200.Engine()
    est = rpe_b200.PoseEstimator(np.eye(3))
    img = np.zeros((480, 640), np.uint8)
    with pytest.raises(L.RpeError):
        est.estimate(img, img)


def test_invalid_config_rejected():
    lib = L.load()
    c = L.RpeConfig()
    lib.rpe_default_config(C.byref(c))
    c.nlevels = 99
    h = C.c_void_p()
    assert lib.rpe_create(C.byref(c), 0, C.byref(h)) == -1 and not h.value
    assert lib.rpe_create(None, 0, C.byref(h)) == -1
    lib.rpe_destroy(None)                                            # no-op


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "relative-pose-estimation_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f
                # cv2 only inside BatchProcessor's load_image (PNG decode stays host work), imported lazily
                assert "import cv2" not in src or f == "batch_processor.py", f
    code = ("import sys; import rpe_b200; "
            "assert not any(m == 'oracle' or m.startswith('oracle.') or m == 'cv2' or m.startswith('bench_support') "
            "for m in sys.modules), [m for m in sys.modules if 'cv2' in m or 'oracle' in m]")
    subprocess.check_call([sys.executable, "-c", code], cwd=ROOT)


def test_constructor_contract():
    """reference: src/core/pose_estimator.py:19-32, :96, :129."""
    K = np.eye(3)
    with pytest.raises(ValueError, match="Unknown feature extraction method"):
        rpe_b200.PoseEstimator(K, feature_method="SURF")
    with pytest.raises(ValueError, match="Unknown norm type"):
        rpe_b200.PoseEstimator(K, norm_type="L1")
    with pytest.raises(NotImplementedError):
        rpe_b200.PoseEstimator(K, feature_method="SIFT")
    with pytest.raises(NotImplementedError):
        rpe_b200.PoseEstimator(K, norm_type="L2")
    e = rpe_b200.PoseEstimator(K, "orb", "hamming", 300, 2000, True)
    assert (e.max_matches, e.nfeatures, e.use_vp_refinement) == (300, 2000, True)
    with pytest.raises(ValueError, match="uint8 HxW grayscale"):
        e._gray(np.zeros((4, 4, 4), np.uint8))
    with pytest.raises(ValueError, match="uint8 HxW grayscale"):
        e._gray(np.zeros((4, 4), np.float32))
    assert e._gray(np.zeros((4, 4, 3), np.uint8)).shape == (4, 4, 3)      # BGR is converted on the GPU
    with pytest.raises(ValueError):
        rpe_b200.EngineConfig(match_mode="flann").to_c()


def test_euler_helpers_equal_reference():
    from rpe_b200 import geometry as g
    for e in G["known"]["euler"]:
        R = g.euler_to_rotation(*e["ypr_in"], e["convention"])
        assert np.allclose(R, e["R"], atol=1e-12)
        assert np.allclose(g.rotation_to_euler(np.array(e["R"]), e["convention"]), e["ypr_out"], atol=1e-10)


def test_vp_refinement_is_not_dropped_silently(monkeypatch):
    """not built (its gates pass on 0 of the reference's 147 dataset pairs, profiles/r02_vp_gate_count.json),
    but asking for it must be audible: one warning on the first call that would have entered the reference's branch."""
    import json
    import warnings
    from rpe_b200.pose_estimator import VPRefinementNotApplied
    rep = json.load(open(os.path.join(ROOT, "profiles", "r02_vp_gate_count.json")))
    assert set(rep) == {"simulator-data", "phone-data", "vo_dataset_salah"}
    assert sum(r["pairs"] for r in rep.values()) == 147
    assert all(r["gate_pass_pairs"] == 0 and r["R_changed_pairs"] == 0 and r["failed_pairs"] == 0 for r in rep.values())
    est = rpe_b200.PoseEstimator(np.eye(3), use_vp_refinement=True)
    monkeypatch.setattr(est, "estimate_batch", lambda a, b: (np.zeros((1, 3, 3)), np.zeros((1, 3, 1)), np.zeros(1, np.int32)))
    est.last = {"info": np.zeros((1, 4), np.int32)}
    img = np.zeros((8, 8), np.uint8)
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        est.estimate(img, img)                       # no R_prev: the reference skips the branch too
        assert not w
        est.estimate(img, img, R_prev=np.eye(3))
        est.estimate(img, img, R_prev=np.eye(3))     # once
        assert len(w) == 1 and issubclass(w[0].category, VPRefinementNotApplied)
    off = rpe_b200.PoseEstimator(np.eye(3), use_vp_refinement=False)
    monkeypatch.setattr(off, "estimate_batch", lambda a, b: (np.zeros((1, 3, 3)), np.zeros((1, 3, 1)), np.zeros(1, np.int32)))
    off.last = {"info": np.zeros((1, 4), np.int32)}
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        off.estimate(img, img, R_prev=np.eye(3))
        assert not w


def test_euler_yup_survives_rounding_above_one():
    """R_prev @ R_rel can leave |R[2,1]| = 1 + 1 ulp near gimbal lock: numpy semantics as the reference (nan + warning),
    never an exception that aborts BatchProcessor.process_sequence."""
    import warnings
    from rpe_b200 import geometry as g
    R = np.eye(3)
    R[2, 1] = 1.0 + 2.3e-16
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        yaw, pitch, roll = g.rotation_to_euler_yup(R)
    assert np.isnan(pitch) and yaw == 0.0
