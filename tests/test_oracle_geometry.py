"""Pins the geometry oracle (oracle/geometry.py): OpenCV's RANSAC driver hypothesis-for-hypothesis,
the five-point solver, recoverPose and the reference's Euler helpers (golden.json)."""
import json
import os

import cv2
import numpy as np
from conftest import GOLDEN, load_gray

from oracle import geometry as og
from oracle import matcher as om
from oracle import orb as oo

G = json.load(open(os.path.join(GOLDEN, "golden.json")))


def _points():
    r1 = oo.orb_detect_and_compute(load_gray("sim_000000.png"))
    r2 = oo.orb_detect_and_compute(load_gray("sim_000015.png"))
    m = om.sort_and_cut(om.match_crosscheck(r1.desc, r2.desc), 500)
    p1 = np.stack([r1.x[m[:, 0]], r1.y[m[:, 0]]], 1).astype(np.float32)
    p2 = np.stack([r2.x[m[:, 1]], r2.y[m[:, 1]]], 1).astype(np.float32)
    return p1, p2, np.array(G["pairs"][0]["K"])


def _cv_minimal(x1, x2):
    """cv2 with exactly 5 points returns every root of its minimal solver stacked."""
    E, _ = cv2.findEssentialMat(x1.astype(np.float64), x2.astype(np.float64), np.eye(3), method=cv2.RANSAC)
    return [] if E is None else [E[3 * i:3 * i + 3] for i in range(E.shape[0] // 3)]


def test_rng_subsets_known_answer():
    rng = og.CvRNG()
    got = [og.draw_subset(rng, 500) for _ in range(8)]
    assert got == G["known"]["rng_subsets_500"]
    assert got[:3] == [[105, 4, 440, 173, 331], [12, 301, 280, 73, 110], [485, 494, 46, 149, 418]]


def test_update_num_iters():
    assert og.ransac_update_num_iters(0.999, 0.0, 5, 1000) == 0
    assert og.ransac_update_num_iters(0.999, 1.0, 5, 1000) == 1000
    assert og.ransac_update_num_iters(0.999, 76 / 500, 5, 1000) == 12      # the cfg1 golden stop (424 inliers)
    assert og.ransac_update_num_iters(0.999, 0.5, 5, 1000) == 218


def test_ransac_driver_identical_to_cv2_with_cv2_solver():
    p1, p2, K = _points()
    E_cv, mask_cv = cv2.findEssentialMat(p1, p2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    tr = []
    E, mask = og.find_essential_ransac(p1, p2, K, solver=_cv_minimal, trace=tr)
    assert np.array_equal(mask, mask_cv.ravel()) and int(mask.sum()) == G["pairs"][0]["canonical"]["ransac_inliers"]
    assert np.abs(E - E_cv).max() < 1e-9
    assert np.abs(E_cv - np.array(G["pairs"][0]["canonical"]["E"])).max() < 1e-9
    assert len(tr) == 12


def test_numpy_five_point_and_full_ransac():
    p1, p2, K = _points()
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    rng = og.CvRNG()
    for _ in range(6):
        s = og.draw_subset(rng, len(x1))
        mine, theirs = og.five_point_models(x1[s], x2[s]), _cv_minimal(x1[s], x2[s])
        assert len(mine) == len(theirs)
        for E in mine:
            assert abs(np.linalg.norm(E) - 1) < 1e-9
            assert min(min(np.abs(E - T).max(), np.abs(E + T).max()) for T in theirs) < 1e-4   # roots of an ill-conditioned degree-10 polynomial
            assert np.abs(np.einsum("ni,ij,nj->n", np.c_[x2[s], np.ones(5)], E, np.c_[x1[s], np.ones(5)])).max() < 1e-9
    E, mask = og.find_essential_ransac(p1, p2, K)
    _, mask_cv = cv2.findEssentialMat(p1, p2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    m, c = mask.astype(bool), mask_cv.ravel().astype(bool)
    assert (m & c).sum() / (m | c).sum() >= 0.95


def test_recover_pose_vs_cv2_and_golden():
    p1, p2, K = _points()
    g = G["pairs"][0]["canonical"]
    E = np.array(g["E"])
    n_cv, R_cv, t_cv, mask_cv = cv2.recoverPose(E, p1, p2, K)
    n, R, t, mask = og.recover_pose(E, p1, p2, K)
    assert n == n_cv == g["inliers"] and np.array_equal(mask, mask_cv.ravel())
    assert np.abs(R - R_cv).max() < 1e-12 and np.abs(t - t_cv).max() < 1e-12
    assert np.abs(R - np.array(g["R"])).max() < 1e-9 and np.abs(t.ravel() - np.array(g["t"])).max() < 1e-9


def test_degenerate_inputs():
    K = np.array(G["pairs"][0]["K"])
    pts = np.random.default_rng(0).uniform(0, 400, (4, 2)).astype(np.float32)
    E, mask = og.find_essential_ransac(pts, pts + 1, K)
    assert E is None and mask.shape == (4,)


def test_euler_equals_reference_helpers():
    for e in G["known"]["euler"]:
        got = og.rotation_to_euler(np.array(e["R"]), e["convention"])
        assert np.allclose(got, e["ypr_out"], atol=1e-10), e
    for p in G["pairs"]:
        R = np.array(p["reference"]["R"])
        assert np.allclose(og.rotation_to_euler_yup(R), p["reference"]["euler_yup"], atol=1e-10)
        assert np.allclose(og.rotation_to_euler_zyx(R), p["reference"]["euler_zyx"], atol=1e-10)


def test_static_pair_convention_matches_cv2():
    """Identical point sets (a camera that did not move: the last five pairs of the reference's simulator sequence).
    The five-point system is rank deficient; cv2 4.13 returns E = [[0,-1,-1],[1,0,-1],[1,1,0]] / sqrt(6) for any such
    input, whatever the points, their number and K, with every point an inlier.  The oracle
    (and the CUDA solver, geom_math.cuh step 0) reproduce exactly that."""
    import cv2
    from oracle import geometry as og
    rng = np.random.default_rng(0)
    E0 = np.array([[0, -1, -1], [1, 0, -1], [1, 1, 0]]) / np.sqrt(6.0)
    for _ in range(6):
        M = int(rng.integers(6, 600))
        K = np.array([[rng.uniform(300, 1200), 0, rng.uniform(200, 700)], [0, rng.uniform(300, 1200), rng.uniform(150, 500)],
                      [0, 0, 1.0]])
        p = np.c_[rng.uniform(0, 1280, M), rng.uniform(0, 720, M)].astype(np.float32)
        E, mask = cv2.findEssentialMat(p, p.copy(), K, method=cv2.RANSAC, prob=0.999, threshold=1.0, maxIters=1000)
        assert E.shape == (3, 3) and np.abs(E - E0).max() < 1e-12 and int(mask.sum()) == M
        Eo, mo = og.find_essential_ransac(p, p.copy(), K)
        assert np.abs(np.asarray(Eo) - E0).max() < 1e-15 and int(np.asarray(mo).sum()) == M
        # recoverPose on it is itself ill-defined: two equal singular values and no point in front of any of the
        # four candidates (count 0), so cv2 returns R = I or the 180-degree twisted rotation depending on 1e-17 noise
        n, R, _, _ = cv2.recoverPose(E, p, p.copy(), K)
        assert n == 0 and min(og.rotation_angle_deg(R, np.eye(3)), abs(180.0 - og.rotation_angle_deg(R, np.eye(3)))) < 1e-4


def test_cv2_tie_order_is_rounding_noise():
    """Why the engine does not reproduce cv2.recoverPose's pick when all four candidates tie:
    the order of {R1, R2} out of cv2.decomposeEssentialMat on a rank-2 E follows the rounding residue of its SVD -
    a 1e-13 perturbation of E swaps them - while the candidate SET is stable."""
    rng = np.random.default_rng(2)
    swapped = 0
    for _ in range(40):
        r, _ = cv2.Rodrigues(rng.uniform(-1, 1, 3))
        t = rng.normal(size=3)
        t /= np.linalg.norm(t)
        tx = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]])
        E = tx @ r
        E /= np.linalg.norm(E)
        R1, R2, _ = cv2.decomposeEssentialMat(E)
        for _ in range(8):
            Q1, Q2, _ = cv2.decomposeEssentialMat(E + rng.normal(size=(3, 3)) * 1e-13)
            same = np.allclose(Q1, R1, atol=1e-6) and np.allclose(Q2, R2, atol=1e-6)
            swap = np.allclose(Q1, R2, atol=1e-6) and np.allclose(Q2, R1, atol=1e-6)
            assert same or swap                      # the set {R1, R2} never changes
            if swap:
                swapped += 1
                break
    assert swapped >= 30, swapped


def test_null_space_basis_and_model_order_are_cv2s():
    """cv2's five-point solver parametrises E over the null-space basis its (Jacobi) SVD returns, and the order of
    the models of a hypothesis is the order in which cv2.solvePoly finds the roots of the polynomial in the third
    coordinate.  The restatement reproduces the basis (vs cv2.SVDecomp) and therefore the ORDER of the models (vs
    cv2.findEssentialMat on exactly five points, which returns all models stacked in its own order)."""
    rng = np.random.default_rng(3)
    K = np.array([[800., 0, 320], [0, 800, 240], [0, 0, 1]])
    n = 300
    X = np.c_[rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(3, 9, n)]
    R, _ = cv2.Rodrigues(np.array([0.02, -0.03, 0.01]))
    t = np.array([0.3, 0.05, 0.1])
    p1 = X @ K.T; p1 = p1[:, :2] / p1[:, 2:]
    X2 = X @ R.T + t
    p2 = X2 @ K.T; p2 = p2[:, :2] / p2[:, 2:]
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    same = 0
    trials = 40
    for _ in range(trials):
        s = rng.choice(n, 5, replace=False)
        a, b = x1[s], x2[s]
        Q = np.stack([b[:, 0] * a[:, 0], b[:, 0] * a[:, 1], b[:, 0], b[:, 1] * a[:, 0], b[:, 1] * a[:, 1], b[:, 1],
                      a[:, 0], a[:, 1], np.ones(5)], axis=1)
        _, _, vt = cv2.SVDecomp(Q, flags=cv2.SVD_FULL_UV)
        assert np.abs(og.cv2_null_space(Q) - vt[5:9]).max() < 1e-12
        Es, _ = cv2.findEssentialMat(a, b, np.eye(3), method=cv2.RANSAC, prob=0.999, threshold=1.0)
        want = [] if Es is None else [Es[3 * i:3 * i + 3] for i in range(Es.shape[0] // 3)]
        got = og.five_point_models(a, b)
        same += len(got) == len(want) and all(min(np.abs(g - w).max(), np.abs(g + w).max()) < 1e-6 for g, w in zip(got, want))
    assert same >= trials - 1, same       # (a complex pair within 1e-10 of the real axis may be kept by one and not the other)
