"""The per-element arithmetic of the CUDA kernels (csrc/orb_math.cuh, geom_math.cuh), instantiated on
the host by tests/hostcheck (TEST-ONLY, g++), against the oracle: bit-exact for integer / f32 work,
tolerance for the FP64 geometry.  Catches arithmetic regressions on a box without a GPU; the GPU
parity tests (test_gpu_parity.py) are the real gate."""
import cv2
import numpy as np
import pytest
from conftest import default_K, load_gray

import _hostcheck as hc
from oracle import geometry as og
from oracle import matcher as om
from oracle import orb as oo


@pytest.fixture(scope="module")
def img():
    return load_gray("sim_000015.png")


def test_resize_bit_exact(img):
    x0, _, _, cx1 = oo.linear_exact_coeffs(640, 582)
    y0, _, _, cy1 = oo.linear_exact_coeffs(480, 436)
    got = hc.resize(img, 582, 436, x0, cx1, y0, cy1)
    assert np.array_equal(got, oo.resize_linear_exact(img, 582, 436))


def test_fast_scores_bit_exact(img):
    assert np.array_equal(hc.fast_scores(img, 15), oo.fast_score_map(img, 15))
    noise = np.random.default_rng(0).integers(0, 256, (64, 96), dtype=np.uint8)
    for t in (1, 15, 60):
        assert np.array_equal(hc.fast_scores(noise, t), oo.fast_score_map(noise, t))
        assert np.array_equal(hc.fast_scores_scalar(noise, t), oo.fast_score_map(noise, t))
    # extremes of the packed form: saturated rings around dark / bright centres
    ext = np.zeros((16, 16), np.uint8); ext[8, 8] = 255
    ext2 = np.full((16, 16), 255, np.uint8); ext2[8, 8] = 0
    for im in (ext, ext2):
        for t in (1, 127, 254):
            assert np.array_equal(hc.fast_scores(im, t), oo.fast_score_map(im, t))


def test_fast_packed_pretest_is_the_scalar_compass_test(img):
    """The 4-pixel packed pre-test of k_fast equals the scalar two-pair compass test exactly, and
    no FAST corner fails it (so compaction never drops a corner)."""
    rng = np.random.default_rng(2)
    hard = rng.integers(0, 256, (96, 128), dtype=np.uint8)
    hard[:, :40] = rng.choice(np.array([0, 1, 14, 15, 16, 17, 127, 128, 129, 239, 240, 241, 254, 255], np.uint8), (96, 40))
    for im, thrs in ((img, (15,)), (hard, (1, 15, 16, 100, 127))):
        I = im.astype(np.int16)
        h, w = im.shape
        for t in thrs:
            got = hc.fast_compass(im, t)
            c = I[3:h - 3, 4:w - 4]
            d0, d8 = c - I[6:h, 4:w - 4], c - I[0:h - 6, 4:w - 4]
            d4, d12 = c - I[3:h - 3, 7:w - 1], c - I[3:h - 3, 1:w - 7]
            want = (((d0 > t) | (d8 > t)) & ((d4 > t) | (d12 > t))) | (((d0 < -t) | (d8 < -t)) & ((d4 < -t) | (d12 < -t)))
            x_hi = 4 + ((w - 8) // 4) * 4
            assert np.array_equal(got[3:h - 3, 4:x_hi].astype(bool), want[:, :x_hi - 4]), t
            corners = oo.fast_score_map(im, t)[3:h - 3, 4:x_hi] > 0
            assert not (corners & ~want[:, :x_hi - 4]).any()


def test_harris_angle_bit_exact(img):
    xs, ys, _ = oo.fast_candidates(img, 15)
    assert len(xs) > 1000
    assert np.array_equal(hc.harris(img, xs, ys).view(np.uint32), oo.harris_response(img, xs, ys).view(np.uint32))
    assert np.array_equal(hc.angle(img, xs, ys).view(np.uint32), oo.ic_angle(img, xs, ys).view(np.uint32))
    rng = np.random.default_rng(1)
    y, x = rng.integers(-5000, 5000, 5000).astype(np.float32), rng.integers(-5000, 5000, 5000).astype(np.float32)
    assert np.array_equal(hc.atan2(y, x), oo.fast_atan2(y, x))


def test_blur_and_brief_bit_exact(img):
    bl = oo.blur7(img)
    assert np.array_equal(hc.blur(img), bl)
    xs, ys, _ = oo.fast_candidates(img, 15)
    ang = oo.ic_angle(img, xs, ys)
    want = oo.brief_descriptors(bl, xs.astype(np.float32), ys.astype(np.float32), ang, np.float32(1.0))
    assert np.array_equal(hc.brief(bl, xs, ys, ang), want)


def test_rng_and_iteration_update():
    rng = og.CvRNG()
    want = [og.draw_subset(rng, 500) for _ in range(64)]
    assert hc.subsets(500, 64).tolist() == want
    rng = og.CvRNG()
    assert hc.subsets(7, 32).tolist() == [og.draw_subset(rng, 7) for _ in range(32)]    # many redraws
    for ep in (0.0, 0.1, 76 / 500, 0.5, 0.9, 0.999, 1.0):
        for mx in (12, 1000, 2000):
            assert hc.update_iters(0.999, ep, 5, mx) == og.ransac_update_num_iters(0.999, ep, 5, mx)


def _points():
    r1 = oo.orb_detect_and_compute(load_gray("sim_000000.png"))
    r2 = oo.orb_detect_and_compute(load_gray("sim_000015.png"))
    m = om.sort_and_cut(om.match_crosscheck(r1.desc, r2.desc), 500)
    p1 = np.stack([r1.x[m[:, 0]], r1.y[m[:, 0]]], 1).astype(np.float32)
    p2 = np.stack([r2.x[m[:, 1]], r2.y[m[:, 1]]], 1).astype(np.float32)
    return p1, p2


def test_five_point_and_sampson():
    p1, p2 = _points()
    K = default_K()
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    rng = og.CvRNG()
    thr = np.float32((1.0 / ((K[0, 0] + K[1, 1]) / 2)) ** 2)
    for _ in range(10):
        s = og.draw_subset(rng, len(x1))
        mine, want = hc.five_point(x1[s], x2[s]), og.five_point_models(x1[s], x2[s])
        assert len(mine) == len(want)
        for E in mine:
            assert abs(np.linalg.norm(E) - 1) < 1e-9
            res = np.einsum("ni,ij,nj->n", np.c_[x2[s], np.ones(5)], E, np.c_[x1[s], np.ones(5)])
            assert np.abs(res).max() < 1e-9
            T = min(want, key=lambda W: min(np.abs(E - W).max(), np.abs(E + W).max()))
            c1 = int((hc.sampson(E, x1, x2) <= thr).sum())
            c2 = int((og.sampson_error_f32(T, x1, x2) <= thr).sum())
            assert abs(c1 - c2) <= 3
    E = want[0]
    assert np.array_equal(hc.sampson(E, x1, x2), og.sampson_error_f32(E, x1, x2))


def test_division_free_inlier_gate_is_the_reference_decision():
    """sampson_is_inlier (no FP64 division) == `(float)(num2 / den) <= t2` for every point, including thresholds placed
    exactly ON a point's error, one float above / below it, and the degenerate den == 0."""
    p1, p2 = _points()
    K = default_K()
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    E, _ = cv2.findEssentialMat(p1, p2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    rng = np.random.default_rng(9)
    for Em in (E, rng.standard_normal((3, 3)), np.zeros((3, 3))):
        err = hc.sampson(Em, x1, x2)
        finite = err[np.isfinite(err)]
        thrs = [np.float32((1.0 / ((K[0, 0] + K[1, 1]) / 2)) ** 2)]
        for e in (finite[:40] if len(finite) else []):
            if e > 0:
                thrs += [np.float32(e), np.nextafter(np.float32(e), np.float32(np.inf)), np.nextafter(np.float32(e), np.float32(0))]
        for t2 in thrs:
            with np.errstate(invalid="ignore"):
                want = err <= t2
            assert np.array_equal(hc.sampson_inlier(Em, x1, x2, t2), want), t2


def test_decompose_cheirality_euler():
    p1, p2 = _points()
    K = default_K()
    E, _ = cv2.findEssentialMat(p1, p2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    R1, R2, t = hc.decompose(E)
    o1, o2, ot = og.decompose_essential(E)
    cands = [o1, o2]
    for R in (R1, R2):
        assert min(np.abs(R - C).max() for C in cands) < 1e-12 and abs(np.linalg.det(R) - 1) < 1e-12
    assert min(np.abs(t - ot).max(), np.abs(t + ot).max()) < 1e-12
    n_cv, R_cv, t_cv, mask_cv = cv2.recoverPose(E, p1, p2, K)
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    ok = hc.cheirality(R_cv, t_cv, x1, x2)
    assert np.array_equal(ok, mask_cv.ravel() > 0) and int(ok.sum()) == n_cv
    for conv, name in ((0, "yup"), (1, "zyx")):
        assert np.allclose(hc.euler(R_cv, conv), og.rotation_to_euler(R_cv, name), atol=1e-12)
    gimbal = np.array([[1.0, 0, 0], [0, 0, -1], [0, 1, 0]])          # pitch = +90 deg
    assert np.allclose(hc.euler(gimbal, 0), og.rotation_to_euler_yup(gimbal), atol=1e-12)


def test_solve_poly10_against_numpy_roots():
    rng = np.random.default_rng(5)
    for _ in range(5):
        roots = np.concatenate([rng.uniform(-3, 3, 4), (lambda z: np.r_[z, z.conj()])(rng.uniform(-2, 2, 3) + 1j * rng.uniform(0.2, 2, 3))])
        c = np.real(np.poly(roots))[::-1].copy()                     # c[i] * x^i
        got = hc.solve_poly10(c)
        for r in roots:
            assert np.abs(got - r).min() < 1e-7


def test_hamming_carry_save_forms_are_exact():
    """match_math.cuh: 3- and 4-adder carry-save popcount == 8 plain popcounts == numpy, incl. extremes."""
    rng = np.random.default_rng(7)
    q = rng.integers(0, 256, (96, 32), dtype=np.uint8)
    t = rng.integers(0, 256, (80, 32), dtype=np.uint8)
    q[0] = 0; q[1] = 255; t[0] = 0; t[1] = 255; t[2] = q[5]
    q[2] = np.unpackbits(np.arange(32, dtype=np.uint8))[:32] * 255     # sparse / dense word mixes
    want = np.unpackbits(q[:, None, :] ^ t[None, :, :], axis=2).sum(axis=2).astype(np.int32)
    for mode in (0, 1, 2, 3, 4, 5, 6):    # 3 / 4: the multiply-add key forms (hamming256_key); 5 / 6: pre-combined operands
        assert np.array_equal(hc.hamming(q, t, mode), want), mode
    assert want[1, 0] == 256 and want[0, 0] == 0 and want[5, 2] == 0


def test_five_point_matches_cv2_minimal_solver():
    """The kernel's five-point solver (cv2's Jacobi-SVD null basis, Gauss-Jordan, Durand-Kerner + Newton)
    against cv2's OWN minimal solver — cv2.findEssentialMat called with exactly five points returns
    all its roots stacked — on a generic 3-D scene: same models, same order."""
    import cv2
    rng = np.random.default_rng(3)
    K = np.array([[800., 0, 320], [0, 800, 240], [0, 0, 1]])
    n = 300
    X = np.c_[rng.uniform(-2, 2, n), rng.uniform(-1.5, 1.5, n), rng.uniform(3, 9, n)]
    R, _ = cv2.Rodrigues(np.array([0.02, -0.03, 0.01]))
    t = np.array([0.3, 0.05, 0.1])
    p1 = X @ K.T
    p1 = p1[:, :2] / p1[:, 2:]
    X2 = X @ R.T + t
    p2 = X2 @ K.T
    p2 = p2[:, :2] / p2[:, 2:]
    p1 = (p1 + rng.normal(0, 0.3, p1.shape)).astype(np.float32)
    p2 = (p2 + rng.normal(0, 0.3, p2.shape)).astype(np.float32)
    x1, x2 = og.normalize_points(p1, K), og.normalize_points(p2, K)
    bad = 0
    for _ in range(120):
        s = rng.choice(n, 5, replace=False)
        Es, _ = cv2.findEssentialMat(p1[s].astype(np.float64), p2[s].astype(np.float64), K, method=cv2.RANSAC,
                                     prob=0.999, threshold=1.0)
        want = [] if Es is None else [Es[3 * i:3 * i + 3] for i in range(Es.shape[0] // 3)]
        got = [e.reshape(3, 3) for e in hc.five_point(x1[s], x2[s])]
        # same models IN THE SAME ORDER (the order decides RANSAC ties between the models of one hypothesis)
        ok = len(want) == len(got) and all(min(np.abs(a - b).max(), np.abs(a + b).max()) < 1e-5 for a, b in zip(got, want))
        bad += not ok
        for e in got:                              # every model really is an essential matrix through the sample
            assert np.abs(2 * e @ e.T @ e - np.trace(e @ e.T) * e).max() < 1e-6
            assert max(abs(np.r_[x2[i], 1] @ e @ np.r_[x1[i], 1]) for i in s) < 1e-12
    assert bad <= 3, bad
