"""CPU restatement of findEssentialMat(RANSAC, 5-point) / recoverPose / Euler (numpy).

TEST INFRASTRUCTURE ONLY (see oracle/orb.py header for the import rule).

Restates, in float64, what the reference obtains from
``cv2.findEssentialMat(pts1, pts2, K, RANSAC, 0.999, 1.0)`` and
``cv2.recoverPose(E, pts1, pts2, K)`` (reference: src/core/pose_estimator.py:522-527, :533)
and ``rotation_to_euler*`` (reference: src/utils/geometry.py:14-27, 48-80, 181-205).
The arithmetic is OpenCV's (third-party, build pinned 4.13.0): Nister five-point solver in
the Stewenius/Li formulation (null space -> 10 cubic constraints -> 10x20 elimination ->
degree-10 polynomial in z -> Durand-Kerner roots), OpenCV's RANSAC driver with its MWC RNG
seeded 2^64-1 and adaptive iteration count, and the four-candidate cheirality test
(SURVEY.md A.10, A.11).  Pinned against live cv2 by tests/test_oracle_geometry.py:
identical RNG subsets, identical inlier masks with cv2's own minimal solver plugged in,
E to ~1e-12 and identical masks (IoU>=0.95 bar) with the numpy solver.
"""
from __future__ import annotations

import math

import numpy as np

# ----------------------------------------------------------------------------- RNG
CV_RNG_COEFF = 4164903690


class CvRNG:
    """OpenCV's multiply-with-carry RNG (cv::RNG), as used by RANSACPointSetRegistrator."""

    def __init__(self, state: int = 0xFFFFFFFFFFFFFFFF):
        self.state = state & 0xFFFFFFFFFFFFFFFF

    def next(self) -> int:
        self.state = ((self.state & 0xFFFFFFFF) * CV_RNG_COEFF + (self.state >> 32)) & 0xFFFFFFFFFFFFFFFF
        return self.state & 0xFFFFFFFF

    def uniform(self, a: int, b: int) -> int:
        return self.next() % (b - a) + a


def draw_subset(rng: CvRNG, count: int, m: int = 5) -> list[int]:
    """RANSACPointSetRegistrator::getSubset: redraw any index equal to an earlier one."""
    idx: list[int] = []
    for _ in range(m):
        while True:
            v = rng.uniform(0, count)
            if v not in idx:
                break
        idx.append(v)
    return idx


def ransac_update_num_iters(p: float, ep: float, model_points: int, max_iters: int) -> int:
    """cv::RANSACUpdateNumIters."""
    DBL_MIN = 2.2250738585072014e-308
    p = min(max(p, 0.0), 1.0)
    ep = min(max(ep, 0.0), 1.0)
    num = max(1.0 - p, DBL_MIN)
    denom = 1.0 - (1.0 - ep) ** model_points
    if denom < DBL_MIN:
        return 0
    num = math.log(num)
    denom = math.log(denom)
    if denom >= 0 or -num >= max_iters * (-denom):
        return max_iters
    return int(np.rint(num / denom))


# ----------------------------------------------------------------------------- 5-point
# monomial order of the 10x20 constraint matrix (exponents of x, y, z)
_MONO = [(3, 0, 0), (0, 3, 0), (2, 1, 0), (1, 2, 0), (2, 0, 1), (2, 0, 0), (0, 2, 1), (0, 2, 0),
         (1, 1, 1), (1, 1, 0),
         (1, 0, 2), (1, 0, 1), (1, 0, 0), (0, 1, 2), (0, 1, 1), (0, 1, 0), (0, 0, 3), (0, 0, 2),
         (0, 0, 1), (0, 0, 0)]
_MONO_IDX = {m: i for i, m in enumerate(_MONO)}


def _pmul(a: dict, b: dict) -> dict:
    out: dict = {}
    for (i1, j1, k1), c1 in a.items():
        for (i2, j2, k2), c2 in b.items():
            key = (i1 + i2, j1 + j2, k1 + k2)
            out[key] = out.get(key, 0.0) + c1 * c2
    return out


def _padd(a: dict, b: dict, s: float = 1.0) -> dict:
    out = dict(a)
    for k, c in b.items():
        out[k] = out.get(k, 0.0) + s * c
    return out


def _constraint_matrix(basis: np.ndarray) -> np.ndarray:
    """10x20 coefficients of det(E)=0 and 2EE'E - tr(EE')E = 0 for E = x E1 + y E2 + z E3 + E4."""
    E = [[{(1, 0, 0): basis[0][3 * r + c], (0, 1, 0): basis[1][3 * r + c],
           (0, 0, 1): basis[2][3 * r + c], (0, 0, 0): basis[3][3 * r + c]} for c in range(3)] for r in range(3)]
    rows = []
    # determinant
    det: dict = {}
    for (a, b, c, s) in [(0, 1, 2, 1), (1, 2, 0, 1), (2, 0, 1, 1), (0, 2, 1, -1), (1, 0, 2, -1), (2, 1, 0, -1)]:
        det = _padd(det, _pmul(_pmul(E[0][a], E[1][b]), E[2][c]), s)
    rows.append(det)
    # EEt
    EEt = [[{} for _ in range(3)] for _ in range(3)]
    for r in range(3):
        for c in range(3):
            acc: dict = {}
            for k in range(3):
                acc = _padd(acc, _pmul(E[r][k], E[c][k]))
            EEt[r][c] = acc
    tr = _padd(_padd(EEt[0][0], EEt[1][1]), EEt[2][2])
    for r in range(3):
        for c in range(3):
            acc = {}
            for k in range(3):
                acc = _padd(acc, _pmul(EEt[r][k], E[k][c]), 2.0)
            acc = _padd(acc, _pmul(tr, E[r][c]), -1.0)
            rows.append(acc)
    A = np.zeros((10, 20))
    for i, poly in enumerate(rows):
        for m, c in poly.items():
            A[i, _MONO_IDX[m]] = c
    return A


def solve_poly_dk(coeffs: np.ndarray, max_iters: int = 300) -> np.ndarray:
    """cv::solvePoly (Durand-Kerner, in-place updates, start 1,(1+i),(1+i)^2,...); coeffs[i]*x^i.
    Matches cv2.solvePoly root order and values (tests/test_oracle_geometry.py)."""
    c = np.asarray(coeffs, dtype=np.float64)
    n = len(c) - 1
    while n > 0 and c[n] == 0:
        n -= 1
    cc = c.astype(np.complex128)
    roots = np.zeros(n, dtype=np.complex128)
    p = complex(1, 0)
    r = complex(1, 1)
    for i in range(n):
        roots[i] = p
        p = p * r
    for _ in range(max_iters):
        maxdiff = 0.0
        for i in range(n):
            p = roots[i]
            num = cc[n]
            den = cc[n]
            for j in range(n):
                num = num * p + cc[n - j - 1]
                if j != i:
                    d = p - roots[j]
                    if d != 0:
                        den = den * d
            num = num / den
            roots[i] = p - num
            maxdiff = max(maxdiff, abs(num))
        if maxdiff <= 0:
            break
    return roots


def _cv_svd_fill_signs() -> np.ndarray:
    """The 36 draws of cv::RNG(0x12345678) that OpenCV's JacobiSVDImpl_ turns into +-1 (bit 8 of each draw)."""
    state, out = 0x12345678, []
    for _ in range(36):
        state = ((state & 0xFFFFFFFF) * 4164903690 + (state >> 32)) & 0xFFFFFFFFFFFFFFFF
        out.append(1.0 if (state & 0xFFFFFFFF) & 256 else -1.0)
    return np.array(out).reshape(4, 9)


_SVD_FILL = _cv_svd_fill_signs()


def cv2_null_space(Q: np.ndarray) -> np.ndarray:
    """Rows 5..8 of Vt as cv2's five-point solver gets them from SVD::compute(Q, FULL_UV) on its 5 x 9 system.
    OpenCV's HAL hands matrices this small to its own one-sided Jacobi SVD even in LAPACK builds, and that routine
    FILLS the rows beyond the rank by Gram-Schmidt: row i = a fixed pseudo-random +-1/9 vector (RNG re-seeded with
    0x12345678 on every call), orthogonalised twice against all previous rows with an L1 rescale after every
    projection, then normalised.  Only the row space of Q enters.  Verified against cv2.SVDecomp to 1e-15
    (tests/test_oracle_geometry.py); this basis fixes the parametrisation E = x E1 + y E2 + z E3 + E4, hence the
    order in which cv2.solvePoly returns the roots in z, hence the ORDER of the models of a hypothesis."""
    eps = np.finfo(np.float64).eps * 10
    rows = list(np.linalg.qr(Q.T, mode="complete")[0][:, :5].T)
    for i in range(4):
        v = _SVD_FILL[i] / 9.0
        for _ in range(2):
            for r in rows:
                v = v - float(v @ r) * r
                asum = np.abs(v).sum()
                v = v * (1.0 / asum if asum > eps * 100 else 0.0)
        rows.append(v / np.sqrt((v * v).sum()))
    return np.array(rows[5:])


def five_point_models(x1: np.ndarray, x2: np.ndarray, roots_fn=None) -> list[np.ndarray]:
    """All real essential matrices through 5 normalised correspondences (x2' E x1 = 0),
    each with ||E||_F = 1.  x1, x2: (5, 2) float64."""
    x1 = np.asarray(x1, dtype=np.float64)
    x2 = np.asarray(x2, dtype=np.float64)
    if np.array_equal(x1, x2):
        # static sample (no camera motion): the system is rank deficient and cv2 4.13 returns this skew-symmetric
        # matrix (R = I) for ANY such sample — pinned against live cv2 in tests/test_oracle_geometry.py
        a = 1.0 / np.sqrt(6.0)
        return [np.array([[0.0, -a, -a], [a, 0.0, -a], [a, a, 0.0]])]
    Q = np.stack([x2[:, 0] * x1[:, 0], x2[:, 0] * x1[:, 1], x2[:, 0],
                  x2[:, 1] * x1[:, 0], x2[:, 1] * x1[:, 1], x2[:, 1],
                  x1[:, 0], x1[:, 1], np.ones(5)], axis=1)
    basis = cv2_null_space(Q)                         # 4 x 9, rows = E1..E4 (row-major 3x3)
    A = _constraint_matrix(basis)
    try:
        G = np.linalg.solve(A[:, :10], A[:, 10:])     # Gauss-Jordan of the first ten monomials
    except np.linalg.LinAlgError:
        return []
    B = np.zeros((3, 13))
    for i in range(3):
        r1 = G[2 * i + 4]
        r2 = G[2 * i + 5]
        row1 = np.zeros(13); row2 = np.zeros(13)
        row1[1:4] = r1[0:3]; row1[5:8] = r1[3:6]; row1[9:13] = r1[6:10]
        row2[0:3] = r2[0:3]; row2[4:7] = r2[3:6]; row2[8:12] = r2[6:10]
        B[i] = row1 - row2
    # det of the 3x3 polynomial matrix [p3 p3 p4] (np.poly: highest power first)
    P = [[B[i, 0:4], B[i, 4:8], B[i, 8:13]] for i in range(3)]

    def pm(a, b):
        return np.polymul(a, b)

    det = np.zeros(11)
    for (a, b, c, s) in [(0, 1, 2, 1), (1, 2, 0, 1), (2, 0, 1, 1), (0, 2, 1, -1), (1, 0, 2, -1), (2, 1, 0, -1)]:
        term = pm(pm(P[0][a], P[1][b]), P[2][c])
        det[11 - len(term):] += s * term
    coeffs_low_first = det[::-1]
    if not np.all(np.isfinite(coeffs_low_first)):
        return []
    roots = (roots_fn or solve_poly_dk)(coeffs_low_first)
    models = []
    for rt in roots:
        if abs(rt.imag) > 1e-10:
            continue
        z = rt.real
        Bz = np.array([[np.polyval(P[j][0], z), np.polyval(P[j][1], z), np.polyval(P[j][2], z)] for j in range(3)])
        _, _, vt = np.linalg.svd(Bz)
        v = vt[2]
        if abs(v[2]) < 1e-10:
            continue
        x, y = v[0] / v[2], v[1] / v[2]
        e = x * basis[0] + y * basis[1] + z * basis[2] + basis[3]
        e = e / np.linalg.norm(e)
        models.append(e.reshape(3, 3))
    return models


def sampson_error_f32(E: np.ndarray, x1: np.ndarray, x2: np.ndarray) -> np.ndarray:
    """EMEstimatorCallback::computeError: f64 Sampson distance stored as float32.  Every product and sum is rounded
    on its own, in cv2's order (Matx row times vector left to right, the dot product, a + b + c + d): cv2 is built
    without FMA contraction, and numpy's elementwise operations round the same way."""
    E = np.asarray(E, dtype=np.float64).reshape(3, 3)
    xa, ya, xb, yb = x1[:, 0], x1[:, 1], x2[:, 0], x2[:, 1]
    Ex0 = (E[0, 0] * xa + E[0, 1] * ya) + E[0, 2]
    Ex1 = (E[1, 0] * xa + E[1, 1] * ya) + E[1, 2]
    Ex2 = (E[2, 0] * xa + E[2, 1] * ya) + E[2, 2]
    Et0 = (E[0, 0] * xb + E[1, 0] * yb) + E[2, 0]
    Et1 = (E[0, 1] * xb + E[1, 1] * yb) + E[2, 1]
    x2tEx1 = (xb * Ex0 + yb * Ex1) + Ex2
    den = ((Ex0 * Ex0 + Ex1 * Ex1) + Et0 * Et0) + Et1 * Et1
    with np.errstate(divide="ignore", invalid="ignore"):
        return (x2tEx1 * x2tEx1 / den).astype(np.float32)


def normalize_points(pts: np.ndarray, K: np.ndarray) -> np.ndarray:
    pts = np.asarray(pts, dtype=np.float64).reshape(-1, 2)
    fx, fy, cx, cy = K[0, 0], K[1, 1], K[0, 2], K[1, 2]
    return np.stack([(pts[:, 0] - cx) / fx, (pts[:, 1] - cy) / fy], axis=1)


def find_essential_ransac(pts1, pts2, K, prob=0.999, threshold=1.0, max_iters=1000,
                          solver=None, trace=None):
    """cv2.findEssentialMat(pts1, pts2, K, RANSAC, prob, threshold, maxIters).

    ``solver(x1_5, x2_5) -> list of 3x3`` defaults to the numpy five-point solver; the tests
    also plug in cv2's own minimal solver to pin the driver hypothesis-for-hypothesis.
    Returns (E or None, mask uint8 (M,) in {0,1}).  ``trace`` (list) receives
    (iteration, subset, [inlier count per model]).
    """
    K = np.asarray(K, dtype=np.float64)
    x1 = normalize_points(pts1, K)
    x2 = normalize_points(pts2, K)
    M = x1.shape[0]
    solver = solver or five_point_models
    thr = threshold / ((K[0, 0] + K[1, 1]) / 2.0)
    t = np.float32(thr * thr)
    if M < 5:
        return None, np.zeros(M, np.uint8)
    rng = CvRNG()
    niters = max_iters
    best = 0
    bestE = None
    bestmask = None
    it = 0
    while it < niters:
        subset = list(range(5)) if M == 5 else draw_subset(rng, M)
        models = solver(x1[subset], x2[subset])
        counts = []
        for E in models:
            err = sampson_error_f32(E, x1, x2)
            mask = err <= t
            cnt = int(mask.sum())
            counts.append(cnt)
            if cnt > max(best, 4):
                best = cnt
                bestE = E
                bestmask = mask
                niters = ransac_update_num_iters(prob, (M - cnt) / M, 5, niters)
        if trace is not None:
            trace.append((it, subset, counts))
        it += 1
    if bestE is None:
        return None, np.zeros(M, np.uint8)
    return bestE, bestmask.astype(np.uint8)


# ----------------------------------------------------------------------------- recoverPose
def decompose_essential(E: np.ndarray):
    U, _, Vt = np.linalg.svd(E)
    if np.linalg.det(U) < 0:
        U = -U
    if np.linalg.det(Vt) < 0:
        Vt = -Vt
    W = np.array([[0.0, 1, 0], [-1, 0, 0], [0, 0, 1]])
    return U @ W @ Vt, U @ W.T @ Vt, U[:, 2].copy()


def _triangulate(P1: np.ndarray, x1: np.ndarray, x2: np.ndarray) -> np.ndarray:
    """DLT with P0 = [I|0]: null vector of the 4x4 system per point.  Returns (4, M)."""
    M = x1.shape[0]
    P0 = np.hstack([np.eye(3), np.zeros((3, 1))])
    A = np.zeros((M, 4, 4))
    A[:, 0] = x1[:, 0:1] * P0[2] - P0[0]
    A[:, 1] = x1[:, 1:2] * P0[2] - P0[1]
    A[:, 2] = x2[:, 0:1] * P1[2] - P1[0]
    A[:, 3] = x2[:, 1:2] * P1[2] - P1[1]
    _, _, vt = np.linalg.svd(A)
    return vt[:, 3, :].T


def recover_pose(E, pts1, pts2, K, distance_thresh=50.0):
    """cv2.recoverPose(E, pts1, pts2, K).  Returns (count, R, t(3,1), mask 0/255)."""
    K = np.asarray(K, dtype=np.float64)
    x1 = normalize_points(pts1, K)
    x2 = normalize_points(pts2, K)
    R1, R2, t = decompose_essential(np.asarray(E, dtype=np.float64))
    cands = [(R1, t), (R2, t), (R1, -t), (R2, -t)]
    masks = []
    for R, tt in cands:
        P1 = np.hstack([R, tt.reshape(3, 1)])
        Q = _triangulate(P1, x1, x2)
        with np.errstate(divide="ignore", invalid="ignore"):
            m = (Q[2] * Q[3]) > 0
            Qn = Q / Q[3]
            m &= Qn[2] < distance_thresh
            Q2 = P1 @ Qn
            m &= (Q2[2] > 0) & (Q2[2] < distance_thresh)
        masks.append(m)
    good = [int(m.sum()) for m in masks]
    g1, g2, g3, g4 = good
    if g1 >= g2 and g1 >= g3 and g1 >= g4:
        k = 0
    elif g2 >= g1 and g2 >= g3 and g2 >= g4:
        k = 1
    elif g3 >= g1 and g3 >= g2 and g3 >= g4:
        k = 2
    else:
        k = 3
    R, tt = cands[k]
    return good[k], R, tt.reshape(3, 1), (masks[k].astype(np.uint8) * 255)


# ----------------------------------------------------------------------------- Euler
def rotation_to_euler_yup(R):
    """reference src/utils/geometry.py:48-80 -> (yaw, pitch, roll) degrees, R = Ry Rx Rz."""
    pitch = math.asin(R[2, 1])
    if abs(R[2, 1]) > 0.9999:
        roll = math.atan2(-R[1, 2], R[1, 1])
        yaw = 0.0
    else:
        yaw = math.atan2(-R[2, 0], R[0, 0])
        roll = math.atan2(R[1, 0], R[1, 1])
    return math.degrees(yaw), math.degrees(pitch), math.degrees(roll)


def rotation_to_euler_zyx(R):
    """reference src/utils/geometry.py:181-205 -> (yaw, pitch, roll) degrees, R = Rz Ry Rx."""
    sy = math.sqrt(R[0, 0] ** 2 + R[1, 0] ** 2)
    if not sy < 1e-6:
        roll = math.atan2(R[2, 1], R[2, 2])
        pitch = math.atan2(-R[2, 0], sy)
        yaw = math.atan2(R[1, 0], R[0, 0])
    else:
        roll = math.atan2(-R[1, 2], R[1, 1])
        pitch = math.atan2(-R[2, 0], sy)
        yaw = 0.0
    return math.degrees(yaw), math.degrees(pitch), math.degrees(roll)


def rotation_to_euler(R, convention="yup"):
    """reference src/utils/geometry.py:14-27."""
    if convention == "zyx":
        return rotation_to_euler_zyx(R)
    return rotation_to_euler_yup(R)


def rotation_angle_deg(Ra, Rb) -> float:
    c = (np.trace(Ra @ Rb.T) - 1.0) / 2.0
    return math.degrees(math.acos(min(1.0, max(-1.0, c))))


def direction_angle_deg(ta, tb) -> float:
    ta = np.asarray(ta).ravel(); tb = np.asarray(tb).ravel()
    c = float(ta @ tb) / (np.linalg.norm(ta) * np.linalg.norm(tb))
    return math.degrees(math.acos(min(1.0, max(-1.0, c))))
