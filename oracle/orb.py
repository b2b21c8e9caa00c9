"""CPU restatement of the ORB stage of the reference hot path (numpy).

TEST INFRASTRUCTURE ONLY.  Nothing under ``relative-pose-estimation_b200/`` may
import this module; only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` do, as the checker.

What it restates
----------------
``cv2.ORB_create(nfeatures, scaleFactor=1.1, nlevels=12, fastThreshold=15,
scoreType=HARRIS).detectAndCompute(image, None)`` exactly as the reference calls
it (reference: src/core/pose_estimator.py:85-91 and :108).  The arithmetic lives
in OpenCV (third-party, un-vendored; reference pin ``opencv-python>=4.8.0``,
requirements.txt:1; build pinned here: 4.13.0), so every function below follows
OpenCV's published ORB algorithm as specified in SURVEY.md Appendix A, and is
pinned against live ``cv2`` 4.13 by ``tests/test_oracle_orb.py`` (0 differing
keypoints / descriptor bits on the committed fixture images).

Keypoint order: cv2's within-level order is an artefact of std::nth_element, so
both this oracle and the CUDA engine emit the CANONICAL order
(octave, y_level, x_level); ``canonicalize_cv2`` reorders cv2's output the same way.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

from .orb_pattern import PATTERN

EDGE_THRESHOLD = 31
PATCH_SIZE = 31
HALF_PATCH = 15
HARRIS_BLOCK = 7
HARRIS_K = np.float32(0.04)

# FAST-9/16 Bresenham ring, (dx, dy), index 0..15
FAST_RING = [(0, 3), (1, 3), (2, 2), (3, 1), (3, 0), (3, -1), (2, -2), (1, -3),
             (0, -3), (-1, -3), (-2, -2), (-3, -1), (-3, 0), (-3, 1), (-2, 2), (-1, 3)]

# circular-patch half widths for v = 0..15
UMAX = [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]

# f32 bit patterns of getGaussianKernel(7, 2)
BLUR_KERNEL = np.array([0x3d8fafb1, 0x3e06387e, 0x3e434a39, 0x3e5d4ae0,
                        0x3e434a39, 0x3e06387e, 0x3d8fafb1], dtype=np.uint32).view(np.float32)


@dataclass
class OrbParams:
    nfeatures: int = 4000
    scale_factor: float = 1.1      # passed to cv2 as a C float
    nlevels: int = 12
    fast_threshold: int = 15


@dataclass
class Level:
    w: int
    h: int
    scale: np.float32
    quota: int


def level_geometry(W: int, H: int, p: OrbParams) -> list[Level]:
    """Level sizes, f32 scales and per-level feature quotas."""
    sf = float(np.float32(p.scale_factor))           # ORB_create takes float scaleFactor
    scales = [np.float32(sf ** l) for l in range(p.nlevels)]
    # quotas: f32 arithmetic as in computeKeyPoints
    factor = np.float32(1.0 / sf)
    nd = np.float32(p.nfeatures) * (np.float32(1) - factor) / \
        (np.float32(1) - np.float32(float(factor) ** p.nlevels))
    nd = np.float32(nd)
    quotas = []
    total = 0
    for l in range(p.nlevels - 1):
        q = int(np.rint(nd))
        quotas.append(q)
        total += q
        nd = np.float32(nd * factor)
    quotas.append(max(p.nfeatures - total, 0))
    out = []
    for l in range(p.nlevels):
        w = int(np.rint(np.float32(W) / scales[l]))
        h = int(np.rint(np.float32(H) / scales[l]))
        out.append(Level(w, h, scales[l], quotas[l]))
    return out


# --------------------------------------------------------------------------- pyramid
def linear_exact_coeffs(src: int, dst: int):
    """Per-axis INTER_LINEAR_EXACT tap table: (ofs0, ofs1, c0, c1), 8.8 fixed."""
    scale = 1.0 / (float(dst) / float(src))
    d = np.arange(dst, dtype=np.float64)
    f = scale * (d + 0.5) - 0.5
    i = np.floor(f).astype(np.int64)
    c1 = np.rint((f - i) * 256.0).astype(np.int64)
    ofs0 = i.copy()
    ofs1 = i + 1
    c0 = 256 - c1
    left = (i < 0) | (src <= 1)
    right = (~left) & (i >= src - 1)
    ofs0[left] = 0; ofs1[left] = 0; c0[left] = 256; c1[left] = 0
    ofs0[right] = src - 1; ofs1[right] = src - 1; c0[right] = 256; c1[right] = 0
    return ofs0, ofs1, c0, c1


def resize_linear_exact(img: np.ndarray, w: int, h: int) -> np.ndarray:
    """cv2.resize(img, (w, h), interpolation=INTER_LINEAR_EXACT) for u8 single channel."""
    sh, sw = img.shape
    x0, x1, cx0, cx1 = linear_exact_coeffs(sw, w)
    y0, y1, cy0, cy1 = linear_exact_coeffs(sh, h)
    s = img.astype(np.int64)
    hline = cx0[None, :] * s[:, x0] + cx1[None, :] * s[:, x1]           # 8.8
    v = cy0[:, None] * hline[y0, :] + cy1[:, None] * hline[y1, :]       # 16.16
    out = (v + 32768) >> 16
    return np.minimum(out, 255).astype(np.uint8)


def build_pyramid(img: np.ndarray, levels: list[Level]) -> list[np.ndarray]:
    """Chained pyramid: level l is resized from level l-1."""
    pyr = [np.ascontiguousarray(img)]
    for l in range(1, len(levels)):
        pyr.append(resize_linear_exact(pyr[-1], levels[l].w, levels[l].h))
    return pyr


# --------------------------------------------------------------------------- FAST
def fast_score_map(img: np.ndarray, threshold: int) -> np.ndarray:
    """FAST-9/16 corner score for every pixel (0 = not a corner).

    score = max(a, b) - 1 with a = max over the 16 contiguous 9-arcs of min(v - p_k),
    b = same for (p_k - v); corner iff a > threshold or b > threshold.
    """
    h, w = img.shape
    out = np.zeros((h, w), dtype=np.int32)
    if h < 7 or w < 7:
        return out
    I = img.astype(np.int16)
    c = I[3:h - 3, 3:w - 3]
    d = np.stack([c - I[3 + dy:h - 3 + dy, 3 + dx:w - 3 + dx] for dx, dy in FAST_RING])  # (16,h-6,w-6)
    d2 = np.concatenate([d, d[:8]], axis=0)  # circular
    a = np.full(c.shape, -32768, dtype=np.int16)
    b = np.full(c.shape, -32768, dtype=np.int16)
    for s in range(16):
        win = d2[s:s + 9]
        a = np.maximum(a, win.min(axis=0))
        b = np.maximum(b, (-win).min(axis=0))
    corner = (a > threshold) | (b > threshold)
    score = np.maximum(a, b).astype(np.int32) - 1
    out[3:h - 3, 3:w - 3] = np.where(corner, score, 0)
    return out


def nms_3x3(score: np.ndarray) -> np.ndarray:
    """Keep a corner iff its score is strictly greater than all 8 neighbours' (SURVEY A.2)."""
    h, w = score.shape
    p = np.zeros((h + 2, w + 2), dtype=score.dtype)
    p[1:-1, 1:-1] = score
    keep = score > 0
    for dy in (-1, 0, 1):
        for dx in (-1, 0, 1):
            if dx == 0 and dy == 0:
                continue
            keep &= score > p[1 + dy:1 + dy + h, 1 + dx:1 + dx + w]
    return keep


def fast_candidates(img: np.ndarray, threshold: int):
    """Raster-ordered (x, y, score) of FAST+NMS corners inside the 31-px border (A.2)."""
    h, w = img.shape
    score = fast_score_map(img, threshold)
    keep = nms_3x3(score)
    e = EDGE_THRESHOLD
    inner = np.zeros_like(keep)
    if w > 2 * e and h > 2 * e:
        inner[e:h - e, e:w - e] = True
    keep &= inner
    ys, xs = np.nonzero(keep)                 # raster order
    return xs.astype(np.int32), ys.astype(np.int32), score[ys, xs].astype(np.int32)


def retain_best(resp: np.ndarray, n: int) -> np.ndarray:
    """KeyPointsFilter::retainBest as a boolean keep mask: ties with the n-th kept (A.4)."""
    if n <= 0:
        return np.zeros(resp.shape[0], dtype=bool)
    if resp.shape[0] <= n:
        return np.ones(resp.shape[0], dtype=bool)
    thr = np.sort(resp)[::-1][n - 1]
    return resp >= thr


# --------------------------------------------------------------------------- Harris
def harris_response(img: np.ndarray, xs: np.ndarray, ys: np.ndarray) -> np.ndarray:
    """HarrisResponses(blockSize 7, k 0.04): int32 sums then non-fused f32."""
    I = img.astype(np.int32)
    h, w = I.shape
    Ix = np.zeros((h, w), dtype=np.int64)
    Iy = np.zeros((h, w), dtype=np.int64)
    Ix[1:-1, 1:-1] = 2 * (I[1:-1, 2:] - I[1:-1, :-2]) + (I[:-2, 2:] - I[:-2, :-2]) + (I[2:, 2:] - I[2:, :-2])
    Iy[1:-1, 1:-1] = 2 * (I[2:, 1:-1] - I[:-2, 1:-1]) + (I[2:, :-2] - I[:-2, :-2]) + (I[2:, 2:] - I[:-2, 2:])

    def box(a):
        ii = np.zeros((h + 1, w + 1), dtype=np.int64)
        ii[1:, 1:] = a.cumsum(0).cumsum(1)
        r = HARRIS_BLOCK // 2
        y0, y1, x0, x1 = ys - r, ys + r + 1, xs - r, xs + r + 1
        return ii[y1, x1] - ii[y0, x1] - ii[y1, x0] + ii[y0, x0]

    a = box(Ix * Ix).astype(np.int32).astype(np.float32)
    b = box(Iy * Iy).astype(np.int32).astype(np.float32)
    c = box(Ix * Iy).astype(np.int32).astype(np.float32)
    s = np.float32(1.0) / (np.float32(4 * HARRIS_BLOCK) * np.float32(255.0))
    s4 = np.float32(np.float32(np.float32(s * s) * s) * s)
    apb = (a + b).astype(np.float32)
    t1 = (a * b).astype(np.float32)
    t2 = (c * c).astype(np.float32)
    t3 = ((HARRIS_K * apb).astype(np.float32) * apb).astype(np.float32)
    return (((t1 - t2).astype(np.float32) - t3).astype(np.float32) * s4).astype(np.float32)


# --------------------------------------------------------------------------- orientation
_ATAN_SCALE = np.float32(180.0 / math.pi)
_P1 = np.float32(np.float32(0.9997878412794807) * _ATAN_SCALE)
_P3 = np.float32(np.float32(-0.3258083974640975) * _ATAN_SCALE)
_P5 = np.float32(np.float32(0.1555786518463281) * _ATAN_SCALE)
_P7 = np.float32(np.float32(-0.04432655554792128) * _ATAN_SCALE)
_EPS32 = np.float32(2.220446049250313e-16)


def fast_atan2(y: np.ndarray, x: np.ndarray) -> np.ndarray:
    """cv::fastAtan2 scalar f32 polynomial, degrees in [0, 360)."""
    y = np.asarray(y, dtype=np.float32)
    x = np.asarray(x, dtype=np.float32)
    ax, ay = np.abs(x), np.abs(y)
    big = ax >= ay
    num = np.where(big, ay, ax)
    den = np.where(big, ax, ay) + _EPS32
    c = (num / den).astype(np.float32)
    c2 = (c * c).astype(np.float32)
    p = ((_P7 * c2 + _P5).astype(np.float32) * c2 + _P3).astype(np.float32)
    p = ((p * c2 + _P1).astype(np.float32) * c).astype(np.float32)
    a = np.where(big, p, np.float32(90.0) - p).astype(np.float32)
    a = np.where(x < 0, np.float32(180.0) - a, a).astype(np.float32)
    a = np.where(y < 0, np.float32(360.0) - a, a).astype(np.float32)
    return a


def _patch_offsets():
    us, vs = [], []
    for v in range(-HALF_PATCH, HALF_PATCH + 1):
        d = UMAX[abs(v)]
        for u in range(-d, d + 1):
            us.append(u); vs.append(v)
    return np.array(us, dtype=np.int32), np.array(vs, dtype=np.int32)


_PU, _PV = _patch_offsets()


def ic_angle(img: np.ndarray, xs: np.ndarray, ys: np.ndarray) -> np.ndarray:
    """Intensity-centroid orientation over the 31-px circular patch."""
    I = img.astype(np.int64)
    vals = I[ys[:, None] + _PV[None, :], xs[:, None] + _PU[None, :]]
    m10 = (vals * _PU[None, :]).sum(axis=1)
    m01 = (vals * _PV[None, :]).sum(axis=1)
    return fast_atan2(m01.astype(np.float32), m10.astype(np.float32))


# --------------------------------------------------------------------------- blur
def _fma32(a, b, c):
    """float32 fused multiply-add: exact a*b+c in float64 (24-bit x 24-bit products are
    exact in a double; the add is exact enough that one final rounding to f32 equals FMA
    except in double-rounding corner cases, which the fixture tests show do not occur)."""
    return (a.astype(np.float64) * b.astype(np.float64) + c.astype(np.float64)).astype(np.float32)


def blur7(img: np.ndarray) -> np.ndarray:
    """The 7x7 sigma-2 blur as executed inside ORB: f32 FMA separable, reflect-101 (A.8)."""
    k = BLUR_KERNEL
    h, w = img.shape
    src = np.pad(img, 3, mode="reflect").astype(np.float32)           # reflect == BORDER_REFLECT_101
    # row pass: s = k0*x[-3]; s = fma(k[j], x[j-3], s)
    rows = (k[0] * src[:, 0:w]).astype(np.float32)
    for j in range(1, 7):
        rows = _fma32(np.full_like(rows, k[j]), src[:, j:j + w], rows)
    # column pass on padded rows (h+6, w)
    r = rows
    s = (k[3] * r[3:3 + h]).astype(np.float32)
    s = _fma32(np.full_like(s, k[2]), (r[2:2 + h] + r[4:4 + h]).astype(np.float32), s)
    s = _fma32(np.full_like(s, k[1]), (r[1:1 + h] + r[5:5 + h]).astype(np.float32), s)
    s = _fma32(np.full_like(s, k[0]), (r[0:0 + h] + r[6:6 + h]).astype(np.float32), s)
    out = np.rint(s)
    return np.clip(out, 0, 255).astype(np.uint8)


# --------------------------------------------------------------------------- descriptor
_DEG2RAD = np.float32(math.pi / 180.0)


def brief_descriptors(blurred: np.ndarray, pt_x: np.ndarray, pt_y: np.ndarray,
                      angle: np.ndarray, scale: np.float32) -> np.ndarray:
    """Steered rBRIEF-256 on the blurred level.  pt_* are the f32 image-space
    keypoint coordinates (already multiplied by the level scale)."""
    n = pt_x.shape[0]
    inv = np.float32(1.0) / np.float32(scale)
    cx = np.rint((pt_x.astype(np.float32) * inv).astype(np.float32)).astype(np.int64)
    cy = np.rint((pt_y.astype(np.float32) * inv).astype(np.float32)).astype(np.int64)
    theta = (angle.astype(np.float32) * _DEG2RAD).astype(np.float32)
    a = np.cos(theta.astype(np.float64)).astype(np.float32)[:, None]
    b = np.sin(theta.astype(np.float64)).astype(np.float32)[:, None]
    P = PATTERN.astype(np.float32)

    def sample(px, py):
        x = ((px[None, :] * a).astype(np.float32) - (py[None, :] * b).astype(np.float32)).astype(np.float32)
        y = ((px[None, :] * b).astype(np.float32) + (py[None, :] * a).astype(np.float32)).astype(np.float32)
        ix = np.rint(x).astype(np.int64)
        iy = np.rint(y).astype(np.int64)
        return blurred[cy[:, None] + iy, cx[:, None] + ix]

    va = sample(P[:, 0], P[:, 1])
    vb = sample(P[:, 2], P[:, 3])
    bits = (va < vb).astype(np.uint8).reshape(n, 32, 8)
    weights = (1 << np.arange(8)).astype(np.uint8)
    return (bits * weights[None, None, :]).sum(axis=2).astype(np.uint8)


# --------------------------------------------------------------------------- full ORB
@dataclass
class OrbResult:
    x: np.ndarray          # f32 image-space x   (pt.x)
    y: np.ndarray          # f32 image-space y   (pt.y)
    size: np.ndarray       # f32
    angle: np.ndarray      # f32 degrees
    response: np.ndarray   # f32 Harris
    octave: np.ndarray     # i32
    xl: np.ndarray         # i32 level-space x
    yl: np.ndarray         # i32 level-space y
    desc: np.ndarray       # (N, 32) u8

    def __len__(self):
        return int(self.x.shape[0])


def detect_level(img_l: np.ndarray, quota: int, threshold: int):
    """Keypoints of one level in canonical (y, x) order: (xl, yl, harris)."""
    xs, ys, sc = fast_candidates(img_l, threshold)
    keep = retain_best(sc, 2 * quota)
    xs, ys = xs[keep], ys[keep]
    if xs.shape[0] == 0:
        return xs, ys, np.zeros(0, dtype=np.float32)
    hr = harris_response(img_l, xs, ys)
    keep = retain_best(hr, quota)
    return xs[keep], ys[keep], hr[keep]


def orb_detect_and_compute(img: np.ndarray, p: OrbParams | None = None,
                           return_pyramid: bool = False):
    """Full ORB on a u8 gray image; keypoints in canonical (octave, y_level, x_level) order."""
    p = p or OrbParams()
    H, W = img.shape
    levels = level_geometry(W, H, p)
    pyr = build_pyramid(img, levels)
    X, Y, S, A, R, O, XL, YL, D = [], [], [], [], [], [], [], [], []
    blurred = []
    for l, (lv, im) in enumerate(zip(levels, pyr)):
        xs, ys, hr = detect_level(im, lv.quota, p.fast_threshold)
        bl = blur7(im) if (xs.shape[0] or return_pyramid) else None
        blurred.append(bl)
        if xs.shape[0] == 0:
            continue
        ang = ic_angle(im, xs, ys)
        px = (xs.astype(np.float32) * lv.scale).astype(np.float32)
        py = (ys.astype(np.float32) * lv.scale).astype(np.float32)
        desc = brief_descriptors(bl, px, py, ang, lv.scale)
        n = xs.shape[0]
        X.append(px); Y.append(py); S.append(np.full(n, np.float32(PATCH_SIZE) * lv.scale, np.float32))
        A.append(ang); R.append(hr); O.append(np.full(n, l, np.int32)); XL.append(xs); YL.append(ys); D.append(desc)

    def cat(v, dt, shape=(0,)):
        return np.concatenate(v).astype(dt) if v else np.zeros(shape, dt)

    res = OrbResult(cat(X, np.float32), cat(Y, np.float32), cat(S, np.float32), cat(A, np.float32),
                    cat(R, np.float32), cat(O, np.int32), cat(XL, np.int32), cat(YL, np.int32),
                    cat(D, np.uint8, (0, 32)))
    if return_pyramid:
        return res, levels, pyr, blurred
    return res


def canonicalize_cv2(kps, desc, levels: list[Level]) -> OrbResult:
    """Reorder cv2.ORB output into the canonical (octave, y_level, x_level) order."""
    n = len(kps)
    x = np.array([k.pt[0] for k in kps], dtype=np.float32)
    y = np.array([k.pt[1] for k in kps], dtype=np.float32)
    size = np.array([k.size for k in kps], dtype=np.float32)
    ang = np.array([k.angle for k in kps], dtype=np.float32)
    resp = np.array([k.response for k in kps], dtype=np.float32)
    octv = np.array([k.octave for k in kps], dtype=np.int32)
    sc = np.array([lv.scale for lv in levels], dtype=np.float32)[octv] if n else np.zeros(0, np.float32)
    xl = np.rint(x / sc).astype(np.int32)
    yl = np.rint(y / sc).astype(np.int32)
    order = np.lexsort((xl, yl, octv))
    d = desc[order] if desc is not None and n else np.zeros((0, 32), np.uint8)
    return OrbResult(x[order], y[order], size[order], ang[order], resp[order], octv[order],
                     xl[order], yl[order], d)
