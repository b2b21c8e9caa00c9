"""ctypes wrapper over tests/hostcheck/libhostcheck.so (TEST-ONLY host instantiation of the
kernels' __host__ __device__ arithmetic; built on demand with g++)."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(_HERE)
_SRC = os.path.join(_HERE, "hostcheck", "hostcheck.cpp")
_SO = os.path.join(_HERE, "hostcheck", "libhostcheck.so")
_CSRC = os.path.join(_ROOT, "relative-pose-estimation_b200", "csrc")


def build(force=False):
    deps = [_SRC] + [os.path.join(_CSRC, f) for f in os.listdir(_CSRC) if f.endswith((".cuh", ".inc"))]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(d) > os.path.getmtime(_SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-shared", "-fPIC", "-x", "c++",
                               "-I", _CSRC, _SRC, "-o", _SO])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def gray(bgr, mode):
    bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
    out = np.zeros(bgr.shape[:-1], np.uint8)
    lib().hc_gray(_p(bgr), out.size, int(mode), _p(out))
    return out


def resize(src, dw, dh, xo, xc, yo, yc):
    src = np.ascontiguousarray(src)
    out = np.zeros((dh, dw), np.uint8)
    a = [np.ascontiguousarray(v, dtype=np.int32) for v in (xo, xc, yo, yc)]
    lib().hc_resize(_p(src), src.shape[1], src.shape[0], _p(out), dw, dh, *[_p(v) for v in a])
    return out


def fast_scores(img, t):
    img = np.ascontiguousarray(img)
    out = np.zeros(img.shape, np.int32)
    lib().hc_fast_scores(_p(img), img.shape[1], img.shape[0], int(t), _p(out))
    return out


def fast_scores_scalar(img, t):
    img = np.ascontiguousarray(img)
    out = np.zeros(img.shape, np.int32)
    lib().hc_fast_scores_scalar(_p(img), img.shape[1], img.shape[0], int(t), _p(out))
    return out


def fast_compass(img, t):
    img = np.ascontiguousarray(img)
    out = np.zeros(img.shape, np.uint8)
    lib().hc_fast_compass(_p(img), img.shape[1], img.shape[0], int(t), _p(out))
    return out


def harris(img, xs, ys):
    img = np.ascontiguousarray(img)
    xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32)
    out = np.zeros(xs.shape[0], np.float32)
    lib().hc_harris(_p(img), img.shape[1], _p(xs), _p(ys), xs.shape[0], _p(out))
    return out


def angle(img, xs, ys):
    img = np.ascontiguousarray(img)
    xs = np.ascontiguousarray(xs, np.int32); ys = np.ascontiguousarray(ys, np.int32)
    out = np.zeros(xs.shape[0], np.float32)
    lib().hc_angle(_p(img), img.shape[1], _p(xs), _p(ys), xs.shape[0], _p(out))
    return out


def atan2(y, x):
    y = np.ascontiguousarray(y, np.float32); x = np.ascontiguousarray(x, np.float32)
    out = np.zeros(x.shape[0], np.float32)
    lib().hc_atan2(_p(y), _p(x), x.shape[0], _p(out))
    return out


def blur(img):
    img = np.ascontiguousarray(img)
    out = np.zeros_like(img)
    lib().hc_blur(_p(img), img.shape[1], img.shape[0], _p(out))
    return out


def brief(blurred, cx, cy, ang):
    blurred = np.ascontiguousarray(blurred)
    cx = np.ascontiguousarray(cx, np.int32); cy = np.ascontiguousarray(cy, np.int32)
    ang = np.ascontiguousarray(ang, np.float32)
    out = np.zeros((cx.shape[0], 32), np.uint8)
    lib().hc_brief(_p(blurred), blurred.shape[1], _p(cx), _p(cy), _p(ang), cx.shape[0], _p(out))
    return out


def subsets(count, n):
    out = np.zeros((n, 5), np.int32)
    lib().hc_subsets(int(count), int(n), _p(out))
    return out


def update_iters(p, ep, mp, mx):
    f = lib().hc_update_iters
    f.argtypes = [C.c_double, C.c_double, C.c_int, C.c_int]
    return f(p, ep, mp, mx)


def five_point(x1, x2):
    x1 = np.ascontiguousarray(x1, np.float64); x2 = np.ascontiguousarray(x2, np.float64)
    E = np.zeros((10, 9), np.float64)
    n = lib().hc_five_point(_p(x1), _p(x2), _p(E))
    return [E[i].reshape(3, 3).copy() for i in range(n)]


def sampson(E, x1, x2):
    E = np.ascontiguousarray(E, np.float64); x1 = np.ascontiguousarray(x1, np.float64)
    x2 = np.ascontiguousarray(x2, np.float64)
    out = np.zeros(x1.shape[0], np.float32)
    lib().hc_sampson(_p(E), _p(x1), _p(x2), x1.shape[0], _p(out))
    return out


def sampson_inlier(E, x1, x2, t2):
    E = np.ascontiguousarray(E, np.float64); x1 = np.ascontiguousarray(x1, np.float64)
    x2 = np.ascontiguousarray(x2, np.float64)
    out = np.zeros(x1.shape[0], np.uint8)
    lib().hc_sampson_inlier(_p(E), _p(x1), _p(x2), x1.shape[0], C.c_float(float(t2)), _p(out))
    return out.astype(bool)


def decompose(E):
    E = np.ascontiguousarray(E, np.float64)
    R1 = np.zeros((3, 3)); R2 = np.zeros((3, 3)); t = np.zeros(3)
    lib().hc_decompose(_p(E), _p(R1), _p(R2), _p(t))
    return R1, R2, t


def cheirality(R, t, x1, x2, dist=50.0):
    R = np.ascontiguousarray(R, np.float64); t = np.ascontiguousarray(t, np.float64).ravel()
    x1 = np.ascontiguousarray(x1, np.float64); x2 = np.ascontiguousarray(x2, np.float64)
    ok = np.zeros(x1.shape[0], np.uint8)
    f = lib().hc_cheirality
    f.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_double, C.c_void_p]
    f(_p(R), _p(t), _p(x1), _p(x2), x1.shape[0], dist, _p(ok))
    return ok.astype(bool)


def euler(R, conv):
    R = np.ascontiguousarray(R, np.float64)
    out = np.zeros(3)
    lib().hc_euler(_p(R), int(conv), _p(out))
    return tuple(out)


def solve_poly10(c):
    c = np.ascontiguousarray(c, np.float64)
    out = np.zeros((10, 2))
    lib().hc_solve_poly10(_p(c), _p(out))
    return out[:, 0] + 1j * out[:, 1]


def hamming(q, t, mode=0):
    """all-pairs 256-bit Hamming distances through the kernel's carry-save arithmetic (match_math.cuh)"""
    q = np.ascontiguousarray(q, np.uint8); t = np.ascontiguousarray(t, np.uint8)
    out = np.zeros((q.shape[0], t.shape[0]), np.int32)
    f = lib().hc_hamming
    f.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    f.restype = None
    f(_p(q), q.shape[0], _p(t), t.shape[0], mode, _p(out))
    return out
