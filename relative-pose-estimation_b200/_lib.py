"""ctypes binding of librpe_b200.so (the C ABI declared in include/rpe.h).

The library is the product; this module only loads it.  There is no fallback: if the shared
library is missing it is built in-tree with nvcc, and if that fails, or no B200 is present at
rpe_create time, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

from . import build as _build

HERE = os.path.dirname(os.path.abspath(__file__))

RPE_T_NAMES = ["h2d", "pyramid", "fast", "select", "blur", "describe", "match", "match_sort",
               "ransac", "pose", "d2h"]
RPE_T_COUNT = len(RPE_T_NAMES)

STATUS_OK, STATUS_NO_DESCRIPTORS, STATUS_INSUFFICIENT_MATCHES, STATUS_E_FAILED = 0, 1, 2, 3
STATUS_FLAG_OVERFLOW = 0x100
MATCH_CROSSCHECK, MATCH_RATIO = 0, 1
EULER_YUP, EULER_ZYX = 0, 1
GRAY_CVTCOLOR, GRAY_IMREAD_PNG = 0, 1

(STAGE_PYRAMID_LEVEL, STAGE_BLURRED_LEVEL, STAGE_KP_X, STAGE_KP_Y, STAGE_KP_SIZE, STAGE_KP_ANGLE,
 STAGE_KP_RESPONSE, STAGE_KP_OCTAVE, STAGE_DESCRIPTORS, STAGE_MATCHES, STAGE_PTS1, STAGE_PTS2, STAGE_E,
 STAGE_RANSAC_MASK, STAGE_POSE_MASK) = range(15)


class RpeConfig(C.Structure):
    _fields_ = [
        ("max_width", C.c_int32), ("max_height", C.c_int32), ("max_pairs", C.c_int32),
        ("nfeatures", C.c_int32), ("nlevels", C.c_int32), ("scale_factor", C.c_float),
        ("fast_threshold", C.c_int32), ("match_mode", C.c_int32), ("ratio", C.c_double),
        ("max_matches", C.c_int32), ("ransac_prob", C.c_double), ("ransac_threshold", C.c_double),
        ("ransac_max_iters", C.c_int32), ("euler_convention", C.c_int32),
    ]


# every symbol include/rpe.h declares: (name, restype, argtypes)
_P = C.c_void_p
_SIGNATURES = [
    ("rpe_abi_version", C.c_int, []),
    ("rpe_default_config", None, [C.POINTER(RpeConfig)]),
    ("rpe_last_error", C.c_char_p, []),
    ("rpe_create", C.c_int, [C.POINTER(RpeConfig), C.c_int, C.POINTER(_P)]),
    ("rpe_destroy", None, [_P]),
    ("rpe_set_stream", C.c_int, [_P, _P]),
    ("rpe_set_pipeline", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("rpe_set_input_format", C.c_int, [_P, C.c_int32, C.c_int32]),
    ("rpe_get_capacity", C.c_int, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    ("rpe_estimate_batch", C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_submit_batch", C.c_int, [_P, _P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_submit_batch_device", C.c_int, [_P, _P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_wait_batch", C.c_int, [_P]),
    ("rpe_estimate_batch_device", C.c_int, [_P, _P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_estimate_sequence", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_submit_sequence", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_estimate_sequence_device", C.c_int, [_P, _P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_submit_sequence_device", C.c_int, [_P, _P, C.c_size_t, C.c_int32, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P]),
    ("rpe_orb_detect", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_size_t, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    ("rpe_bgr_to_gray", C.c_int, [_P, _P, C.c_int32, C.c_int32, C.c_size_t, C.c_int32, _P]),
    ("rpe_match", C.c_int, [_P, _P, C.c_int32, _P, C.c_int32, _P, _P]),
    ("rpe_find_essential", C.c_int, [_P, _P, _P, C.c_int32, _P, _P, _P, _P]),
    ("rpe_ransac_trace", C.c_int, [_P, C.c_int32, _P, _P, _P, _P, _P]),
    ("rpe_recover_pose", C.c_int, [_P, _P, _P, _P, C.c_int32, _P, _P, _P, _P, _P, _P]),
    ("rpe_estimate_from_matches", C.c_int, [_P, _P, _P, _P, C.c_int32, C.c_int32, _P, _P, _P, _P, _P, _P]),
    ("rpe_evaluate_sequence", C.c_int, [_P, C.c_int32, _P, _P, _P, _P, _P, C.c_int32, _P, _P]),
    ("rpe_summarize_errors", C.c_int, [_P, C.c_int32, _P, _P]),
    ("rpe_evaluate_sequence_noalloc", C.c_int, [C.c_int32, _P, _P, _P, _P, _P, C.c_int32, _P, _P]),
    ("rpe_get_stage", C.c_int, [_P, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _P, C.POINTER(C.c_size_t)]),
    ("rpe_set_profiling", C.c_int, [_P, C.c_int32]),
    ("rpe_get_stage_times", C.c_int, [_P, _P, _P]),
]
EXPORTED_SYMBOLS = [s[0] for s in _SIGNATURES]

_lib = None


def library_path() -> str:
    return _build.LIB


def load(build_if_missing: bool = True) -> C.CDLL:
    """Load librpe_b200.so (building it in-tree if absent).  Raises on failure — never falls back."""
    global _lib
    if _lib is not None:
        return _lib
    path = library_path()
    if build_if_missing and _build.needs_build():
        _build.build()
    if not os.path.exists(path):
        raise RuntimeError(f"{path} is missing and could not be built; the CUDA library is required "
                           "(there is no CPU fallback)")
    lib = C.CDLL(path)
    for name, res, args in _SIGNATURES:
        fn = getattr(lib, name)          # AttributeError if the ABI drifted
        fn.restype = res
        fn.argtypes = args
    if lib.rpe_abi_version() != 1:
        raise RuntimeError("librpe_b200.so ABI version mismatch")
    _lib = lib
    return lib


class RpeError(RuntimeError):
    pass


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = load().rpe_last_error()
        raise RpeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
