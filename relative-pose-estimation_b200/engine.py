"""Thin Python wrapper over one rpe_handle (one per GPU).  All arithmetic happens in
librpe_b200.so; this file only marshals numpy buffers through the C ABI."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass

import numpy as np

from . import _lib as L


@dataclass
class EngineConfig:
    """Mirrors rpe_config; defaults are the reference's values
    (reference: src/core/pose_estimator.py:19-32, :85-91, :131, :150, :522-527)."""
    max_width: int = 640
    max_height: int = 480
    max_pairs: int = 1
    nfeatures: int = 4000
    nlevels: int = 12
    scale_factor: float = 1.1
    fast_threshold: int = 15
    match_mode: str = "crosscheck"        # "crosscheck" (reference) | "ratio" (BASELINE configs 3-5)
    ratio: float = 0.75
    max_matches: int | None = 500
    ransac_prob: float = 0.999
    ransac_threshold: float = 1.0
    ransac_max_iters: int = 1000
    euler_convention: str = "yup"

    def to_c(self) -> L.RpeConfig:
        if self.match_mode not in ("crosscheck", "ratio"):
            raise ValueError(f"Unknown match mode: {self.match_mode}")
        if self.euler_convention not in ("yup", "zyx"):
            raise ValueError(f"Unknown euler convention: {self.euler_convention}")
        return L.RpeConfig(
            self.max_width, self.max_height, self.max_pairs, self.nfeatures, self.nlevels,
            self.scale_factor, self.fast_threshold,
            L.MATCH_CROSSCHECK if self.match_mode == "crosscheck" else L.MATCH_RATIO, float(self.ratio),
            0 if self.max_matches is None else int(self.max_matches), self.ransac_prob,
            self.ransac_threshold, self.ransac_max_iters,
            L.EULER_YUP if self.euler_convention == "yup" else L.EULER_ZYX)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Engine:
    def __init__(self, cfg: EngineConfig | None = None, device: int = 0):
        self.cfg = cfg or EngineConfig()
        self.device = device
        self._lib = L.load()
        self._h = C.c_void_p()
        ccfg = self.cfg.to_c()
        L.check(self._lib.rpe_create(C.byref(ccfg), device, C.byref(self._h)), "rpe_create")
        kc, mc = C.c_int32(), C.c_int32()
        L.check(self._lib.rpe_get_capacity(self._h, C.byref(kc), C.byref(mc)), "rpe_get_capacity")
        self.kp_cap, self.match_cap = kc.value, mc.value
        self._fmt = (1, L.GRAY_CVTCOLOR)
        self._pending = []                    # submitted, not yet waited-for batches (oldest first; at most 2)
        self.gray_mode = L.GRAY_CVTCOLOR      # arithmetic used when BGR images are passed (see set_input_format)

    def close(self):
        if getattr(self, "_h", None) and self._h.value:
            self._lib.rpe_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ------------------------------------------------------------------ batch path
    def set_stream(self, cuda_stream_ptr: int | None):
        L.check(self._lib.rpe_set_stream(self._h, C.c_void_p(cuda_stream_ptr or 0)), "rpe_set_stream")

    def set_pipeline(self, chunk_pairs: int, n_streams: int):
        """Sub-batch pipeline of the batch calls (see include/rpe.h: rpe_set_pipeline)."""
        L.check(self._lib.rpe_set_pipeline(self._h, int(chunk_pairs), int(n_streams)), "rpe_set_pipeline")

    def set_input_format(self, channels: int, gray_mode: int | None = None):
        """1 = gray images (default); 3 = interleaved BGR as cv2.imread returns it, converted on the GPU with
        cv2.cvtColor's arithmetic (GRAY_CVTCOLOR) or IMREAD_GRAYSCALE's for PNG (GRAY_IMREAD_PNG)."""
        fmt = (int(channels), self.gray_mode if gray_mode is None else int(gray_mode))
        if fmt != self._fmt:
            L.check(self._lib.rpe_set_input_format(self._h, fmt[0], fmt[1]), "rpe_set_input_format")
            self._fmt = fmt

    def _prep(self, imgs):
        """-> (list of C-contiguous uint8 arrays, H, W, channels); sets the input format to match."""
        arrs = [np.ascontiguousarray(i, dtype=np.uint8) for i in imgs]
        a0 = arrs[0]
        if a0.ndim == 3 and a0.shape[2] == 3:
            ch = 3
        elif a0.ndim == 2:
            ch = 1
        else:
            raise ValueError("images must be HxW uint8 gray or HxWx3 uint8 BGR")
        for a in arrs:
            if a.shape != a0.shape:
                raise ValueError("all images of a batch must have one size and format")
        self.set_input_format(ch)
        return arrs, a0.shape[0], a0.shape[1], ch

    def bgr_to_gray(self, bgr, gray_mode: int = L.GRAY_CVTCOLOR):
        """Stage entry: the GPU gray conversion alone (rpe_bgr_to_gray)."""
        bgr = np.ascontiguousarray(bgr, dtype=np.uint8)
        H, W, _ = bgr.shape
        out = np.zeros((H, W), np.uint8)
        L.check(self._lib.rpe_bgr_to_gray(self._h, _ptr(bgr), H, W, 3 * W, int(gray_mode), _ptr(out)), "rpe_bgr_to_gray")
        return out

    def set_profiling(self, on: bool):
        L.check(self._lib.rpe_set_profiling(self._h, int(on)), "rpe_set_profiling")

    def stage_times(self):
        ms = np.zeros(L.RPE_T_COUNT, np.float32)
        ln = np.zeros(L.RPE_T_COUNT, np.int32)
        L.check(self._lib.rpe_get_stage_times(self._h, _ptr(ms), _ptr(ln)), "rpe_get_stage_times")
        return dict(zip(L.RPE_T_NAMES, ms.tolist())), dict(zip(L.RPE_T_NAMES, ln.tolist()))

    @staticmethod
    def _outputs(B):
        return (np.zeros((B, 3, 3)), np.zeros((B, 3, 1)), np.zeros((B, 3)), np.zeros(B, np.int32),
                np.zeros((B, 4), np.int32))

    def estimate_batch(self, imgs1, imgs2, K):
        """imgs1 / imgs2: sequences (or (B,H,W) arrays) of HxW uint8 gray host images.
        Returns dict(R (B,3,3), t (B,3,1), euler (B,3) = yaw,pitch,roll deg, status (B,), info (B,4))."""
        B = len(imgs1)
        if B != len(imgs2) or B == 0:
            raise ValueError("imgs1 and imgs2 must be non-empty and of equal length")
        arrs, H, W, ch = self._prep(list(imgs1) + list(imgs2))
        a1, a2 = arrs[:B], arrs[B:]
        p1 = (C.c_void_p * B)(*[a.ctypes.data for a in a1])
        p2 = (C.c_void_p * B)(*[a.ctypes.data for a in a2])
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_estimate_batch(self._h, p1, p2, B, H, W, W * ch, _ptr(K), _ptr(R), _ptr(t), _ptr(e),
                                             _ptr(st), _ptr(info)), "rpe_estimate_batch")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    def estimate_batch_interleaved(self, frames: np.ndarray, K):
        """frames: (2B, H, W) uint8 host array, view 1 of pair i at 2i and view 2 at 2i+1 (one
        H2D copy when the array is contiguous and W is a multiple of 16; pin it for async copies)."""
        B = frames.shape[0] // 2
        H, W = frames.shape[1:3]
        self.set_input_format(3 if frames.ndim == 4 else 1)
        base = frames.ctypes.data
        step = frames.strides[0]
        p1 = (C.c_void_p * B)(*[base + (2 * i) * step for i in range(B)])
        p2 = (C.c_void_p * B)(*[base + (2 * i + 1) * step for i in range(B)])
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_estimate_batch(self._h, p1, p2, B, H, W, frames.strides[1], _ptr(K), _ptr(R), _ptr(t),
                                             _ptr(e), _ptr(st), _ptr(info)), "rpe_estimate_batch")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    def submit_interleaved(self, frames: np.ndarray, K):
        """Asynchronous estimate_batch_interleaved (rpe_submit_batch): returns at once; ``wait()``
        returns the result dict.  ``frames`` must stay alive (and unchanged) until then."""
        B = frames.shape[0] // 2
        H, W = frames.shape[1:3]
        self.set_input_format(3 if frames.ndim == 4 else 1)
        base = frames.ctypes.data
        step = frames.strides[0]
        p1 = (C.c_void_p * B)(*[base + (2 * i) * step for i in range(B)])
        p2 = (C.c_void_p * B)(*[base + (2 * i + 1) * step for i in range(B)])
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_submit_batch(self._h, p1, p2, B, H, W, frames.strides[1], _ptr(K), _ptr(R), _ptr(t),
                                           _ptr(e), _ptr(st), _ptr(info)), "rpe_submit_batch")
        self._pending.append(({"R": R, "t": t, "euler": e, "status": st, "info": info}, frames, K))
        return self

    def submit_device(self, dev_ptr: int, img_stride: int, B: int, H: int, W: int, pitch: int, K):
        """Asynchronous estimate_batch_device (rpe_submit_batch_device); ``wait()`` returns the results."""
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_submit_batch_device(self._h, C.c_void_p(dev_ptr), img_stride, B, H, W, pitch, _ptr(K),
                                                  _ptr(R), _ptr(t), _ptr(e), _ptr(st), _ptr(info)),
                "rpe_submit_batch_device")
        self._pending.append(({"R": R, "t": t, "euler": e, "status": st, "info": info}, None, K))
        return self

    def wait(self):
        """Block until the OLDEST submitted batch is done (rpe_wait_batch) and return its result dict.
        Two batches may be pending per engine: the second one's upload and ORB / matcher kernels run under
        the first one's RANSAC / recoverPose tail."""
        L.check(self._lib.rpe_wait_batch(self._h), "rpe_wait_batch")
        return self._pending.pop(0)[0]

    def estimate_batch_device(self, dev_ptr: int, img_stride: int, B: int, H: int, W: int, pitch: int, K):
        """Inputs resident in device memory (e.g. a torch uint8 CUDA tensor's data_ptr())."""
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_estimate_batch_device(self._h, C.c_void_p(dev_ptr), img_stride, B, H, W, pitch,
                                                    _ptr(K), _ptr(R), _ptr(t), _ptr(e), _ptr(st), _ptr(info)),
                "rpe_estimate_batch_device")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    # ------------------------------------------------------------------ consecutive frames (feature cache)
    def _frame_ptrs(self, frames):
        if (isinstance(frames, np.ndarray) and frames.dtype == np.uint8 and frames.ndim in (3, 4)
                and frames[0].flags.c_contiguous and (frames.ndim == 3 or frames.shape[3] == 3)):
            F, H, W = frames.shape[:3]
            self.set_input_format(3 if frames.ndim == 4 else 1)
            base, step = frames.ctypes.data, frames.strides[0]
            return (C.c_void_p * F)(*[base + i * step for i in range(F)]), F, H, W, frames.strides[1], frames
        if len(frames) < 2:
            raise ValueError("a sequence needs at least 2 frames")
        arrs, H, W, ch = self._prep(frames)
        return (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs]), len(arrs), H, W, W * ch, arrs

    def estimate_sequence(self, frames, K):
        """frames: (F, H, W) uint8 host array or list of F gray images -> results of the F-1 consecutive
        pairs (k, k+1); ORB runs once per frame (rpe_estimate_sequence).  Same dict as estimate_batch."""
        ptrs, F, H, W, pitch, keep = self._frame_ptrs(frames)
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(F - 1)
        L.check(self._lib.rpe_estimate_sequence(self._h, ptrs, F, H, W, pitch, _ptr(K), _ptr(R), _ptr(t), _ptr(e),
                                                _ptr(st), _ptr(info)), "rpe_estimate_sequence")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    def submit_sequence(self, frames, K):
        """Asynchronous estimate_sequence (rpe_submit_sequence); ``wait()`` returns the result dict."""
        ptrs, F, H, W, pitch, keep = self._frame_ptrs(frames)
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(F - 1)
        L.check(self._lib.rpe_submit_sequence(self._h, ptrs, F, H, W, pitch, _ptr(K), _ptr(R), _ptr(t), _ptr(e),
                                              _ptr(st), _ptr(info)), "rpe_submit_sequence")
        self._pending.append(({"R": R, "t": t, "euler": e, "status": st, "info": info}, keep, K))
        return self

    def submit_sequence_device(self, dev_ptr: int, img_stride: int, F: int, H: int, W: int, pitch: int, K):
        """Asynchronous sequence call on device-resident frames (rpe_submit_sequence_device)."""
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(F - 1)
        L.check(self._lib.rpe_submit_sequence_device(self._h, C.c_void_p(dev_ptr), img_stride, F, H, W, pitch,
                                                     _ptr(K), _ptr(R), _ptr(t), _ptr(e), _ptr(st), _ptr(info)),
                "rpe_submit_sequence_device")
        self._pending.append(({"R": R, "t": t, "euler": e, "status": st, "info": info}, None, K))
        return self

    def estimate_sequence_device(self, dev_ptr: int, img_stride: int, F: int, H: int, W: int, pitch: int, K):
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(F - 1)
        L.check(self._lib.rpe_estimate_sequence_device(self._h, C.c_void_p(dev_ptr), img_stride, F, H, W, pitch,
                                                       _ptr(K), _ptr(R), _ptr(t), _ptr(e), _ptr(st), _ptr(info)),
                "rpe_estimate_sequence_device")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    # ------------------------------------------------------------------ stage entry points
    def orb_detect(self, img):
        img = np.ascontiguousarray(img, dtype=np.uint8)
        H, W = img.shape
        n = self.kp_cap
        x = np.zeros(n, np.float32); y = np.zeros(n, np.float32); size = np.zeros(n, np.float32)
        ang = np.zeros(n, np.float32); resp = np.zeros(n, np.float32); octv = np.zeros(n, np.int32)
        desc = np.zeros((n, 32), np.uint8)
        cnt, fl = C.c_int32(), C.c_int32()
        L.check(self._lib.rpe_orb_detect(self._h, _ptr(img), H, W, W, _ptr(x), _ptr(y), _ptr(size), _ptr(ang),
                                         _ptr(resp), _ptr(octv), _ptr(desc), C.byref(cnt), C.byref(fl)),
                "rpe_orb_detect")
        k = cnt.value
        return {"x": x[:k], "y": y[:k], "size": size[:k], "angle": ang[:k], "response": resp[:k],
                "octave": octv[:k], "desc": desc[:k], "flags": fl.value}

    def match(self, d1, d2):
        d1 = np.ascontiguousarray(d1, dtype=np.uint8).reshape(-1, 32)
        d2 = np.ascontiguousarray(d2, dtype=np.uint8).reshape(-1, 32)
        out = np.zeros((self.match_cap, 3), np.int32)
        m = C.c_int32()
        L.check(self._lib.rpe_match(self._h, _ptr(d1), d1.shape[0], _ptr(d2), d2.shape[0], _ptr(out), C.byref(m)),
                "rpe_match")
        return out[:m.value].copy()

    def find_essential(self, pts1, pts2, K):
        pts1 = np.ascontiguousarray(pts1, dtype=np.float32).reshape(-1, 2)
        pts2 = np.ascontiguousarray(pts2, dtype=np.float32).reshape(-1, 2)
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        m = pts1.shape[0]
        E = np.zeros((3, 3)); mask = np.zeros(max(m, 1), np.uint8); info = np.zeros(4, np.int32)
        L.check(self._lib.rpe_find_essential(self._h, _ptr(pts1), _ptr(pts2), m, _ptr(K), _ptr(E), _ptr(mask),
                                             _ptr(info)), "rpe_find_essential")
        return {"ok": bool(info[0]), "E": E, "mask": mask[:m], "best": int(info[1]), "iters": int(info[2]),
                "solved": int(info[3])}

    def ransac_trace(self, max_hyp=4096):
        subset = np.zeros((max_hyp, 5), np.int32); nm = np.zeros(max_hyp, np.int32)
        counts = np.zeros((max_hyp, 10), np.int32); models = np.zeros((max_hyp, 10, 3, 3))
        n = C.c_int32()
        L.check(self._lib.rpe_ransac_trace(self._h, max_hyp, _ptr(subset), _ptr(nm), _ptr(counts), _ptr(models),
                                           C.byref(n)), "rpe_ransac_trace")
        k = n.value
        return {"subset": subset[:k], "nmodels": nm[:k], "counts": counts[:k], "models": models[:k]}

    def recover_pose(self, E, pts1, pts2, K):
        E = np.ascontiguousarray(E, dtype=np.float64).reshape(3, 3)
        pts1 = np.ascontiguousarray(pts1, dtype=np.float32).reshape(-1, 2)
        pts2 = np.ascontiguousarray(pts2, dtype=np.float32).reshape(-1, 2)
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        m = pts1.shape[0]
        R = np.zeros((3, 3)); t = np.zeros((3, 1)); e = np.zeros(3); cnt = C.c_int32()
        mask = np.zeros(max(m, 1), np.uint8)
        L.check(self._lib.rpe_recover_pose(self._h, _ptr(E), _ptr(pts1), _ptr(pts2), m, _ptr(K), _ptr(R), _ptr(t),
                                           _ptr(e), C.byref(cnt), _ptr(mask)), "rpe_recover_pose")
        return {"R": R, "t": t, "euler": e, "count": cnt.value, "mask": mask[:m]}

    def estimate_from_matches(self, pts1, pts2, counts, K):
        """rpe_estimate_from_matches: findEssentialMat + recoverPose + Euler for B pairs of already matched points.
        pts1 / pts2: (B, cap, 2) float32, counts: (B,) matches per pair.  Same dict as estimate_batch."""
        pts1 = np.ascontiguousarray(pts1, dtype=np.float32)
        pts2 = np.ascontiguousarray(pts2, dtype=np.float32)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        B, cap = pts1.shape[0], pts1.shape[1]
        if pts1.shape != (B, cap, 2) or pts2.shape != pts1.shape or counts.shape != (B,):
            raise ValueError("pts1 / pts2 must be (B, cap, 2) and counts (B,)")
        K = np.ascontiguousarray(K, dtype=np.float64).reshape(3, 3)
        R, t, e, st, info = self._outputs(B)
        L.check(self._lib.rpe_estimate_from_matches(self._h, _ptr(pts1), _ptr(pts2), _ptr(counts), B, cap, _ptr(K),
                                                    _ptr(R), _ptr(t), _ptr(e), _ptr(st), _ptr(info)),
                "rpe_estimate_from_matches")
        return {"R": R, "t": t, "euler": e, "status": st, "info": info}

    def evaluate_sequence(self, est_R, est_t, est_euler, gt_euler, gt_pos, euler_convention: int = L.EULER_YUP):
        """rpe_evaluate_sequence: per-frame errors (n, 5) = roll, pitch, yaw, rotation, translation-direction (deg)
        and summary (5, 5) = per metric mean, std, median, max, min.  Angles are (yaw, pitch, roll) degrees."""
        est_R = np.ascontiguousarray(est_R, dtype=np.float64).reshape(-1, 9)
        n = est_R.shape[0]
        a = [np.ascontiguousarray(v, dtype=np.float64).reshape(n, 3) for v in (est_t, est_euler, gt_euler, gt_pos)]
        errors = np.zeros((n, 5)); summary = np.zeros((5, 5))
        L.check(self._lib.rpe_evaluate_sequence(self._h, n, _ptr(est_R), _ptr(a[0]), _ptr(a[1]), _ptr(a[2]), _ptr(a[3]),
                                                int(euler_convention), _ptr(errors), _ptr(summary)),
                "rpe_evaluate_sequence")
        return errors, summary

    def summarize_errors(self, errors):
        errors = np.ascontiguousarray(errors, dtype=np.float64).reshape(-1, 5)
        summary = np.zeros((5, 5))
        L.check(self._lib.rpe_summarize_errors(self._h, errors.shape[0], _ptr(errors), _ptr(summary)), "rpe_summarize_errors")
        return summary

    def get_stage(self, pair: int, view: int, stage: int, arg: int = 0, dtype=np.uint8, nbytes: int | None = None):
        if nbytes is None:
            nbytes = max(self.kp_cap * 32, self.cfg.max_width * self.cfg.max_height)
        buf = np.zeros(nbytes, np.uint8)
        n = C.c_size_t(nbytes)
        L.check(self._lib.rpe_get_stage(self._h, pair, view, stage, arg, _ptr(buf), C.byref(n)), "rpe_get_stage")
        return buf[:n.value].view(dtype)

    def keypoints(self, pair: int, view: int):
        g = lambda s, dt: self.get_stage(pair, view, s, 0, dt).copy()
        return {"x": g(L.STAGE_KP_X, np.float32), "y": g(L.STAGE_KP_Y, np.float32),
                "size": g(L.STAGE_KP_SIZE, np.float32), "angle": g(L.STAGE_KP_ANGLE, np.float32),
                "response": g(L.STAGE_KP_RESPONSE, np.float32), "octave": g(L.STAGE_KP_OCTAVE, np.int32),
                "desc": g(L.STAGE_DESCRIPTORS, np.uint8).reshape(-1, 32)}
