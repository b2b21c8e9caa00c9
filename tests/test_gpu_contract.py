"""GPU tests of the failure contract and of the ABI corners the round-1 suite left out (a9 on the
device, overflow flag, device Euler 'zyx', host rows with pitch > W, the pending-batch guard of the blocking
sequence calls, max_matches=None at 720p), all through the C ABI."""
import numpy as np
import pytest

import cv2
from conftest import default_K, load_gray

from oracle import geometry as og
from oracle import matcher as om
from oracle import orb as oo
from oracle.reference_path import PathConfig, ReferencePath

pytestmark = pytest.mark.gpu


def _dots(points, val=200, shape=(240, 320)):
    a = np.full(shape, 100, np.uint8)
    for x, y in points:
        a[y, x] = val
    return a


@pytest.mark.parametrize("n_dots,expected", [(1, 2), (2, 4)])
def test_insufficient_matches_status_and_message(n_dots, expected):
    """reference: src/core/pose_estimator.py:514-515 — fewer than 5 matches raise
    "Insufficient matches: n (minimum 5 required)" with the match count n; the batch call reports
    RPE_STATUS_INSUFFICIENT_MATCHES and n in info[0].  Isolated bright pixels give a handful of keypoints each
    (one per pyramid level while the contrast lasts); the expected n is what the reference's cv2 path counts."""
    import rpe_b200
    from rpe_b200 import _lib as L
    K = np.array([[300.0, 0, 160], [0, 300, 120], [0, 0, 1]])
    dots = [(60 + 40 * i, 80 + 17 * i) for i in range(n_dots)]
    a = _dots(dots + [(250, 200), (280, 60)])
    b = _dots([(x + 2, y + 1) for x, y in dots])
    with pytest.raises(RuntimeError) as ref_err:
        ReferencePath(K, PathConfig(), canonical=True).estimate(a, b)
    assert str(ref_err.value) == f"Insufficient matches: {expected} (minimum 5 required)"
    est = rpe_b200.PoseEstimator(K)
    with pytest.raises(RuntimeError) as err:
        est.estimate(a, b)
    assert str(err.value) == str(ref_err.value)
    with pytest.raises(RuntimeError, match=f"Insufficient matches: {expected} "):
        est.estimate_with_debug(a, b)
    # batched: no exception, status code + count; a good pair in the same batch is unaffected
    s1, s2 = load_gray("sim_000000.png")[:240, :320].copy(), load_gray("sim_000015.png")[:240, :320].copy()
    R, t, status = est.estimate_batch([a, s1], [b, s2])
    assert status[0] == L.STATUS_INSUFFICIENT_MATCHES and est.last["info"][0, 0] == expected
    assert not R[0].any() and not t[0].any()
    assert status[1] == L.STATUS_OK and abs(np.linalg.det(R[1]) - 1) < 1e-9


def test_candidate_overflow_sets_the_flag():
    """RPE_STATUS_FLAG_OVERFLOW (include/rpe.h): a level whose FAST + NMS corners outnumber the candidate capacity
    (w*h/16) is truncated and flagged — per image (rpe_orb_detect flags) and per pair (status bit 0x100).
    Isolated bright pixels on a staggered 4 x 3 grid: every one is a FAST corner and an NMS maximum (density 1/12)."""
    import rpe_b200
    from rpe_b200 import _lib as L
    img = np.full((480, 640), 60, np.uint8)
    img[0::6, 0::4] = 200
    img[3::6, 2::4] = 200
    n_corners = int((oo.nms_3x3(oo.fast_score_map(img, 15))[31:-31, 31:-31] > 0).sum())
    assert n_corners > 640 * 480 // 16                      # more than level 0's capacity
    est = rpe_b200.PoseEstimator(default_K())
    res = est._detect_and_compute(img)
    assert res["flags"] & 1
    R, t, status = est.estimate_batch([img, load_gray("sim_000000.png")], [img, load_gray("sim_000015.png")])
    assert status[0] & L.STATUS_FLAG_OVERFLOW
    assert status[1] == L.STATUS_OK
    # the wrapper raises only for the status CODE; the flag alone is not an error
    code = int(status[0]) & 0xFF
    assert code in (L.STATUS_OK, L.STATUS_INSUFFICIENT_MATCHES, L.STATUS_E_FAILED)


def test_device_euler_zyx():
    """euler_convention='zyx' (the reference's phone / salah runs): the Euler angles computed in k_pose against
    rotation_to_euler_zyx of the returned R (reference: src/utils/geometry.py:181-205), plus the host twin."""
    import rpe_b200
    a, b, c = load_gray("phone_000000.png"), load_gray("phone_000005.png"), load_gray("sim_000000.png")
    K = default_K(a.shape[1], a.shape[0])
    for conv, fn in (("zyx", og.rotation_to_euler_zyx), ("yup", og.rotation_to_euler_yup)):
        eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=a.shape[1], max_height=a.shape[0], max_pairs=2,
                                                     euler_convention=conv))
        try:
            out = eng.estimate_batch([a, b], [b, a], K)
            assert (out["status"] == 0).all()
            for i in range(2):
                assert np.allclose(out["euler"][i], fn(out["R"][i]), atol=1e-9), conv
                assert np.allclose(rpe_b200.rotation_to_euler(out["R"][i], conv), out["euler"][i], atol=1e-9)
        finally:
            eng.close()
    assert not np.allclose(og.rotation_to_euler_zyx(out["R"][0]), og.rotation_to_euler_yup(out["R"][0]), atol=1e-3)
    del c


def test_host_rows_with_pitch_larger_than_width():
    """The ABI takes a row pitch: frames that sit in a wider host buffer (pitch = W + 48, and an odd W whose device
    pitch is padded) give bit-identical results to tight copies."""
    import rpe_b200
    a, b = load_gray("sim_000000.png"), load_gray("sim_000015.png")
    K = default_K()
    for W in (640, 601):
        H = 480
        tight = np.ascontiguousarray(np.stack([a[:, :W], b[:, :W]]))
        wide = np.full((2, H, W + 48), 255, np.uint8)
        wide[:, :, :W] = tight
        view = wide[:, :, :W]
        assert view.strides[1] == W + 48
        eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=1))
        try:
            want = eng.estimate_batch_interleaved(tight, K)
            kp_want = eng.keypoints(0, 1)
            got = eng.estimate_batch_interleaved(view, K)
            kp_got = eng.keypoints(0, 1)
            assert want["status"][0] == 0
            for k in ("R", "t", "euler", "status", "info"):
                assert np.array_equal(got[k], want[k]), (W, k)
            for k in kp_want:
                assert np.array_equal(kp_got[k], kp_want[k]), (W, k)
        finally:
            eng.close()


def test_blocking_calls_are_refused_while_a_batch_is_pending():
    """include/rpe.h: while a submitted batch is pending every other call fails with RPE_ERR_INVALID — including the
    blocking sequence calls, which would otherwise wait for the OLDER batch and return zeros."""
    import rpe_b200
    from rpe_b200._lib import RpeError
    a, b, c = load_gray("sim_000000.png"), load_gray("sim_000015.png"), load_gray("sim_000030.png")
    K = default_K()
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=640, max_height=480, max_pairs=2))
    try:
        pair = np.ascontiguousarray(np.stack([a, b]))
        want = eng.estimate_batch_interleaved(pair, K)
        seq = np.ascontiguousarray(np.stack([a, b, c]))
        want_seq = eng.estimate_sequence(seq, K)
        eng.submit_interleaved(pair, K)
        with pytest.raises(RpeError, match="pending"):
            eng.estimate_sequence(seq, K)
        with pytest.raises(RpeError, match="pending"):
            eng.estimate_batch_interleaved(pair, K)
        with pytest.raises(RpeError, match="pending"):
            eng.orb_detect(a)
        assert len(eng._pending) == 1                      # the host-side queue stayed in step with the C side
        got = eng.wait()
        for k in ("R", "t", "euler", "status", "info"):
            assert np.array_equal(got[k], want[k]), k
        with pytest.raises(RpeError):
            eng.wait()                                      # nothing pending any more
        assert len(eng._pending) == 0
        again = eng.estimate_sequence(seq, K)               # and the handle is usable
        for k in ("R", "t", "status"):
            assert np.array_equal(again[k], want_seq[k]), k
    finally:
        eng.close()


def test_no_cut_cross_check_at_720p():
    """max_matches=None (no cut) on a 1280x720 pair with 5000 features, cross-check mode (the reference's matcher):
    the whole match list equals cv2 BFMatcher(crossCheck=True) + the reference's stable sort."""
    import rpe_b200
    from bench_support.synth import synth_K, synth_pair
    from rpe_b200 import _lib as L
    a, b = synth_pair(3, 5, 1280, 720)[:2]
    K = synth_K(1280, 720)
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=1280, max_height=720, max_pairs=1, nfeatures=5000,
                                                 match_mode="crosscheck", max_matches=None))
    try:
        out = eng.estimate_batch([a], [b], K)
        assert out["status"][0] == 0
        kp1, kp2 = eng.keypoints(0, 0), eng.keypoints(0, 1)
        got = eng.get_stage(0, 0, L.STAGE_MATCHES, 0, np.int32).reshape(-1, 3)
        bf = cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True)
        ms = sorted(bf.match(kp1["desc"], kp2["desc"]), key=lambda m: m.distance)
        want = np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], np.int32)
        assert len(want) > 2500 and np.array_equal(got, want)
        assert np.array_equal(got, om.sort_and_cut(om.match_crosscheck(kp1["desc"], kp2["desc"]), None))
        assert out["info"][0, 0] == len(want)
    finally:
        eng.close()


def _corrupt(pts2, frac, seed, W, H):
    """replace a fraction of the correspondences by uniform random points (gross outliers)"""
    rng = np.random.default_rng(seed)
    out = pts2.copy()
    idx = rng.permutation(len(out))[:int(round(frac * len(out)))]
    out[idx] = np.stack([rng.uniform(0, W, len(idx)), rng.uniform(0, H, len(idx))], axis=1).astype(np.float32)
    return out


def test_estimate_from_matches_batch_and_hard_ransac():
    """rpe_estimate_from_matches = findEssentialMat + recoverPose over a batch of matched point sets (reference:
    src/core/pose_estimator.py:518-533).  (i) equals the single-pair stage calls bit for bit, (ii) a short point set
    reports INSUFFICIENT_MATCHES without disturbing its neighbours, (iii) on correspondences with 30 / 50 / 70 % gross
    outliers - where cv2 and the engine run tens to thousands of RANSAC iterations instead of the dozen of clean data
    - every pair agrees with cv2 on the same points: same iteration count path, R within 0.1 deg, t within 0.5 deg,
    inlier IoU >= 0.95."""
    import rpe_b200
    from rpe_b200 import _lib as L
    from test_gpu_parity import _matched_points
    K = default_K()
    pts1, pts2 = _matched_points()
    M = len(pts1)
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=640, max_height=480, max_pairs=8, ransac_max_iters=2000))
    try:
        cases = [(pts1, pts2), (pts1[:300], pts2[:300]), (pts1[:4], pts2[:4])]
        for k, frac in enumerate((0.3, 0.5, 0.5, 0.7, 0.7)):
            cases.append((pts1, _corrupt(pts2, frac, 100 + k, 640, 480)))
        B = len(cases)
        P1 = np.zeros((B, M, 2), np.float32); P2 = np.zeros((B, M, 2), np.float32)
        counts = np.array([len(c[0]) for c in cases], np.int32)
        for i, (a, b) in enumerate(cases):
            P1[i, :len(a)] = a; P2[i, :len(b)] = b
        out = eng.estimate_from_matches(P1, P2, counts, K)
        assert out["status"][2] == L.STATUS_INSUFFICIENT_MATCHES and not out["R"][2].any()
        ok = [i for i in range(B) if i != 2]
        assert (out["status"][ok] == 0).all(), out["status"]
        masks = {i: eng.get_stage(i, 0, L.STAGE_RANSAC_MASK, 0, np.uint8).astype(bool).copy() for i in ok}
        iters = out["info"][:, 3]
        assert iters[0] <= 64 and iters[3:].min() > 16 and iters[6:].min() == 2000, iters      # clean vs hard: cap reached at 70 %
        for i in ok:
            a, b = cases[i]
            # (i) the single-pair stage entry points run the same kernels
            one = eng.find_essential(a, b, K)
            assert one["iters"] == iters[i] and one["best"] == out["info"][i, 2]
            rp = eng.recover_pose(one["E"], a, b, K)
            assert np.array_equal(rp["R"], out["R"][i]) and np.array_equal(rp["t"], out["t"][i])
            assert np.array_equal(one["mask"].astype(bool), masks[i])
            # (iii) cv2 on the same points
            E_cv, m_cv = cv2.findEssentialMat(a, b, K, method=cv2.RANSAC, prob=0.999, threshold=1.0, maxIters=2000)
            n_cv, R_cv, t_cv, _ = cv2.recoverPose(E_cv, a, b, K)
            m_cv = m_cv.ravel().astype(bool)
            iou = (masks[i] & m_cv).sum() / max(1, (masks[i] | m_cv).sum())
            rot, tdir = og.rotation_angle_deg(out["R"][i], R_cv), og.direction_angle_deg(out["t"][i], t_cv)
            assert rot < 0.1 and tdir < 0.5 and iou >= 0.95, (i, rot, tdir, iou, int(iters[i]))
            assert abs(int(out["info"][i, 1]) - n_cv) <= 2
    finally:
        eng.close()


def test_device_resident_frames_are_used_in_place_or_copied_with_identical_results():
    """rpe_estimate_batch_device / rpe_estimate_sequence_device: frames whose device layout already is the pyramid's
    (pitch = align16(W), 16-byte aligned) serve as level 0 in place (no copy); any other pitch / alignment is copied.
    Same bits either way, and the same as the host-buffer call; the caller's buffer is not modified."""
    import torch
    import rpe_b200
    from rpe_b200 import _lib as L
    a, b, c = load_gray("sim_000000.png"), load_gray("sim_000015.png"), load_gray("sim_000030.png")
    K = default_K()
    H, W = a.shape
    frames = np.ascontiguousarray(np.stack([a, b, b, c]))
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=3))
    try:
        want = eng.estimate_batch_interleaved(frames, K)
        assert (want["status"] == 0).all()
        want_seq = eng.estimate_sequence(np.ascontiguousarray(np.stack([a, b, c])), K)
        dev = torch.from_numpy(frames).cuda()
        keep = dev.clone()
        got = eng.estimate_batch_device(dev.data_ptr(), H * W, 2, H, W, W, K)                   # in place
        lvl0 = eng.get_stage(1, 1, L.STAGE_PYRAMID_LEVEL, 0, np.uint8, nbytes=H * W).reshape(H, W)
        assert np.array_equal(lvl0, c)
        lvl1 = eng.get_stage(0, 0, L.STAGE_PYRAMID_LEVEL, 1, np.uint8, nbytes=H * W)
        host1 = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=1))
        try:
            host1.estimate_batch([a], [b], K)
            assert np.array_equal(lvl1, host1.get_stage(0, 0, L.STAGE_PYRAMID_LEVEL, 1, np.uint8, nbytes=H * W))
        finally:
            host1.close()
        wide = torch.zeros((4, H, W + 16), dtype=torch.uint8, device="cuda")                     # other pitch: copied
        wide[:, :, :W] = dev
        got_wide = eng.estimate_batch_device(wide.data_ptr(), H * (W + 16), 2, H, W, W + 16, K)
        raw = torch.zeros(4 * H * W + 64, dtype=torch.uint8, device="cuda")                       # unaligned base: copied
        raw[4:4 + 4 * H * W] = dev.reshape(-1)
        got_off = eng.estimate_batch_device(raw.data_ptr() + 4, H * W, 2, H, W, W, K)
        for g in (got, got_wide, got_off):
            for k in ("R", "t", "euler", "status", "info"):
                assert np.array_equal(g[k], want[k]), k
        seq = torch.from_numpy(np.ascontiguousarray(np.stack([a, b, c]))).cuda()
        got_seq = eng.estimate_sequence_device(seq.data_ptr(), H * W, 3, H, W, W, K)
        for k in ("R", "t", "euler", "status", "info"):
            assert np.array_equal(got_seq[k], want_seq[k]), k
        assert torch.equal(dev, keep)
    finally:
        eng.close()


def test_a_second_smaller_handle_does_not_break_the_first():
    """The dynamic shared-memory limit of a kernel belongs to the function, not to a handle: creating a handle with a
    smaller configuration (fewer features, lower images, fewer matches) must not lower the limit under a larger handle
    created before it."""
    import rpe_b200
    big_img = load_gray("salah_000000.png")                 # 1920 x 1080
    big = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=1920, max_height=1080, max_pairs=1, nfeatures=8000, max_matches=None))
    try:
        first = big.orb_detect(big_img)
        small = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=320, max_height=240, max_pairs=1, nfeatures=500, max_matches=100))
        try:
            small.orb_detect(np.ascontiguousarray(big_img[:240, :320]))
            again = big.orb_detect(big_img)
            out = big.estimate_batch([big_img], [big_img], default_K(1920, 1080))
        finally:
            small.close()
        assert np.array_equal(first["desc"], again["desc"]) and len(first["x"]) > 4000
        assert out["status"][0] == 0 and out["info"][0, 0] > 4000
    finally:
        big.close()
