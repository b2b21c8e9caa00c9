"""GPU parity tests: the CUDA path, called through the C ABI (librpe_b200.so), against the
oracle and live cv2 on the same inputs.  Bars: bit-exact keypoints / descriptors / matches;
R within 0.1 deg, t direction within 0.5 deg, inlier IoU >= 0.95 (north_star)."""
import numpy as np
import pytest

import cv2
from conftest import default_K, dump, load_gray

from oracle import geometry as og
from oracle import matcher as om
from oracle import orb as oo
from oracle.reference_path import PathConfig, ReferencePath

pytestmark = pytest.mark.gpu

FIXTURES = ["sim_000000.png", "sim_000015.png", "sim_000030.png", "phone_000000.png", "phone_000005.png",
            "salah_000000.png"]


@pytest.fixture(scope="module")
def engine():
    import rpe_b200
    e = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=1920, max_height=1080, max_pairs=4))
    yield e
    e.close()


@pytest.fixture(scope="module")
def engine_ratio():
    import rpe_b200
    e = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=640, max_height=480, max_pairs=2, match_mode="ratio",
                                              max_matches=None, ransac_max_iters=2000))
    yield e
    e.close()


def _orb_diff(res, ref):
    out = {"n_gpu": len(res["x"]), "n_ref": len(ref)}
    if out["n_gpu"] == out["n_ref"]:
        for f in ["x", "y", "size", "angle", "response", "octave"]:
            out[f] = int((res[f] != getattr(ref, f)).sum())
        out["desc_bits"] = int(np.unpackbits(res["desc"] ^ ref.desc).sum())
    return out


@pytest.mark.parametrize("name", FIXTURES)
def test_orb_bit_exact(engine, name):
    img = load_gray(name)
    res = engine.orb_detect(img)
    ref = oo.orb_detect_and_compute(img, oo.OrbParams())
    d = _orb_diff(res, ref)
    if any(v != 0 for k, v in d.items() if k not in ("n_gpu", "n_ref")) or d["n_gpu"] != d["n_ref"]:
        dump("orb_mismatch_" + name.replace(".png", ""), **{k: np.asarray(v) for k, v in res.items()},
             ref_x=ref.x, ref_y=ref.y, ref_octave=ref.octave, ref_angle=ref.angle, ref_response=ref.response,
             ref_desc=ref.desc)
    assert res["flags"] == 0
    assert d["n_gpu"] == d["n_ref"], d
    assert all(v == 0 for k, v in d.items() if k not in ("n_gpu", "n_ref")), d


def test_orb_matches_cv2_canonical(engine):
    img = load_gray("sim_000000.png")
    p = oo.OrbParams()
    kps, desc = cv2.ORB_create(nfeatures=4000, scaleFactor=1.1, nlevels=12, fastThreshold=15,
                               scoreType=cv2.ORB_HARRIS_SCORE).detectAndCompute(img, None)
    ref = oo.canonicalize_cv2(kps, desc, oo.level_geometry(img.shape[1], img.shape[0], p))
    d = _orb_diff(engine.orb_detect(img), ref)
    assert d["n_gpu"] == d["n_ref"] == 2990 and all(v == 0 for k, v in d.items() if k not in ("n_gpu", "n_ref")), d


def test_pyramid_and_blur_levels(engine):
    import rpe_b200._lib as L
    a, b = load_gray("phone_000000.png"), load_gray("phone_000005.png")
    engine.estimate_batch([a], [b], default_K(848, 478))
    for view, img in enumerate((a, b)):
        _, levels, pyr, blurred = oo.orb_detect_and_compute(img, oo.OrbParams(), return_pyramid=True)
        for l, lv in enumerate(levels):
            got = engine.get_stage(0, view, L.STAGE_PYRAMID_LEVEL, l).reshape(lv.h, lv.w)
            assert np.array_equal(got, pyr[l]), f"pyramid level {l} view {view}: {(got != pyr[l]).sum()} px differ"
            got = engine.get_stage(0, view, L.STAGE_BLURRED_LEVEL, l).reshape(lv.h, lv.w)
            assert np.array_equal(got, blurred[l]), f"blur level {l} view {view}: {(got != blurred[l]).sum()} px differ"


def _oracle_features(name):
    r = oo.orb_detect_and_compute(load_gray(name), oo.OrbParams())
    return r


def test_match_crosscheck_bit_exact(engine):
    r1, r2 = _oracle_features("sim_000000.png"), _oracle_features("sim_000015.png")
    got = engine.match(r1.desc, r2.desc)
    want = om.sort_and_cut(om.match_crosscheck(r1.desc, r2.desc), 500)
    ms = sorted(cv2.BFMatcher(cv2.NORM_HAMMING, crossCheck=True).match(r1.desc, r2.desc), key=lambda m: m.distance)[:500]
    want_cv = np.array([[m.queryIdx, m.trainIdx, int(m.distance)] for m in ms], np.int32)
    assert np.array_equal(want, want_cv)
    assert got.shape == want.shape and np.array_equal(got, want), (got[:5], want[:5])


def test_match_ratio_bit_exact(engine_ratio):
    r1, r2 = _oracle_features("sim_000000.png"), _oracle_features("sim_000015.png")
    got = engine_ratio.match(r1.desc, r2.desc)
    want = om.sort_and_cut(om.match_ratio(r1.desc, r2.desc, 0.75), None)
    assert got.shape == want.shape and np.array_equal(got, want)


def test_match_edge_cases(engine):
    rng = np.random.default_rng(0)
    d = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    assert engine.match(d[:0], d).shape == (0, 3)
    assert engine.match(d, d[:0]).shape == (0, 3)
    # duplicates force ties: lowest index must win both ways
    d2 = np.concatenate([d[:100], d[:100], d[100:]])
    got = engine.match(d, d2)
    want = om.sort_and_cut(om.match_crosscheck(d, d2), 500)
    assert np.array_equal(got, want)
    one = engine.match(d[:1], d[5:6])
    assert one.shape == (1, 3) and one[0, 0] == 0 and one[0, 1] == 0


def _matched_points(name1="sim_000000.png", name2="sim_000015.png"):
    r1, r2 = _oracle_features(name1), _oracle_features(name2)
    m = om.sort_and_cut(om.match_crosscheck(r1.desc, r2.desc), 500)
    pts1 = np.stack([r1.x[m[:, 0]], r1.y[m[:, 0]]], axis=1).astype(np.float32)
    pts2 = np.stack([r2.x[m[:, 1]], r2.y[m[:, 1]]], axis=1).astype(np.float32)
    return pts1, pts2


def test_ransac_parity(engine):
    pts1, pts2 = _matched_points()
    K = default_K()
    got = engine.find_essential(pts1, pts2, K)
    E_cv, mask_cv = cv2.findEssentialMat(pts1, pts2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    mask_cv = mask_cv.ravel().astype(bool)
    tr = engine.ransac_trace()
    # hypothesis stream: identical subsets to OpenCV's RNG
    rng = og.CvRNG()
    for i in range(min(20, len(tr["subset"]))):
        assert tr["subset"][i].tolist() == og.draw_subset(rng, len(pts1))
    # per-hypothesis inlier counts against the oracle solver (5-pt roots are ill-conditioned:
    # allow a small count jitter on a few models)
    x1, x2 = og.normalize_points(pts1, K), og.normalize_points(pts2, K)
    thr = np.float32((1.0 / ((K[0, 0] + K[1, 1]) / 2)) ** 2)
    bad = 0
    for i in range(min(12, len(tr["subset"]))):
        s = tr["subset"][i]
        want = sorted(int((og.sampson_error_f32(E, x1, x2) <= thr).sum()) for E in og.five_point_models(x1[s], x2[s]))
        have = sorted(tr["counts"][i][:tr["nmodels"][i]].tolist())
        if len(want) != len(have) or max(abs(a - b) for a, b in zip(want, have)) > 3:
            bad += 1
    assert bad <= 1, bad
    m = got["mask"].astype(bool)
    iou = (m & mask_cv).sum() / max(1, (m | mask_cv).sum())
    dump("ransac", E=got["E"], E_cv=E_cv, mask=m, mask_cv=mask_cv, counts=tr["counts"], nmodels=tr["nmodels"])
    assert got["ok"] and iou >= 0.95, iou
    assert min(np.abs(got["E"] - E_cv).max(), np.abs(got["E"] + E_cv).max()) < 1e-3
    assert got["iters"] == 12


def test_ransac_degenerate_inputs(engine):
    K = default_K()
    pts = np.random.default_rng(1).uniform(50, 400, (4, 2)).astype(np.float32)
    assert not engine.find_essential(pts, pts + 1, K)["ok"]          # < 5 points
    pts5 = np.random.default_rng(2).uniform(50, 400, (5, 2)).astype(np.float32)
    assert not engine.find_essential(pts5, pts5 + 3, K)["ok"]        # M == 5 -> E_FAILED (documented)


def test_recover_pose_parity(engine):
    pts1, pts2 = _matched_points()
    K = default_K()
    E_cv, _ = cv2.findEssentialMat(pts1, pts2, K, method=cv2.RANSAC, prob=0.999, threshold=1.0)
    n_cv, R_cv, t_cv, mask_cv = cv2.recoverPose(E_cv, pts1, pts2, K)
    got = engine.recover_pose(E_cv, pts1, pts2, K)
    assert got["count"] == n_cv
    assert np.array_equal(got["mask"], mask_cv.ravel())
    assert np.abs(got["R"] - R_cv).max() < 1e-12 and np.abs(got["t"] - t_cv).max() < 1e-12
    want = og.rotation_to_euler_yup(R_cv)
    assert np.allclose(got["euler"], want, atol=1e-9)


PAIRS = [("sim_000000.png", "sim_000015.png"), ("sim_000015.png", "sim_000030.png"),
         ("phone_000000.png", "phone_000005.png")]


@pytest.mark.parametrize("a,b", PAIRS)
def test_end_to_end_vs_canonical_reference(a, b):
    import rpe_b200
    img1, img2 = load_gray(a), load_gray(b)
    K = default_K(img1.shape[1], img1.shape[0])
    est = rpe_b200.PoseEstimator(K, nfeatures=4000)
    dbg = est.estimate_with_debug(img1, img2)
    ref = ReferencePath(K, PathConfig(), canonical=True).estimate_with_debug(img1, img2)
    assert dbg["num_matches"] == ref["num_matches"]
    assert np.array_equal(dbg["pts1"], ref["pts1"]) and np.array_equal(dbg["pts2"], ref["pts2"])
    rot = og.rotation_angle_deg(dbg["R"], ref["R"])
    tdir = og.direction_angle_deg(dbg["t"], ref["t"])
    assert rot < 0.1 and tdir < 0.5, (rot, tdir)
    assert abs(dbg["inliers"] - ref["inliers"]) <= 2
    R, t = est.estimate(img1, img2)
    assert np.array_equal(R, dbg["R"]) and np.array_equal(t, dbg["t"])
    assert R.shape == (3, 3) and t.shape == (3, 1) and R.dtype == np.float64
    assert abs(np.linalg.det(R) - 1) < 1e-9 and abs(np.linalg.norm(t) - 1) < 1e-9


def test_batch_equals_single(engine):
    import rpe_b200
    a, b, c = load_gray("sim_000000.png"), load_gray("sim_000015.png"), load_gray("sim_000030.png")
    K = default_K()
    out = engine.estimate_batch([a, b, a], [b, c, c], K)
    assert (out["status"] == 0).all()
    for i, (x, y) in enumerate([(a, b), (b, c), (a, c)]):
        one = engine.estimate_batch([x], [y], K)
        assert np.array_equal(one["R"][0], out["R"][i]) and np.array_equal(one["t"][0], out["t"][i])
        eu = og.rotation_to_euler_yup(out["R"][i])
        assert np.allclose(out["euler"][i], eu, atol=1e-9)


def test_failure_statuses(engine):
    import rpe_b200
    K = default_K()
    blank = np.full((480, 640), 128, np.uint8)
    a = load_gray("sim_000000.png")
    out = engine.estimate_batch([blank, a], [a, a], K)
    assert out["status"][0] & 0xFF == 1 and out["status"][1] == 0
    est = rpe_b200.PoseEstimator(K)
    with pytest.raises(RuntimeError, match="Could not compute descriptors"):
        est.estimate(blank, a)
    noise = np.random.default_rng(0).integers(0, 255, (480, 640), dtype=np.uint8)
    noise2 = np.random.default_rng(1).integers(0, 255, (480, 640), dtype=np.uint8)
    # unrelated noise images: whatever happens must mirror the reference path's outcome class
    ref = ReferencePath(K, PathConfig(), canonical=True)
    try:
        ref.estimate(noise, noise2)
        ref_ok = True
    except RuntimeError:
        ref_ok = False
    try:
        est.estimate(noise, noise2)
        ok = True
    except RuntimeError:
        ok = False
    assert ok == ref_ok


def test_pipeline_chunks_do_not_change_results(engine):
    """rpe_set_pipeline only changes scheduling (chunks over streams): results must be identical,
    including with a ragged last chunk and per-pair failures inside a chunk."""
    a, b, c = load_gray("sim_000000.png"), load_gray("sim_000015.png"), load_gray("sim_000030.png")
    blank = np.full((480, 640), 128, np.uint8)
    K = default_K()
    v1, v2 = [a, b, blank, a], [b, c, a, c]
    engine.set_pipeline(4, 1)
    want = engine.estimate_batch(v1, v2, K)
    try:
        for chunk, streams in [(1, 3), (3, 2), (2, 4)]:
            engine.set_pipeline(chunk, streams)
            got = engine.estimate_batch(v1, v2, K)
            for k in ("R", "t", "euler", "status", "info"):
                assert np.array_equal(got[k], want[k]), (chunk, streams, k)
            for i in range(4):
                m = engine.get_stage(i, 0, 9, 0, np.int32)
                engine.set_pipeline(4, 1)
                engine.estimate_batch(v1, v2, K)
                assert np.array_equal(m, engine.get_stage(i, 0, 9, 0, np.int32))
                engine.set_pipeline(chunk, streams)
                engine.estimate_batch(v1, v2, K)
    finally:
        engine.set_pipeline(32, 3)
    assert want["status"][2] & 0xFF == 1 and (want["status"][[0, 1, 3]] == 0).all()


def test_submit_wait_equals_blocking_call(engine):
    """rpe_submit_batch / rpe_wait_batch, two batches pending on one handle plus a second handle in flight ==
    the blocking calls; a third submit on a busy handle and stage calls while a batch is pending are refused."""
    import rpe_b200
    from rpe_b200._lib import RpeError
    a, b, c = load_gray("sim_000000.png"), load_gray("sim_000015.png"), load_gray("sim_000030.png")
    K = default_K()
    frames1 = np.ascontiguousarray(np.stack([a, b, b, c]))
    frames2 = np.ascontiguousarray(np.stack([a, c, b, a]))
    want1 = engine.estimate_batch_interleaved(frames1, K)
    want2 = engine.estimate_batch_interleaved(frames2, K)
    other = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=640, max_height=480, max_pairs=2))
    try:
        engine.submit_interleaved(frames1, K)
        other.submit_interleaved(frames2, K)
        engine.submit_interleaved(frames2, K)             # second pending batch on the same handle
        with pytest.raises(RpeError):
            engine.submit_interleaved(frames2, K)
        with pytest.raises(RpeError):
            engine.orb_detect(a)
        with pytest.raises(RpeError):
            engine.estimate_batch_interleaved(frames1, K)
        got1, got2, got3 = engine.wait(), other.wait(), engine.wait()
        for k in ("R", "t", "euler", "status", "info"):
            assert np.array_equal(got1[k], want1[k]) and np.array_equal(got2[k], want2[k]), k
            assert np.array_equal(got3[k], want2[k]), k
        with pytest.raises(RpeError):
            engine.wait()
        # a long alternating run with two batches always in flight: every result is its own batch's
        engine.set_pipeline(1, 2)
        q = []
        for i in range(9):
            if len(q) == 2:
                j, got = q.pop(0), engine.wait()
                assert np.array_equal(got["R"], (want1 if j % 2 == 0 else want2)["R"]), j
                assert np.array_equal(got["info"], (want1 if j % 2 == 0 else want2)["info"]), j
            engine.submit_interleaved(frames1 if i % 2 == 0 else frames2, K)
            q.append(i)
        for j in q:
            got = engine.wait()
            assert np.array_equal(got["t"], (want1 if j % 2 == 0 else want2)["t"]), j
        engine.set_pipeline(32, 3)
        # stage accessors read the last completed batch
        m = engine.get_stage(0, 0, 9, 0, np.int32).copy()
        engine.estimate_batch_interleaved(frames1, K)
        assert np.array_equal(m, engine.get_stage(0, 0, 9, 0, np.int32))
    finally:
        other.close()


def test_sequence_equals_pair_batch(engine):
    """rpe_estimate_sequence (ORB once per frame) == rpe_estimate_batch on the consecutive pairs, bit for bit,
    for blocking, chunked-pipeline and submit/wait forms; stage accessors address frames by (pair, view)."""
    from rpe_b200 import _lib as L
    names = ["sim_000000.png", "sim_000015.png", "sim_000030.png", "sim_000015.png", "sim_000000.png"]
    frames = [load_gray(n) for n in names]
    K = default_K()
    want = engine.estimate_batch(frames[:-1], frames[1:], K)
    want_kp = [engine.keypoints(i, v) for i in range(4) for v in (0, 1)]
    want_m = [engine.get_stage(i, 0, L.STAGE_MATCHES, 0, np.int32).copy() for i in range(4)]
    assert (want["status"] == 0).all()
    stack = np.ascontiguousarray(np.stack(frames))
    for chunk, streams, src in [(32, 1, frames), (1, 3, frames), (1, 4, stack), (2, 2, stack)]:
        engine.set_pipeline(chunk, streams)
        got = engine.estimate_sequence(src, K)
        for k in ("R", "t", "euler", "status", "info"):
            assert np.array_equal(got[k], want[k]), (k, chunk, streams)
        for i in range(4):
            assert np.array_equal(engine.get_stage(i, 0, L.STAGE_MATCHES, 0, np.int32), want_m[i])
            for v in (0, 1):
                kp = engine.keypoints(i, v)
                for f, a in want_kp[2 * i + v].items():
                    assert np.array_equal(kp[f], a), (i, v, f)
    engine.set_pipeline(32, 4)
    got = engine.submit_sequence(stack, K).wait()
    for k in ("R", "t", "euler", "status", "info"):
        assert np.array_equal(got[k], want[k]), k
    with pytest.raises(RuntimeError):          # max_pairs = 4 pair slots
        engine.estimate_sequence([frames[0]] * 6, K)
    with pytest.raises(Exception):
        engine.estimate_sequence([frames[0]], K)


def test_bgr_ingest_bit_exact_and_equal_to_host_gray(engine):
    """k_bgr2gray == cv2.cvtColor(BGR2GRAY) / IMREAD_GRAYSCALE(PNG) bit for bit (widths that are and are not
    multiples of 4 / 16), and BGR frames through the batch / sequence calls give exactly the results of the
    host-converted gray frames (load_image, image_loader.py:23-28)."""
    from rpe_b200 import _lib as L
    from oracle.gray import bgr_to_gray
    rng = np.random.default_rng(5)
    for (H, W) in [(48, 64), (33, 61), (50, 130), (480, 640), (37, 645)]:
        img = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        assert np.array_equal(engine.bgr_to_gray(img, L.GRAY_CVTCOLOR), cv2.cvtColor(img, cv2.COLOR_BGR2GRAY)), (H, W)
        ok, buf = cv2.imencode(".png", img)
        assert np.array_equal(engine.bgr_to_gray(img, L.GRAY_IMREAD_PNG), cv2.imdecode(buf, cv2.IMREAD_GRAYSCALE))
        assert np.array_equal(engine.bgr_to_gray(img, L.GRAY_IMREAD_PNG), bgr_to_gray(img, 1))
    # colourised fixtures: gray frame g -> BGR with distinct channels, so the conversion matters
    frames = []
    for n in ("sim_000000.png", "sim_000015.png", "sim_000030.png"):
        g = load_gray(n)
        frames.append(np.stack([g, np.roll(g, 1, 1) // 2 + 60, 255 - g // 2], axis=2))
    grays = [cv2.cvtColor(f, cv2.COLOR_BGR2GRAY) for f in frames]
    K = default_K()
    engine.set_pipeline(32, 4)
    want = engine.estimate_sequence(grays, K)
    assert (want["status"] == 0).all()
    want_desc = engine.keypoints(1, 1)["desc"]
    for src in (frames, np.ascontiguousarray(np.stack(frames))):
        got = engine.estimate_sequence(src, K)
        for k in ("R", "t", "euler", "status", "info"):
            assert np.array_equal(got[k], want[k]), k
        assert np.array_equal(engine.keypoints(1, 1)["desc"], want_desc)
        assert np.array_equal(engine.get_stage(0, 1, L.STAGE_PYRAMID_LEVEL, 0).reshape(480, 640), grays[1])
    got = engine.estimate_batch(frames[:-1], frames[1:], K)
    for k in ("R", "t", "euler", "status", "info"):
        assert np.array_equal(got[k], want[k]), k
    # back to gray input on the same handle
    again = engine.estimate_sequence(grays, K)
    assert np.array_equal(again["R"], want["R"])


@pytest.mark.parametrize("size", [(637, 475), (333, 251), (129, 97), (70, 70), (62, 200), (40, 40)])
def test_orb_ragged_sizes_bit_exact(engine, size):
    """Widths / heights that are not multiples of 4 or 16 (row pitch padding, partial tiles, partial quads), levels
    that shrink below ORB's 63-px minimum on the way up the pyramid, and images too small for any keypoint: the
    keypoint set, every field and the descriptors must equal the oracle's (= cv2's) for each size."""
    w, h = size
    img = np.ascontiguousarray(load_gray("sim_000015.png")[37:37 + h, 101:101 + w])
    if h > 480 - 37:      # taller than the crop allows: tile the fixture
        img = np.ascontiguousarray(np.tile(load_gray("sim_000015.png"), (2, 1))[37:37 + h, 101:101 + w])
    res = engine.orb_detect(img)
    ref = oo.orb_detect_and_compute(img, oo.OrbParams())
    kps, desc = cv2.ORB_create(nfeatures=4000, scaleFactor=1.1, nlevels=12, fastThreshold=15,
                               scoreType=cv2.ORB_HARRIS_SCORE).detectAndCompute(img, None)
    assert len(ref) == len(kps)                      # the oracle itself against live cv2 at this size
    d = _orb_diff(res, ref)
    assert res["flags"] == 0 and d["n_gpu"] == d["n_ref"], (size, d)
    assert all(v == 0 for k, v in d.items() if k not in ("n_gpu", "n_ref")), (size, d)
    if min(w, h) < 63:
        assert d["n_gpu"] == 0


@pytest.mark.parametrize("scale,nlevels", [(1.2, 8), (1.3, 6), (1.5, 5), (2.0, 4)])
def test_orb_other_scale_factors_bit_exact(scale, nlevels):
    """The reference hard-codes scaleFactor 1.1 (pose_estimator.py:85-91), which takes the TMA-staged resize kernel;
    rpe_config.scale_factor is a parameter, though, and larger factors take the other two resize kernels
    (k_resize_strip up to 4/3, the generic k_resize beyond).  Same bar: pyramid levels and ORB output bit-exact
    against the oracle, and the oracle's keypoint count against live cv2."""
    import rpe_b200
    import rpe_b200._lib as L
    img = load_gray("salah_000000.png")
    h, w = img.shape
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=w, max_height=h, max_pairs=1, scale_factor=scale, nlevels=nlevels))
    try:
        p = oo.OrbParams(scale_factor=scale, nlevels=nlevels)
        ref, levels, pyr, blurred = oo.orb_detect_and_compute(img, p, return_pyramid=True)
        kps, _ = cv2.ORB_create(nfeatures=4000, scaleFactor=scale, nlevels=nlevels, fastThreshold=15,
                                scoreType=cv2.ORB_HARRIS_SCORE).detectAndCompute(img, None)
        assert len(ref) == len(kps)
        eng.estimate_batch([img], [img], default_K(w, h))
        for l, lv in enumerate(levels):
            got = eng.get_stage(0, 0, L.STAGE_PYRAMID_LEVEL, l).reshape(lv.h, lv.w)
            assert np.array_equal(got, pyr[l]), f"scale {scale}: pyramid level {l}: {(got != pyr[l]).sum()} px differ"
            got = eng.get_stage(0, 0, L.STAGE_BLURRED_LEVEL, l).reshape(lv.h, lv.w)
            assert np.array_equal(got, blurred[l]), f"scale {scale}: blur level {l}"
        d = _orb_diff(eng.orb_detect(img), ref)
        assert d["n_gpu"] == d["n_ref"], (scale, d)
        assert all(v == 0 for k, v in d.items() if k not in ("n_gpu", "n_ref")), (scale, d)
    finally:
        eng.close()
