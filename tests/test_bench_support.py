"""Host-side logic of bench.py and its support package (no GPU): the workload description both arms print, the
config-5 call schedule, the parallel input producer, and the geometry of the synthetic generator (version 2)."""
import os
import sys

import cv2
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench                                              # noqa: E402
from bench_support import hostgen, synth                  # noqa: E402


def test_both_arms_describe_the_same_workload():
    """`same_config` was false because the two arms printed different `config` keys.  Both now print
    workload_config(cfg, B): nothing arm-specific in it."""
    for c, B in ((3, 256), (4, 16)):
        cfg = bench.CFGS[c]
        a, b = bench.workload_config(cfg, B), bench.workload_config(dict(cfg), B)
        assert a == b and a["pairs_per_gpu_per_step"] == B and a["width"] == cfg["W"] and "workload" in a
        assert not any(k in a for k in ("pipeline", "pairs_per_step", "host_numa_binding", "l2_policy"))
    src = open(os.path.join(ROOT, "bench.py")).read()
    assert src.count('"config": workload_config(') == 2 + 1         # reference arm, B200 arm, the cfg4 line item


def test_stream_call_schedule_covers_every_pair_once_per_lap():
    for per_rank, ring, C in ((100000, 4096, 256), (12500, 4096, 256), (300, 256, 256), (100, 100, 100), (1000, 768, 256)):
        calls = bench.stream_calls(per_rank, ring, C)
        assert sum(n for _, n in calls) == per_rank
        pos_expected = 0
        for pos, n in calls:
            assert 0 <= pos and pos + n <= ring and 1 <= n <= C
            assert pos == pos_expected                  # chained: starts at the previous call's last frame, or restarts the ring
            pos_expected = (pos + n) % ring
        first_lap = [c for c in calls[:max(1, ring // C)]]
        assert sum(n for _, n in first_lap) == min(per_rank, ring)


def test_parallel_producer_equals_the_serial_generator():
    W, H = 160, 96
    out = np.zeros((6, H, W), np.uint8)
    rate = hostgen.generate("pairs", out, 2, cfg=3, first=40)
    assert rate > 0
    for i in range(3):
        a, b = synth.synth_pair(3, 40 + i, W, H)[:2]
        assert np.array_equal(out[2 * i], a) and np.array_equal(out[2 * i + 1], b)
    frames = np.zeros((7, H, W), np.uint8)
    hostgen.generate("stream", frames, 3, cfg=5, first=11)
    assert np.array_equal(frames, synth.synth_stream(5, 11, 7, W, H))
    # frames can be produced independently (rank blocks need no state)
    assert np.array_equal(frames[3:6], synth.synth_stream(5, 14, 3, W, H))


def test_generator_v2_scene_is_two_planes_meeting_in_a_ridge():
    """The two plane homographies agree on the seam (the planes meet there), differ off it (so the scene is not a
    plane and the essential matrix is unique), every depth is well inside recoverPose's 50-baseline gate, and the
    ground-truth motion reproduces the rendered view: a point of view 1 re-projected with its plane's depth lands on
    the same grey value in view 2."""
    W, H = 640, 360
    K = synth.synth_K(W, H)
    Ki = np.linalg.inv(K)
    rng = np.random.default_rng(7)
    R, t = synth.synth_motion(rng)
    assert 0.3 <= np.linalg.norm(t) <= 0.6
    Hs = [K @ (R + np.outer(t, n) / d) @ Ki for n, d in synth._PLANES]
    for y in (0.0, H / 2.0, H - 1.0):
        p = np.array([W / 2.0, y, 1.0])
        a, b = Hs[0] @ p, Hs[1] @ p
        assert np.abs(a[:2] / a[2] - b[:2] / b[2]).max() < 1e-9          # same image of a seam point
    p = np.array([0.9 * W, H / 3.0, 1.0])
    a, b = Hs[0] @ p, Hs[1] @ p
    assert np.abs(a[:2] / a[2] - b[:2] / b[2]).max() > 2.0                # not one plane
    depths = []
    for x in np.linspace(0, W - 1, 9):
        ray = Ki @ np.array([x, H / 2.0, 1.0])
        n, d = synth._PLANES[0] if x < W / 2 else synth._PLANES[1]
        depths.append(d / (n @ ray))
    assert 4.9 < min(depths) and max(depths) < 8.5 and max(depths) / np.linalg.norm(t) < 50 * 0.6
    img1, img2, _, Rg, tg = synth.synth_pair(3, 5, W, H)
    rng = np.random.default_rng(1000 * 3 + 5)
    synth.synth_texture(rng, W, H)
    R, t = synth.synth_motion(rng)
    assert np.array_equal(R, Rg) and np.allclose(t / np.linalg.norm(t), tg)
    blur1 = cv2.GaussianBlur(img1, (0, 0), 2.0).astype(np.float64)
    blur2 = cv2.GaussianBlur(img2, (0, 0), 2.0).astype(np.float64)
    errs = []
    for x, y in [(0.2 * W, 0.3 * H), (0.4 * W, 0.7 * H), (0.6 * W, 0.4 * H), (0.85 * W, 0.6 * H)]:
        ray = Ki @ np.array([x, y, 1.0])
        n, d = synth._PLANES[0] if x < W / 2 else synth._PLANES[1]
        X = ray * (d / (n @ ray))
        q = K @ (R @ X + t)
        u, v = q[0] / q[2], q[1] / q[2]
        if 4 <= u < W - 4 and 4 <= v < H - 4:
            errs.append(abs(blur1[int(round(y)), int(round(x))] - blur2[int(round(v)), int(round(u))]))
    assert len(errs) >= 3 and max(errs) < 12.0, errs
