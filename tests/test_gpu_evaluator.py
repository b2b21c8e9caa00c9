"""GPU parity of the on-device sequence evaluation (rpe_evaluate_sequence) against the golden vectors
of the unmodified reference PoseEvaluator and, at BASELINE config-5 length (100 000 frames), against the oracle's
vectorised form.

Tolerances (FP64; the reference's formulas are ill-conditioned only where stated):
  * wrapped angle errors: 1e-9 deg (subtract, fmod, abs — exact up to the last bit of fmod);
  * rotation / translation-direction errors: 2e-6 deg.  acos near +-1 amplifies a 1-ulp difference of the cosine
    (CUDA sincos / product order vs numpy) to sqrt(2 ulp) ~ 1.5e-8 rad ~ 1e-6 deg;
  * mean / std: 1e-9 relative (summation order); median / max / min: exactly elements of the error table."""
import os

import numpy as np
import pytest

from conftest import GOLDEN

from oracle import evaluator as oe

pytestmark = pytest.mark.gpu

TOL = np.array([1e-9, 1e-9, 1e-9, 2e-6, 2e-6])


@pytest.fixture(scope="module")
def engine():
    import rpe_b200
    e = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=64, max_height=64, max_pairs=1, nfeatures=64, nlevels=1))
    yield e
    e.close()


def _cases():
    g = np.load(os.path.join(GOLDEN, "eval_golden.npz"))
    ci = 0
    while f"c{ci}_errors" in g:
        p = f"c{ci}_"
        yield {k[len(p):]: g[k] for k in g.files if k.startswith(p)}
        ci += 1


def _check_summary(summary, errors):
    want = oe.summary_statistics(errors)
    assert np.allclose(summary[:, :2], want[:, :2], rtol=1e-9, atol=1e-12)
    assert np.array_equal(summary[:, 2:], want[:, 2:])        # median / max / min of the SAME table: exact


def test_evaluate_sequence_matches_reference_golden(engine):
    from rpe_b200 import _lib as L
    for c in _cases():
        conv = L.EULER_ZYX if str(c["convention"]) == "zyx" else L.EULER_YUP
        errors, summary = engine.evaluate_sequence(c["est_R"], c["est_t"], c["est_euler"], c["gt_euler"], c["gt_pos"], conv)
        assert (np.abs(errors - c["errors"]) <= TOL).all(), np.abs(errors - c["errors"]).max(axis=0)
        _check_summary(summary, errors)
        assert np.allclose(summary, c["summary"], rtol=1e-9, atol=2e-6)


def test_pose_evaluator_twin_returns_the_reference_dicts(engine):
    import rpe_b200
    c = list(_cases())[0]
    frames = [int(f) for f in c["frames"]]

    class GT:
        def get_pose(self, f):
            i = frames.index(f)
            return {"x": c["gt_pos"][i, 0], "y": c["gt_pos"][i, 1], "z": c["gt_pos"][i, 2],
                    "yaw": c["gt_euler"][i, 0], "pitch": c["gt_euler"][i, 1], "roll": c["gt_euler"][i, 2]}
    ev = rpe_b200.PoseEvaluator(GT(), euler_convention="yup", engine=engine)
    res = ev.evaluate_sequence({"frames": frames, "roll": c["est_euler"][:, 2], "pitch": c["est_euler"][:, 1],
                                "yaw": c["est_euler"][:, 0], "R": list(c["est_R"]), "t": list(c["est_t"])})
    assert list(res) == ["frames", "roll_error", "pitch_error", "yaw_error", "rotation_error", "translation_dir_error",
                         "gt_roll", "gt_pitch", "gt_yaw", "est_roll", "est_pitch", "est_yaw"]
    for k, m in enumerate(oe.METRICS):
        assert (np.abs(res[m] - c["errors"][:, k]) <= TOL[k]).all()
    stats = ev.compute_summary_statistics(res)
    assert list(stats)[:5] == ["roll_error_mean", "roll_error_std", "roll_error_median", "roll_error_max", "roll_error_min"]
    for i, m in enumerate(oe.METRICS):
        for j, s in enumerate(oe.STATS):
            assert abs(stats[f"{m}_{s}"] - c["summary"][i, j]) <= 2e-6
    # a dict that did not come from evaluate_sequence is reduced on the device too
    other = {m: res[m][::-1].copy() for m in oe.METRICS}
    stats2 = ev.compute_summary_statistics(other)
    assert stats2["rotation_error_median"] == stats["rotation_error_median"]


def test_stream_length_and_nan_propagation(engine):
    """100 000 frames (config 5): vectorised numpy form of the oracle; a zero ground-truth step gives NaN in the
    translation metric (0/0 in translation_direction_error) and NaN statistics for that metric only."""
    rng = np.random.default_rng(5)
    n = 100_000
    gt_euler = np.c_[rng.uniform(-179, 179, n), rng.uniform(-80, 80, n), rng.uniform(-179, 179, n)]
    est_euler = gt_euler + rng.normal(0, 1.0, (n, 3))
    gt_pos = np.cumsum(rng.normal(0, 0.3, (n, 3)), axis=0)
    est_R = np.stack([oe.euler_to_rotation(*est_euler[i]) for i in range(0, n, 1000)])
    est_R = np.repeat(est_R, 1000, axis=0)
    est_t = rng.normal(0, 1, (n, 3))
    errors, summary = engine.evaluate_sequence(est_R, est_t, est_euler, gt_euler, gt_pos)
    idx = rng.integers(0, n, 300)
    want = oe.evaluate_sequence(est_R[idx[0]:idx[0] + 50], est_t[idx[0]:idx[0] + 50], est_euler[idx[0]:idx[0] + 50],
                                gt_euler[idx[0]:idx[0] + 50], gt_pos[idx[0]:idx[0] + 50])
    got = errors[idx[0]:idx[0] + 50]
    assert (np.abs(got[1:] - want[1:]) <= TOL).all()
    # vectorised check of the whole table
    d = np.abs(est_euler - gt_euler)
    w = np.abs(((d + 180) % 360) - 180)
    assert np.abs(errors[:, 0] - w[:, 2]).max() <= 1e-9 and np.abs(errors[:, 2] - w[:, 0]).max() <= 1e-9
    delta = np.diff(gt_pos, axis=0)
    cosang = np.einsum("ij,ij->i", est_t[1:] / np.linalg.norm(est_t[1:], axis=1, keepdims=True),
                       delta / np.linalg.norm(delta, axis=1, keepdims=True))
    assert np.abs(errors[1:, 4] - np.rad2deg(np.arccos(np.clip(cosang, -1, 1)))).max() <= 2e-6
    assert errors[0, 4] == 0.0
    _check_summary(summary, errors)
    gt_pos[501] = gt_pos[500]
    errors2, summary2 = engine.evaluate_sequence(est_R, est_t, est_euler, gt_euler, gt_pos)
    assert np.isnan(errors2[501, 4]) and np.isnan(summary2[4]).all() and not np.isnan(summary2[:4]).any()


def test_pipeline_call_sequence_on_the_twin(engine, capsys):
    """INTEGRATION.md swaps the twin in where src/pipeline.py constructs PoseEvaluator; the pipeline then calls
    print_summary (:160) and create_comparison_dataframe (:163).  Same text / columns as the unmodified reference."""
    import rpe_b200
    from test_oracle_evaluator import _TableGT, _pipeline_call_sequence, check_report_against_reference
    c = list(_cases())[0]
    ev = rpe_b200.PoseEvaluator(_TableGT(c), euler_convention="yup", engine=engine)
    text, df, res = _pipeline_call_sequence(ev, c, capsys)
    check_report_against_reference(text, df, res)
