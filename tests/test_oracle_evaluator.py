"""oracle/evaluator.py (numpy restatement of the reference's PoseEvaluator) against the golden
vectors produced by the UNMODIFIED reference class (tests/golden/make_golden_eval.py)."""
import os

import numpy as np

from conftest import GOLDEN

from oracle import evaluator as oe


def cases():
    g = np.load(os.path.join(GOLDEN, "eval_golden.npz"))
    ci = 0
    while f"c{ci}_errors" in g:
        p = f"c{ci}_"
        yield {k[len(p):]: g[k] for k in g.files if k.startswith(p)}
        ci += 1


def test_oracle_evaluator_equals_reference_golden():
    n_cases = 0
    for c in cases():
        conv = str(c["convention"])
        errors = oe.evaluate_sequence(c["est_R"], c["est_t"], c["est_euler"], c["gt_euler"], c["gt_pos"], conv)
        # same numpy calls in the same order as the reference: identical doubles
        assert np.array_equal(errors, c["errors"]), conv
        assert np.array_equal(oe.summary_statistics(errors), c["summary"])
        n_cases += 1
    assert n_cases == 4


def test_golden_covers_the_edge_cases():
    c = list(cases())[0]
    e = c["errors"]
    assert e[0, 4] == 0.0                       # first frame: no previous position (pose_evaluator.py:116-118)
    assert e[5, 3] < 1e-5 and e[9, 4] > 150.0   # identical rotation; opposite translation direction
    assert (e[:, :3] <= 180.0).all() and (e[:, :3] >= 0).all()   # wrapped
    assert e[::7, 2].max() < 30.0               # +360 deg yaw offsets wrap back


def test_host_wrapper_never_falls_back(monkeypatch):
    """rpe_b200.PoseEvaluator has no CPU path: without a GPU the engine creation raises."""
    import pytest
    import torch

    import rpe_b200
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    c = list(cases())[2]

    class GT:
        def get_pose(self, f):
            return {"x": 0.0, "y": 0.0, "z": 0.0, "yaw": 0.0, "pitch": 0.0, "roll": 0.0}
    ev = rpe_b200.PoseEvaluator(GT())
    with pytest.raises(RuntimeError):
        ev.evaluate_sequence({"frames": [1, 2], "roll": np.zeros(2), "pitch": np.zeros(2), "yaw": np.zeros(2),
                              "R": [np.eye(3)] * 2, "t": [np.ones((3, 1))] * 2})


def _pipeline_call_sequence(ev, c, capsys):
    """What src/pipeline.py:156-163 does with the evaluator: evaluate_sequence -> print_summary ->
    create_comparison_dataframe.  Returns (printed text, dataframe)."""
    frames = [int(f) for f in c["frames"]]
    res = ev.evaluate_sequence({"frames": frames, "roll": c["est_euler"][:, 2], "pitch": c["est_euler"][:, 1],
                                "yaw": c["est_euler"][:, 0], "R": list(c["est_R"]), "t": list(c["est_t"])})
    capsys.readouterr()
    ev.print_summary(res)
    text = capsys.readouterr().out
    return text, ev.create_comparison_dataframe(res), res


class _TableGT:
    def __init__(self, c):
        self.c, self.frames = c, [int(f) for f in c["frames"]]

    def get_pose(self, f):
        i, c = self.frames.index(f), self.c
        return {"x": c["gt_pos"][i, 0], "y": c["gt_pos"][i, 1], "z": c["gt_pos"][i, 2],
                "yaw": c["gt_euler"][i, 0], "pitch": c["gt_euler"][i, 1], "roll": c["gt_euler"][i, 2]}


def check_report_against_reference(text, df, res):
    """The console summary and the table columns of the UNMODIFIED reference for case 0
    (tests/golden/make_golden_eval.py -> eval_report_c0*.txt)."""
    assert text == open(os.path.join(GOLDEN, "eval_report_c0.txt")).read()
    assert ",".join(df.columns) == open(os.path.join(GOLDEN, "eval_report_c0_columns.txt")).read().strip()
    assert len(df) == len(res["frames"]) and list(df["frame"]) == list(res["frames"])
    assert np.array_equal(df["rotation_error"].to_numpy(), res["rotation_error"])


def test_reporting_methods_reproduce_the_reference_text(capsys):
    """Host-side logic of the evaluator twin (print_summary / create_comparison_dataframe / _wrap_angle_error) with
    the device call replaced by the oracle (TEST ONLY: the product has no such path)."""
    import rpe_b200
    c = list(cases())[0]

    class OracleEngine:
        def evaluate_sequence(self, est_R, est_t, est_euler, gt_euler, gt_pos, conv):
            e = oe.evaluate_sequence(est_R.reshape(-1, 3, 3), est_t, est_euler, gt_euler, gt_pos, "zyx" if conv else "yup")
            return e, oe.summary_statistics(e)

        def summarize_errors(self, errors):
            return oe.summary_statistics(errors)
    ev = rpe_b200.PoseEvaluator(_TableGT(c), euler_convention="yup", engine=OracleEngine())
    text, df, res = _pipeline_call_sequence(ev, c, capsys)
    check_report_against_reference(text, df, res)
    assert ev._wrap_angle_error(350.0) == 10.0 and ev._wrap_angle_error(-190.0) == 170.0 and ev._wrap_angle_error(180.0) == 180.0
