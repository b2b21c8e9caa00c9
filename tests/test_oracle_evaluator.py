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
