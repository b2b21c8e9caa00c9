"""Multi-process path on CPU: pair sharding + host gather with torch.distributed (gloo, world 2),
and the batched BatchProcessor twin against the reference's loop semantics (fake estimator)."""
import os
import socket
import sys

import numpy as np
import pytest
from conftest import ROOT

from rpe_b200 import sharding as sh
from rpe_b200.batch_processor import BatchProcessor
from rpe_b200.geometry import euler_to_rotation, rotation_to_euler


def test_shard_range_partitions_exactly():
    for n in (0, 1, 7, 58, 100000):
        for world in (1, 2, 3, 4, 8):
            spans = [sh.shard_range(n, world, r) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        sh.shard_range(10, 2, 2)
    frames = list(range(0, 150, 15))                       # 10 frames -> 9 pairs
    assert sh.shard_frames(frames, 2, 0) == frames[0:5] and sh.shard_frames(frames, 2, 1) == frames[4:10]
    assert sh.shard_frames([0], 2, 0) == []


def _fake_out(lo, hi):
    idx = np.arange(lo, hi, dtype=np.float64)
    B = hi - lo
    return {"R": (idx[:, None, None] + np.arange(9).reshape(3, 3)[None]), "t": idx[:, None, None] * np.ones((B, 3, 1)),
            "euler": np.stack([idx, -idx, 2 * idx], 1), "status": (np.arange(lo, hi) % 4).astype(np.int32)}


def test_pack_roundtrip_and_single_process_gather():
    out = _fake_out(3, 11)
    rows = sh.pack_results(out)
    back = sh.unpack_results(rows)
    for k in out:
        assert np.array_equal(back[k], out[k]), k
    assert sh.gather_results(rows, 8) is rows


def _worker(rank, world, port, n_pairs, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        lo, hi = sh.shard_range(n_pairs, world, rank)
        got = sh.gather_results(sh.pack_results(_fake_out(lo, hi)), n_pairs)
        q.put((rank, None if got is None else got.tolist()))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_pairs", [7, 58])
def test_gloo_world2_gather_in_pair_order(n_pairs):
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n_pairs, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[1] is None
    want = sh.pack_results(_fake_out(0, n_pairs))
    assert np.array_equal(np.array(res[0]), want)


class _FakeGT:
    def __init__(self, n):
        self.poses = {15 * i: {"yaw": 3.0 * i, "pitch": -1.0 * i, "roll": 0.5 * i} for i in range(n)}

    def get_pose(self, f):
        return self.poses[f]

    def get_frame_indices(self, step=15):
        return [f for f in sorted(self.poses) if f % step == 0]


class _FakeEstimator:
    """estimate_batch / estimate_sequence with a deterministic R_rel per pair; status 2 on request."""
    def __init__(self, fail_at=None):
        self.fail_at, self.calls = fail_at, 0

    def estimate_batch(self, a, b):
        B = len(a)
        R = np.stack([euler_to_rotation(1.0 + int(x[0, 0]), 0.5, -0.25) for x in a])
        st = np.zeros(B, np.int32)
        if self.fail_at is not None:
            st[[i for i, x in enumerate(a) if int(x[0, 0]) == self.fail_at]] = 2
        self.last = {"info": np.full((B, 4), 3, np.int32)}
        self.calls += 1
        return R, np.tile(np.array([[0.0], [0.0], [1.0]]), (B, 1, 1)), st

    def estimate_sequence(self, frames):
        return self.estimate_batch(frames[:-1], frames[1:])

    def _raise_for(self, status, n):
        if int(status) & 0xFF == 2:
            raise RuntimeError(f"Insufficient matches: {n} (minimum 5 required)")


def test_batch_processor_twin(tmp_path, monkeypatch):
    import rpe_b200.batch_processor as bp
    n = 7
    monkeypatch.setattr(bp, "load_image", lambda p, to_gray=True: np.full((4, 4), int(os.path.basename(str(p))[:6]) // 15, np.uint8))
    gt = _FakeGT(n)
    est = _FakeEstimator()
    proc = BatchProcessor(tmp_path, est, gt, "yup", chunk_pairs=4)
    out = proc.process_at_interval(15)
    assert out["frames"] == [15 * i for i in range(1, n)] and est.calls == 2
    for k, f in enumerate(out["frames"]):                  # the reference loop, pair by pair (batch_processor.py:82-101)
        g = gt.get_pose(f - 15)
        Rw = euler_to_rotation(g["yaw"], g["pitch"], g["roll"], "yup") @ euler_to_rotation(1.0 + k, 0.5, -0.25)
        assert np.allclose(out["R"][k], Rw)
        assert np.allclose((out["yaw"][k], out["pitch"][k], out["roll"][k]), rotation_to_euler(Rw, "yup"))
    assert out["roll"].shape == (n - 1,) and proc.get_image_path(15).name == "000015.png"
    with pytest.raises(ValueError, match="at least 2 frames"):
        proc.process_sequence([0])
    with pytest.raises(RuntimeError, match="Insufficient matches: 3"):
        BatchProcessor(tmp_path, _FakeEstimator(fail_at=2), gt).process_at_interval(15)


def test_numa_binding_helper_is_safe_without_a_gpu(tmp_path):
    """sharding.bind_host_to_gpu: cpulist parsing, and a no-op ([]) for an unknown PCI device or a node that
    already covers the whole affinity mask."""
    from rpe_b200 import sharding
    assert sharding._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert sharding._parse_cpulist("") == set()
    import os
    before = os.sched_getaffinity(0)
    assert sharding.bind_host_to_gpu("ffff:ff:1f.0") == []
    assert os.sched_getaffinity(0) == before
