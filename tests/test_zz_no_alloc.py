"""rpe_evaluate_sequence_noalloc (include/rpe.h): the allocation-free form of rpe_evaluate_sequence for callers that
may not call cudaMalloc.  No counterpart in the reference; the checker is the existing entry point.

Called straight through the C ABI with caller-owned device buffers (torch tensors are only the caller's memory here),
twice in a row on the shared static scratch: a long sequence first, then a SHORTER one with other values, so that the
second result would be wrong if rows left over from the first call were read.  Tolerance: the one
tests/test_gpu_evaluator.py uses for the summary (mean / std 1e-9 relative; median / max / min exact) — and, since
both entry points launch the same two kernels, bit equality is asserted as well.

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _sequence(n, seed):
    from oracle import evaluator as oe
    rng = np.random.default_rng(seed)
    gt_euler = np.c_[rng.uniform(-179, 179, n), rng.uniform(-80, 80, n), rng.uniform(-179, 179, n)]
    est_euler = gt_euler + rng.normal(0, 1.5, (n, 3))
    gt_pos = np.cumsum(rng.normal(0, 0.3, (n, 3)), axis=0)
    step = max(1, n // 64)
    est_R = np.repeat(np.stack([oe.euler_to_rotation(*est_euler[i]) for i in range(0, n, step)]), step, axis=0)[:n]
    est_t = rng.normal(0, 1, (n, 3))
    return [np.ascontiguousarray(a, dtype=np.float64) for a in (est_R.reshape(n, 9), est_t, est_euler, gt_euler, gt_pos)]


def _noalloc(lib, L, arrays, conv, torch, stream):
    n = arrays[0].shape[0]
    dev = [torch.from_numpy(a).to("cuda:0") for a in arrays]            # the CALLER's buffers
    d_summary = torch.full((25,), float("nan"), dtype=torch.float64, device="cuda:0")
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    with torch.cuda.stream(stream):
        rc = lib.rpe_evaluate_sequence_noalloc(n, *[C.c_void_p(t.data_ptr()) for t in dev], conv,
                                               C.c_void_p(d_summary.data_ptr()), C.c_void_p(stream.cuda_stream))
    L.check(rc, "rpe_evaluate_sequence_noalloc")
    stream.synchronize()
    assert torch.cuda.memory_allocated() == before
    return d_summary.cpu().numpy().reshape(5, 5)


def test_noalloc_entry_point_equals_evaluate_sequence_twice_in_a_row():
    import torch
    import rpe_b200
    from rpe_b200 import _lib as L
    lib = L.load()
    torch.cuda.set_device(0)
    eng = rpe_b200.Engine(rpe_b200.EngineConfig(max_width=64, max_height=64, max_pairs=1, nfeatures=64, nlevels=1))
    stream = torch.cuda.Stream(device="cuda:0")
    try:
        # largest size the static scratch takes, then a shorter sequence in the other convention
        for n, seed, conv in ((16384, 11, L.EULER_YUP), (777, 12, L.EULER_ZYX)):
            arrays = _sequence(n, seed)
            want_errors, want = eng.evaluate_sequence(*arrays, conv)
            got = _noalloc(lib, L, arrays, conv, torch, stream)
            assert np.allclose(got[:, :2], want[:, :2], rtol=1e-9, atol=1e-12), (n, got, want)
            assert np.array_equal(got[:, 2:], want[:, 2:]), (n, got, want)
            assert np.array_equal(got, want)                      # same kernels on the same table
            assert np.array_equal(got[:, 3], want_errors.max(axis=0))
    finally:
        eng.close()


def test_noalloc_entry_point_rejects_what_it_cannot_serve():
    import torch
    from rpe_b200 import _lib as L
    lib = L.load()
    torch.cuda.set_device(0)
    d = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    p = C.c_void_p(d.data_ptr())
    rc = lib.rpe_evaluate_sequence_noalloc(16385, p, p, p, p, p, L.EULER_YUP, p, None)
    assert rc == -4 and b"RPE_NOALLOC_MAX_FRAMES = 16384" in lib.rpe_last_error()
    host = np.zeros(64)
    rc = lib.rpe_evaluate_sequence_noalloc(1, C.c_void_p(host.ctypes.data), p, p, p, p, L.EULER_YUP, p, None)
    assert rc == -1 and b"device pointers" in lib.rpe_last_error()
    assert lib.rpe_evaluate_sequence_noalloc(0, p, p, p, p, p, L.EULER_YUP, p, None) == -1
    assert lib.rpe_evaluate_sequence_noalloc(1, p, p, p, p, p, 7, p, None) == -1
    torch.cuda.synchronize()
