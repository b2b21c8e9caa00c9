"""The CPU arm of bench.py: the reference's own cv2 call sequence (oracle/reference_path.py = the four cv2 calls
PoseEstimator.estimate makes, with the same arguments) timed on the host cores in the two modes SURVEY 8(d) names:
cv2's own thread pool, and one single-threaded process per core with the pairs sharded over the processes (the
CPU's best mode for a stream of independent pairs).  Same synthetic pairs as the GPU arm: pair indices
first_pair, first_pair + 1, ... of the workload.  Bench support only."""
from __future__ import annotations

import multiprocessing as mp
import queue as _queue
import time

from .hostgen import host_cores


def _path(cfg):
    from bench_support.synth import synth_K
    from oracle.reference_path import PathConfig, ReferencePath
    return ReferencePath(synth_K(cfg["W"], cfg["H"]),
                         PathConfig(nfeatures=cfg["nfeatures"], nlevels=cfg["nlevels"], match_mode=cfg["match_mode"],
                                    ratio=cfg["ratio"], max_matches=cfg["max_matches"],
                                    ransac_max_iters=cfg["ransac_max_iters"]))


def _time_pairs(ref, pairs):
    ok = 0
    t0 = time.perf_counter()
    for a, b in pairs:
        try:
            ref.estimate(a, b)
            ok += 1
        except RuntimeError:
            pass
    return time.perf_counter() - t0, ok


def rate_cv2_threads(cfg, n_pairs: int, first_pair: int = 0):
    """cv2 with all its threads, pairs one after the other -> (pairs/s, seconds, threads)."""
    import cv2
    from bench_support.synth import synth_pair
    cv2.setNumThreads(-1)
    ref = _path(cfg)
    pairs = [synth_pair(cfg["synth_cfg"], first_pair + i, cfg["W"], cfg["H"])[:2] for i in range(n_pairs)]
    ref.estimate(*pairs[0])                       # warm-up call
    dt, _ = _time_pairs(ref, pairs)
    return n_pairs / dt, dt, cv2.getNumThreads()


def _shard_worker(wid, cfg, steps, n, first_pair, barrier, q):
    import cv2
    from bench_support.synth import synth_pair
    cv2.setNumThreads(1)
    ref = _path(cfg)
    pairs = [synth_pair(cfg["synth_cfg"], first_pair + wid * n + i, cfg["W"], cfg["H"])[:2] for i in range(n)]
    ref.estimate(*pairs[0])                       # warm-up call
    walls = []
    for _ in range(steps):
        barrier.wait()
        walls.append(_time_pairs(ref, pairs)[0])
    q.put((wid, walls))


def rate_process_per_core(cfg, steps: int, n_per_proc: int, first_pair: int = 0, timeout_s: float = 600.0):
    """One single-threaded process per host core; process w takes pairs first_pair + w * n .. + n - 1; every step
    starts on a barrier and lasts until the slowest core is done -> (pairs/s, processes, seconds)."""
    procs = host_cores()
    ctx = mp.get_context("spawn")                 # CUDA may be initialised in this process: never fork
    barrier, q = ctx.Barrier(procs), ctx.Queue()
    ps = [ctx.Process(target=_shard_worker, args=(w, cfg, steps, n_per_proc, first_pair, barrier, q), daemon=True)
          for w in range(procs)]
    for p_ in ps:
        p_.start()
    res, t_end = [], time.perf_counter() + timeout_s
    while len(res) < procs:
        try:
            res.append(q.get(timeout=2))
        except _queue.Empty:
            dead = [p_.exitcode for p_ in ps if p_.exitcode not in (None, 0)]
            if dead or time.perf_counter() > t_end:      # a worker died (or hangs): do not wait for it
                for p_ in ps:
                    if p_.is_alive():
                        p_.terminate()
                raise RuntimeError(f"process-per-core CPU sample failed (exit codes {dead})")
    for p_ in ps:
        p_.join(timeout=60)
    wall = sum(max(r[1][s] for r in res) for s in range(steps))
    return steps * n_per_proc * procs / wall, procs, wall


def cpu_baseline(cfg, name: str, threads_pairs: int, per_proc_pairs: int, steps: int = 1):
    """Both modes on a bounded sample; the better one is the reported value."""
    rate, dt, threads = rate_cv2_threads(cfg, threads_pairs)
    out = {"value": rate, "unit": "pairs/s", "cores": threads, "kind": "port",
           "sample": f"pairs 0..{threads_pairs - 1} of the {name} workload in {dt:.1f}s, cv2 with its own {threads} "
                     f"threads (same cv2 calls as PoseEstimator.estimate)",
           "modes": {"cv2_threads": rate}}
    try:
        r2, procs, wall2 = rate_process_per_core(cfg, steps, per_proc_pairs)
        out["modes"]["process_per_core"] = r2
        if r2 > rate:
            out.update(value=r2, cores=procs,
                       sample=f"{procs} single-threaded processes x {per_proc_pairs} pairs (pairs 0..{procs * per_proc_pairs - 1} "
                              f"of the {name} workload) x {steps} step(s) in {wall2:.1f}s (same cv2 calls as "
                              f"PoseEstimator.estimate); cv2's own {threads} threads: {rate:.1f} pairs/s")
    except Exception as e:      # the thread mode above always exists
        out["modes"]["process_per_core_error"] = repr(e)[:200]
    return out
