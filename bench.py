#!/usr/bin/env python
"""bench.py — pose pairs/sec of the two-view hot path on N B200s (BASELINE.json metric).

A "step" is one pass of the whole hot path (ORB x2 -> Hamming match -> 5-pt RANSAC -> recoverPose
-> Euler) over one batch of synthetic pairs per GPU.  Workload = BASELINE config 3
("synthetic 1280x720 pair batch, ORB 5000 features, BF-Hamming ratio 0.75, RANSAC 5-pt 2000
iters"; other ORB parameters as the reference hard-codes them; max_matches 500).

  value : device-timed throughput, inputs resident in HBM (rpe_estimate_batch_device)
  e2e   : same metric through the host-buffer C ABI call (rpe_estimate_batch): pinned host
          images -> H2D -> kernels -> D2H of R|t|euler|status, all inside the timed region
  --impl reference : the reference's own CPU path (the same cv2 calls PoseEstimator.estimate
          makes; oracle/reference_path.py) on the host cores, bounded sample per step.

Pairs are independent, so N GPUs = N shards with no data-path collective ("scaling": "weak").
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CFG = dict(W=1280, H=720, nfeatures=5000, nlevels=12, match_mode="ratio", ratio=0.75, max_matches=500,
           ransac_max_iters=2000, synth_cfg=3)
METRIC = "pose pairs/sec (ORB 5k -> BF-Hamming ratio 0.75 -> 5-pt RANSAC 2000 -> recoverPose), synthetic 1280x720"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="pairs per GPU per step (default 256, as SURVEY 8d suggests for config 3; config 4: 16)")
    ap.add_argument("--config", type=int, default=3, choices=[3, 4],
                    help="BASELINE.json config: 3 = 1280x720 / 5000 features (the headline workload, default); "
                         "4 = 3840x2160 / 20000 features / 8 levels (stress line item, not the bench line)")
    ap.add_argument("--cpu-pairs", type=int, default=24, help="pairs in the bounded CPU sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--inflight", type=int, default=1, help="engine handles used alternately")
    ap.add_argument("--depth", type=int, default=2, help="batches pending per handle (rpe_submit_* / rpe_wait_batch; max 2)")
    ap.add_argument("--chunk", type=int, default=32, help="pairs per pipeline chunk (rpe_set_pipeline)")
    ap.add_argument("--streams", type=int, default=1, help="pipeline streams per handle (rpe_set_pipeline); 1 = no chunking")
    ap.add_argument("--no-stream", action="store_true", help="skip the extra consecutive-frame (config 5) pass")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.lines: list[str] = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# ------------------------------------------------------------------------------------------ data
def make_batch(rank: int, B: int):
    from bench_support.synth import synth_batch, synth_K   # input generator (host), not the checker
    frames = synth_batch(CFG["synth_cfg"], rank * 100000, B, CFG["W"], CFG["H"])
    return frames, synth_K(CFG["W"], CFG["H"])


def level_pixels():
    """Pyramid pixel counts for the algorithmic-byte model."""
    W, H, L = CFG["W"], CFG["H"], CFG["nlevels"]
    sf = float(np.float32(1.1))
    px = []
    for l in range(L):
        s = np.float32(sf ** l)
        px.append(int(np.rint(np.float32(W) / s)) * int(np.rint(np.float32(H) / s)))
    return px


def stage_models(B: int, nkp_mean: float):
    """Algorithmic bytes per STEP (B pairs = 2B images) for each stage."""
    px = level_pixels()
    P, P0, Plast = sum(px), px[0], px[-1]
    n_img = 2 * B
    N = nkp_mean
    return {
        "pyramid": n_img * ((P - Plast) + (P - P0)),
        "fast": n_img * P,
        "blur": n_img * 2 * P,
        "describe": n_img * N * (32 + 24),
        "match": B * (32 * 2 * N + 8 * 2 * N),
    }


# ------------------------------------------------------------------------------------------ reference arm
def cpu_reference_rate(n_pairs: int, first_pair: int = 900000):
    """The reference's CPU path (cv2, all host threads) on a bounded sample of the same workload."""
    import cv2
    from oracle.reference_path import PathConfig, ReferencePath
    from bench_support.synth import synth_K, synth_pair
    cv2.setNumThreads(-1)
    K = synth_K(CFG["W"], CFG["H"])
    ref = ReferencePath(K, PathConfig(nfeatures=CFG["nfeatures"], nlevels=CFG["nlevels"], match_mode=CFG["match_mode"],
                                      ratio=CFG["ratio"], max_matches=CFG["max_matches"],
                                      ransac_max_iters=CFG["ransac_max_iters"]))
    pairs = [synth_pair(CFG["synth_cfg"], first_pair + i, CFG["W"], CFG["H"])[:2] for i in range(n_pairs)]
    ref.estimate(*pairs[0])                       # warm-up call
    t0 = time.perf_counter()
    ok = 0
    for a, b in pairs:
        try:
            ref.estimate(a, b)
            ok += 1
        except RuntimeError:
            pass
    dt = time.perf_counter() - t0
    return n_pairs / dt, dt, cv2.getNumThreads(), ok


def _shard_worker(wid, steps, n, barrier, q):
    """One host core of the process-per-core mode (SURVEY 8d ii): cv2 single-threaded, own pairs, timed loops
    started together on a barrier."""
    import cv2
    from oracle.reference_path import PathConfig, ReferencePath
    from bench_support.synth import synth_K, synth_pair
    cv2.setNumThreads(1)
    K = synth_K(CFG["W"], CFG["H"])
    ref = ReferencePath(K, PathConfig(nfeatures=CFG["nfeatures"], nlevels=CFG["nlevels"], match_mode=CFG["match_mode"],
                                      ratio=CFG["ratio"], max_matches=CFG["max_matches"],
                                      ransac_max_iters=CFG["ransac_max_iters"]))
    pairs = [synth_pair(CFG["synth_cfg"], 800000 + wid * 64 + i, CFG["W"], CFG["H"])[:2] for i in range(n)]
    ref.estimate(*pairs[0])                       # warm-up call
    walls = []
    for _ in range(steps):
        barrier.wait()
        t0 = time.perf_counter()
        for a, b in pairs:
            try:
                ref.estimate(a, b)
            except RuntimeError:
                pass
        walls.append(time.perf_counter() - t0)
    q.put((wid, walls))


def cpu_reference_rate_sharded(steps: int, n_per_proc: int):
    """The reference's CPU path as one single-threaded process per host core, pairs sharded over the processes:
    the CPU's best throughput mode for a stream of independent pairs.  Returns (pairs/s, procs, seconds)."""
    import multiprocessing as mp
    procs = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    ctx = mp.get_context("spawn")                 # CUDA may be initialised in this process: never fork
    barrier, q = ctx.Barrier(procs), ctx.Queue()
    ps = [ctx.Process(target=_shard_worker, args=(w, steps, n_per_proc, barrier, q), daemon=True) for w in range(procs)]
    for p_ in ps:
        p_.start()
    import queue as _queue
    res, t_end = [], time.perf_counter() + 300
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
    wall = sum(max(r[1][s] for r in res) for s in range(steps))     # per step: until the slowest core is done
    return steps * n_per_proc * procs / wall, procs, wall


def cv2_threads_str():
    import cv2
    return str(cv2.getNumThreads())


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rates, t_all = [], 0.0
    n = max(2, args.cpu_pairs // 4)
    for _ in range(args.warmup):
        cpu_reference_rate(2)
    t0 = time.perf_counter()
    for s in range(args.steps):
        r, dt, threads, _ = cpu_reference_rate(n, 900000 + s * n)
        rates.append(r); t_all += dt
    value = args.steps * n / t_all
    sample = f"{n} synthetic 1280x720 pairs per step x {args.steps} steps, cv2 threads={threads}"
    modes = {"cv2_threads": value}
    try:
        v2, procs, wall2 = cpu_reference_rate_sharded(args.steps, 2)
        modes["process_per_core"] = v2
        if v2 > value:      # report the CPU's better mode
            value, threads, t_all = v2, procs, wall2
            n = 2 * procs
            sample = (f"{procs} single-threaded processes x 2 synthetic 1280x720 pairs per step x {args.steps} steps "
                      f"(cv2 {modes['cv2_threads']:.1f} pairs/s with its own {cv2_threads_str()} threads)")
    except Exception as e:      # the thread mode above always exists
        modes["process_per_core_error"] = repr(e)[:200]
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8/f32/f64", "data": "synthetic",
           "config": {"workload": workload_name(),
                      "pairs_per_step": n, **{k: CFG[k] for k in ("nfeatures", "nlevels", "match_mode", "max_matches")}},
           "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample,
                            "modes": modes},
           "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch
    import torch.distributed as dist

    import rpe_b200
    from rpe_b200 import _lib as L

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the CUDA library has no CPU fallback")
    torch.cuda.set_device(local)
    numa_cpus = []
    if world > 1 and os.environ.get("RPE_NUMA_BIND", "1") != "0":
        # one process per GPU: run on, and allocate the pinned frame buffers from, the GPU's own NUMA node
        try:
            numa_cpus = rpe_b200.sharding.bind_host_to_gpu(rpe_b200.sharding.gpu_pci_bus_id(local))
        except Exception:
            numa_cpus = []
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B = args.batch
    W, H = CFG["W"], CFG["H"]

    frames_np, K = make_batch(rank, B)
    host = torch.from_numpy(frames_np).pin_memory()
    dev = host.cuda(non_blocking=False)
    ecfg = rpe_b200.EngineConfig(max_width=W, max_height=H, max_pairs=B, nfeatures=CFG["nfeatures"],
                                 nlevels=CFG["nlevels"], match_mode=CFG["match_mode"], ratio=CFG["ratio"],
                                 max_matches=CFG["max_matches"], ransac_max_iters=CFG["ransac_max_iters"])
    # batch n+1 is submitted before the results of batch n are waited for (rpe_submit_* / rpe_wait_batch, two
    # pending per handle): its upload and ORB / matcher kernels run under the RANSAC / pose tail of batch n
    engines = [rpe_b200.Engine(ecfg, device=local) for _ in range(max(1, args.inflight))]
    eng = engines[0]
    stream = torch.cuda.Stream()
    streams = [stream] + [torch.cuda.Stream() for _ in engines[1:]]
    for e, st in zip(engines, streams):
        e.set_stream(st.cuda_stream)
        e.set_pipeline(args.chunk, args.streams)
    host_np = host.numpy()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def run_steps(n, submit):
        """n steps over the handles, at most len(engines) batches in flight; returns launches, last result"""
        launches, out, q = 0, None, []
        cap = len(engines) * max(1, min(2, args.depth))
        for i in range(n):
            e = engines[i % len(engines)]
            if len(q) == cap:
                d = q.pop(0)
                out = d.wait()
                launches += sum(d.stage_times()[1].values())
            submit(e)
            q.append(e)
        for d in q:
            out = d.wait()
            launches += sum(d.stage_times()[1].values())
        return launches, out

    submit_device = lambda e: e.submit_device(dev.data_ptr(), H * W, B, H, W, W, K)
    submit_host = lambda e: e.submit_interleaved(host_np, K)

    _, out = run_steps(max(args.warmup, len(engines)), submit_device)
    nkp = float(np.mean([eng.keypoints(i, v)["x"].shape[0] for i in range(min(B, 4)) for v in (0, 1)]))
    ok_frac = float((out["status"] & 0xFF == 0).mean())

    def timed(submit):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        launches, _ = run_steps(args.steps, submit)
        torch.cuda.synchronize()
        e1.record(stream)
        barrier()
        ms_total = e0.elapsed_time(e1)
        if world > 1:
            tt = torch.tensor([ms_total], device="cuda")
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            ms_total = float(tt.item())
        return ms_total, launches

    def profiled():
        """per-stage CUDA events: one handle, one chunk, stages serial (rpe_set_profiling)"""
        eng.set_profiling(True)
        acc = {k: 0.0 for k in L.RPE_T_NAMES}
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(args.steps):
            eng.estimate_batch_device(dev.data_ptr(), H * W, B, H, W, W, K)
            ms, _ = eng.stage_times()
            for k in acc:
                acc[k] += ms[k]
        e1.record(stream)
        barrier()
        eng.set_profiling(False)
        return e0.elapsed_time(e1), acc

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms_dev, launches = timed(submit_device)                   # headline: frames resident in HBM
    run_steps(max(2, len(engines)), submit_host)
    ms_e2e, _ = timed(submit_host)                            # pinned host frames -> results on host
    ms_prof, stage_ms = profiled()
    clocks = sampler.stop() if rank == 0 else None

    # extra line item (not the headline): BASELINE config 5, a video stream of consecutive frames through the
    # sequence ABI (rpe_submit_sequence*: ORB once per frame).  One step = B + 1 frames -> B pairs; a caller
    # chains steps by starting at the previous step's last frame, so the steady-state rate is B pairs / step.
    stream = None
    if not args.no_stream:
        from bench_support.synth import synth_stream
        F = B + 1
        shost = torch.from_numpy(synth_stream(5, rank * 100000, F, W, H)).pin_memory()
        sdev = shost.cuda(non_blocking=False)
        shost_np = shost.numpy()
        sub_dev = lambda e: e.submit_sequence_device(sdev.data_ptr(), H * W, F, H, W, W, K)
        sub_host = lambda e: e.submit_sequence(shost_np, K)
        _, sout = run_steps(max(2, len(engines)), sub_dev)
        ms_sdev, _ = timed(sub_dev)
        run_steps(max(2, len(engines)), sub_host)
        ms_se2e, _ = timed(sub_host)
        npairs = world * args.steps * (F - 1)
        stream = {"workload": "cfg5: synthetic 1280x720 video stream, consecutive frame pairs, cfg3 parameters, "
                              "frame-level feature cache (ORB once per frame)",
                  "value": npairs / (ms_sdev * 1e-3), "e2e": npairs / (ms_se2e * 1e-3), "unit": "pairs/s",
                  "frames_per_gpu_per_step": F, "pairs_per_gpu_per_step": F - 1,
                  "ms_per_step": ms_sdev / args.steps, "e2e_ms_per_step": ms_se2e / args.steps,
                  "ok_fraction": float((sout["status"] & 0xFF == 0).mean())}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    total_pairs = world * args.steps * B
    value = total_pairs / (ms_dev * 1e-3)
    e2e = total_pairs / (ms_e2e * 1e-3)

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    model = stage_models(B, nkp)
    stages = {}
    for k, byts in model.items():
        ms = stage_ms[k] / args.steps
        gbs = byts / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        stages[k] = {"ms_per_step": ms, "algorithmic_bytes_per_step": int(byts), "gbs": gbs, "frac_hbm": gbs / hbm_peak}
    for k in ("select", "match_sort", "ransac", "pose", "h2d", "d2h"):
        stages[k] = {"ms_per_step": stage_ms[k] / args.steps}
    # matcher extras: operand-stream-equivalent bytes and popc32 rate
    mm = stages["match"]["ms_per_step"] * 1e-3
    if mm > 0:
        stages["match"]["operand_stream_gbs"] = B * 64 * nkp * nkp / mm / 1e9
        stages["match"]["popc32_per_s"] = B * 8 * nkp * nkp / mm
    kernel_stages = ["pyramid", "fast", "select", "blur", "describe", "match", "match_sort", "ransac", "pose"]
    dom = max(kernel_stages, key=lambda k: stages[k]["ms_per_step"])
    # roofline object: the slowest HBM-bound kernel (pyramid / FAST / blur stream the images).  The
    # matcher, when dominant, is bound by the popc (XU) pipe, not by HBM: its pipe fraction is in `xu_pipe`.
    rf_stage = max(("pyramid", "fast", "blur"), key=lambda k: stages[k]["ms_per_step"])
    traffic = None      # DRAM bytes per launch from the committed ncu --set full capture, scaled to this step's images
    try:
        per_img = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))[rf_stage]["dram_bytes_per_image"]
        traffic = per_img * 2 * B if CFG["synth_cfg"] == 3 else None      # the capture is of the cfg3 workload
    except Exception:
        pass
    sm_hz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    popc_peak = 148 * 16 * sm_hz * 1e6          # 16 POPC lanes / clk / SM (XU pipe), 148 SMs
    # the kernel issues 4.5 POPC per 256-bit distance (carry-save adders fold the 8 xor words into 5 / 4 words on
    # alternate train descriptors), not 8: the pipe fraction counts ISSUED popcounts
    popc_issued = 4.5 / 8.0 * (stages["match"].get("popc32_per_s") or 0.0)
    xu = {"kernel": "match", "distances_per_s": (stages["match"].get("popc32_per_s") or 0.0) / 8.0,
          "popc_issued_per_distance": 4.5, "achieved_popc32_per_s": popc_issued,
          "plain_form_equivalent_popc32_per_s": stages["match"].get("popc32_per_s"),
          "peak_popc32_per_s": popc_peak, "frac": popc_issued / popc_peak,
          "note": "peak = 148 SMs x 16 popc/clk x median SM clock; ncu sm__inst_executed_pipe_xu = 93 % and pipe_alu = "
                  "83 % on this kernel (profiles/r01_ncu_full_b32_summary_v4.txt)"}
    roofline = {"kernel": rf_stage, "bound": "hbm", "achieved": stages[rf_stage]["gbs"], "peak": hbm_peak,
                "unit": "GB/s", "frac": stages[rf_stage]["frac_hbm"], "traffic": traffic, "peak_source": peak_src,
                "xu_pipe": xu, "dominant_stage": dom, "share_of_step": stages[dom]["ms_per_step"] / (ms_prof / args.steps),
                "serial_ms_per_step": ms_prof / args.steps,
                "note": "achieved = algorithmic bytes per step (SURVEY 8d) / CUDA-event time of that stage, measured "
                        "in a separate pass of the same K steps with an event between stages (stages serial, one "
                        "chunk); `value` is the pipelined pass. The matcher is XU-pipe (popc) bound, see "
                        "stages.match.popc32_per_s"}

    cpu = None
    if world == 1 and not args.no_cpu_baseline:
        rate, dt, threads, ok = cpu_reference_rate(args.cpu_pairs)
        cpu = {"value": rate, "unit": "pairs/s", "cores": threads, "kind": "port",
               "sample": f"{args.cpu_pairs} synthetic 1280x720 cfg3 pairs in {dt:.1f}s, cv2 {threads} threads "
                         f"(same cv2 calls as PoseEstimator.estimate)", "modes": {"cv2_threads": rate}}
        try:        # the CPU's other mode: one single-threaded process per core (SURVEY 8d ii); the better one is reported
            r2, procs, wall2 = cpu_reference_rate_sharded(1, 8)
            cpu["modes"]["process_per_core"] = r2
            if r2 > rate:
                cpu.update(value=r2, cores=procs,
                           sample=f"{procs} single-threaded processes x 8 synthetic 1280x720 cfg3 pairs in {wall2:.1f}s "
                                  f"(same cv2 calls as PoseEstimator.estimate); cv2's own {threads} threads: {rate:.1f} pairs/s")
        except Exception as e:
            cpu["modes"]["process_per_core_error"] = repr(e)[:200]

    out = {"metric": METRIC, "value": value, "unit": "pairs/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms_dev / args.steps, "higher_is_better": True, "scaling": "weak",
           "vs_baseline": None, "dtype": "u8/i32 (ORB, match), f32 (Harris/angle/blur), f64 (RANSAC, pose)",
           "data": "synthetic",
           "config": {"workload": workload_name(),
                      "pairs_per_gpu_per_step": B, "image_bytes_per_step": int(2 * B * H * W),
                      "l2_policy": f"inputs larger than L2 ({2 * B * H * W / 1e6:.0f} MB of frames + "
                                   f"{2 * B * sum(level_pixels()) / 1e9:.1f} GB of pyramids per step)",
                      "parallelism": f"{world} independent shards, no collective",
                      "host_numa_binding": f"{len(numa_cpus)} cpus of the GPU's node" if numa_cpus else "none",
                      "pipeline": {"handles": len(engines), "batches_pending_per_handle": max(1, min(2, args.depth)),
                                   "chunk_pairs": args.chunk, "streams": args.streams},
                      "mean_keypoints_per_image": nkp, "ok_fraction": ok_frac},
           "e2e": {"value": e2e, "unit": "pairs/s", "h2d_bytes_per_step": int(2 * B * H * W + 72),
                   "d2h_bytes_per_step": int(B * (15 * 8 + 3 * 4)), "ms_per_step": ms_e2e / args.steps},
           "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks,
           "stages": stages, "stream": stream}
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


def workload_name():
    return (f"cfg{CFG['synth_cfg']}: synthetic {CFG['W']}x{CFG['H']} pair batch, ORB {CFG['nfeatures']} features (sf 1.1, "
            f"{CFG['nlevels']} levels, FAST 15, HARRIS), BF-Hamming knn2 + ratio 0.75, max_matches {CFG['max_matches']}, "
            f"RANSAC 5-pt {CFG['ransac_max_iters']} iters")


def main():
    global METRIC
    args = parse()
    if args.config == 4:
        CFG.update(W=3840, H=2160, nfeatures=20000, nlevels=8, synth_cfg=4)
        METRIC = METRIC.replace("ORB 5k", "ORB 20k").replace("1280x720", "3840x2160")
        args.no_stream = True
    if args.batch is None:
        args.batch = 256 if args.config == 3 else 16
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
