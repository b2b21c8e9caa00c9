#!/usr/bin/env python
"""bench.py — pose pairs/sec of the two-view hot path on N B200s (BASELINE.json metric).

A "step" is one pass of the whole hot path (ORB x2 -> Hamming match -> 5-pt RANSAC -> recoverPose
-> Euler) over one batch of synthetic pairs per GPU.  Workload = BASELINE config 3
("synthetic 1280x720 pair batch, ORB 5000 features, BF-Hamming ratio 0.75, RANSAC 5-pt 2000
iters"; other ORB parameters as the reference hard-codes them; max_matches 500).

  value : device-timed throughput, inputs resident in HBM (rpe_submit_batch_device)
  e2e   : same metric through the host-buffer C ABI call (rpe_submit_batch): pinned host
          images -> H2D -> kernels -> D2H of R|t|euler|status, all inside the timed region
  --impl reference : the reference's own CPU path (the same cv2 calls PoseEstimator.estimate
          makes; oracle/reference_path.py) on the host cores, bounded sample per step.

Line items of the same JSON line (not the headline; `--lean` skips them all, `--items a,b` picks):
  stream      BASELINE config 5 at its named size: 100 000 consecutive frame pairs of a synthetic video stream,
              sharded over the ranks, chained rpe_submit_sequence calls (ORB once per frame), plus the 1 000-pair
              subsample against the reference's cv2 path
  cfg4        BASELINE config 4 (3840x2160, 20 000 features, 8 levels) with its own stages / cpu_baseline (N = 1)
  hard_ransac config-3 matches with 50 % / 70 % of the correspondences replaced by gross outliers through
              rpe_estimate_from_matches: RANSAC runs hundreds of iterations / to the 2 000 cap (N = 1)
  cfg2        BASELINE config 2 (the reference's simulator sequence, 58 pairs) when its frames are staged (N = 1)

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

CFGS = {
    3: dict(W=1280, H=720, nfeatures=5000, nlevels=12, match_mode="ratio", ratio=0.75, max_matches=500,
            ransac_max_iters=2000, synth_cfg=3, batch=256),
    4: dict(W=3840, H=2160, nfeatures=20000, nlevels=8, match_mode="ratio", ratio=0.75, max_matches=500,
            ransac_max_iters=2000, synth_cfg=4, batch=16),
}
ITEMS = ("stream", "cfg4", "hard_ransac", "cfg2")


def metric_name(cfg):
    return (f"pose pairs/sec (ORB {cfg['nfeatures'] // 1000}k -> BF-Hamming ratio 0.75 -> 5-pt RANSAC 2000 -> recoverPose), "
            f"synthetic {cfg['W']}x{cfg['H']}")


def workload_name(cfg):
    return (f"cfg{cfg['synth_cfg']}: synthetic {cfg['W']}x{cfg['H']} pair batch (two-plane scene, generator v2), ORB {cfg['nfeatures']} features (sf 1.1, "
            f"{cfg['nlevels']} levels, FAST 15, HARRIS), BF-Hamming knn2 + ratio 0.75, max_matches {cfg['max_matches']}, "
            f"RANSAC 5-pt {cfg['ransac_max_iters']} iters")


def workload_config(cfg, B):
    """What names the workload — identical in both arms (the run shape of an arm goes under `run`)."""
    return {"workload": workload_name(cfg), "width": cfg["W"], "height": cfg["H"], "nfeatures": cfg["nfeatures"],
            "nlevels": cfg["nlevels"], "match_mode": cfg["match_mode"], "ratio": cfg["ratio"],
            "max_matches": cfg["max_matches"], "ransac_max_iters": cfg["ransac_max_iters"],
            "pairs_per_gpu_per_step": B, "pair_indices": "rank * 100000 + batch * pairs_per_gpu_per_step + i"}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--batch", type=int, default=None, help="pairs per GPU per step (default 256 for config 3 as SURVEY 8d suggests; config 4: 16)")
    ap.add_argument("--config", type=int, default=3, choices=[3, 4, 5],
                    help="BASELINE.json config: 3 = 1280x720 / 5000 features (the headline workload, default); "
                         "4 = 3840x2160 / 20000 features / 8 levels as the main line; 5 = only the stream line item")
    ap.add_argument("--batches", type=int, default=4, help="distinct batches cycled through the timed loop")
    ap.add_argument("--pairs", type=int, default=100000, help="config 5: consecutive frame pairs of the stream, all ranks together")
    ap.add_argument("--subsample", type=int, default=1000, help="config 5: pairs compared with the reference's cv2 path (0 = skip)")
    ap.add_argument("--cpu-pairs", type=int, default=24, help="pairs in the bounded CPU sample (cv2-threads mode)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--lean", action="store_true", help="headline only: no line items")
    ap.add_argument("--items", default=None, help="comma-separated subset of " + ",".join(ITEMS))
    ap.add_argument("--inflight", type=int, default=1, help="engine handles used alternately")
    ap.add_argument("--depth", type=int, default=2, help="batches pending per handle (rpe_submit_* / rpe_wait_batch; max 2)")
    ap.add_argument("--chunk", type=int, default=32, help="pairs per pipeline chunk (rpe_set_pipeline)")
    ap.add_argument("--streams", type=int, default=1, help="pipeline streams per handle (rpe_set_pipeline); 1 = no chunking")
    a = ap.parse_args()
    a.item_set = set() if a.lean else set(a.items.split(",")) if a.items else set(ITEMS)
    return a


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


# ------------------------------------------------------------------------------------------ models
def level_pixels(cfg):
    """Pyramid pixel counts for the algorithmic-byte model."""
    W, H, L = cfg["W"], cfg["H"], cfg["nlevels"]
    sf = float(np.float32(1.1))
    px = []
    for l in range(L):
        s = np.float32(sf ** l)
        px.append(int(np.rint(np.float32(W) / s)) * int(np.rint(np.float32(H) / s)))
    return px


def stage_models(cfg, B: int, nkp_mean: float):
    """Algorithmic bytes per STEP (B pairs = 2B images) for each stage."""
    px = level_pixels(cfg)
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


def rotation_angle_deg(Ra, Rb):
    return float(np.degrees(np.arccos(np.clip((np.trace(np.asarray(Ra).T @ np.asarray(Rb)) - 1.0) / 2.0, -1.0, 1.0))))


def direction_angle_deg(ta, tb):
    a, b = np.asarray(ta, float).ravel(), np.asarray(tb, float).ravel()
    return float(np.degrees(np.arccos(np.clip(a @ b / (np.linalg.norm(a) * np.linalg.norm(b)), -1.0, 1.0))))


def h2d_ceiling(world):
    """Upload-only ceiling of this rank count on the 8-GPU box (tools/h2d_ceiling.py, committed under profiles/)."""
    try:
        for ln in open(os.path.join(ROOT, "profiles", "r02_h2d_ceiling.jsonl")):
            d = json.loads(ln)
            if d["ranks"] == world and not d["write_combined"]:
                return d
    except Exception:
        pass
    return None


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


def pipe_rates():
    """POPC / REDUX / ... issue rates measured by tools/pipe_rates.cu on a B200 (committed under profiles/)."""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "r02_pipe_rates.json")))
    except Exception:
        return {}


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from bench_support import cpu_arm
    cfg = CFGS[4 if args.config == 4 else 3]
    B = args.batch or cfg["batch"]
    t0 = time.perf_counter()
    n = max(2, args.cpu_pairs // 4)
    for _ in range(min(args.warmup, 2)):
        cpu_arm.rate_cv2_threads(cfg, 2)
    t_all, threads = 0.0, 0
    for s in range(args.steps):
        _, dt, threads = cpu_arm.rate_cv2_threads(cfg, n, first_pair=(s * n) % max(B - n, 1))
        t_all += dt
    value = args.steps * n / t_all
    sample = f"{n} pairs of the workload per step x {args.steps} steps, cv2 with its own {threads} threads"
    modes = {"cv2_threads": value}
    per_proc = 8 if cfg["synth_cfg"] == 3 else 1
    try:
        v2, procs, wall2 = cpu_arm.rate_process_per_core(cfg, args.steps, per_proc)
        modes["process_per_core"] = v2
        if v2 > value:      # report the CPU's better mode
            value, t_all, n = v2, wall2, per_proc * procs
            sample = (f"{procs} single-threaded processes x {per_proc} pairs (pairs 0..{per_proc * procs - 1} of rank 0's "
                      f"first batch) per step x {args.steps} steps; cv2 with its own {threads} threads: "
                      f"{modes['cv2_threads']:.1f} pairs/s")
            threads = procs
    except Exception as e:      # the thread mode above always exists
        modes["process_per_core_error"] = repr(e)[:200]
    out = {"impl": "reference", "metric": metric_name(cfg), "value": value, "unit": "pairs/s", "n_gpus": args.gpus,
           "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_all / args.steps,
           "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
           "dtype": "u8/i32 (ORB, match), f32 (Harris/angle/blur), f64 (RANSAC, pose)", "data": "synthetic",
           "config": workload_config(cfg, B),
           "run": {"pairs_per_step": n, "note": "bounded sample of the workload's pairs on the host cores"},
           "cpu_baseline": {"value": value, "unit": "pairs/s", "cores": threads, "kind": "port", "sample": sample,
                            "modes": modes},
           "e2e": {"value": value, "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "gpu_launches": 0, "wall_s": time.perf_counter() - t0}
    print(json.dumps(out), flush=True)


# ------------------------------------------------------------------------------------------ B200 arm
class Ctx:
    """Per-rank state shared by the headline measurement and the line items."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        import rpe_b200
        self.torch, self.dist, self.rpe = torch, dist, rpe_b200
        self.args = args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a B200: the CUDA library has no CPU fallback")
        torch.cuda.set_device(self.local)
        self.numa_cpus = []
        if self.world > 1 and os.environ.get("RPE_NUMA_BIND", "1") != "0":
            # one process per GPU: run on, and allocate the pinned frame buffers from, the GPU's own NUMA node
            try:
                self.numa_cpus = rpe_b200.sharding.bind_host_to_gpu(rpe_b200.sharding.gpu_pci_bus_id(self.local))
            except Exception:
                self.numa_cpus = []
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        from bench_support.hostgen import host_cores
        self.procs = max(1, min(32, host_cores() // self.world))     # host processes this rank may use for input production
        self.stream = torch.cuda.Stream()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world > 1:
            tt = self.torch.tensor([ms], device="cuda")
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
            ms = float(tt.item())
        return ms

    def timed(self, fn):
        """fn() enqueues and drains its work; CUDA events on the engines' stream, barrier + sync on both sides,
        max over ranks.  Returns (ms, fn's return value)."""
        torch = self.torch
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        ret = fn()
        torch.cuda.synchronize()
        e1.record(self.stream)
        self.barrier()
        return self.max_over_ranks(e0.elapsed_time(e1)), ret

    def engine(self, cfg, max_pairs, **kw):
        e = self.rpe.Engine(self.rpe.EngineConfig(max_width=cfg["W"], max_height=cfg["H"], max_pairs=max_pairs,
                                                  nfeatures=cfg["nfeatures"], nlevels=cfg["nlevels"],
                                                  match_mode=cfg["match_mode"], ratio=cfg["ratio"],
                                                  max_matches=cfg["max_matches"],
                                                  ransac_max_iters=cfg["ransac_max_iters"], **kw), device=self.local)
        e.set_stream(self.stream.cuda_stream)
        return e

    def pinned(self, shape):
        return self.torch.empty(shape, dtype=self.torch.uint8).pin_memory()


def pipelined(engines, depth, n, submit):
    """n submissions over the handles, at most len(engines) * depth batches in flight (rpe_submit_* /
    rpe_wait_batch); returns (kernel launches, results in submission order)."""
    launches, outs, q = 0, [], []
    cap = len(engines) * max(1, min(2, depth))
    for i in range(n):
        e = engines[i % len(engines)]
        if len(q) == cap:
            d = q.pop(0)
            outs.append(d.wait())
            launches += sum(d.stage_times()[1].values())
        submit(e, i)
        q.append(e)
    for d in q:
        outs.append(d.wait())
        launches += sum(d.stage_times()[1].values())
    return launches, outs


def measure_pairs(ctx: Ctx, cfg, B, n_batches, steps, warmup, with_cpu, name):
    """The headline measurement on one pair-batch workload: device-resident, end-to-end, per-stage."""
    from bench_support import hostgen
    from bench_support.synth import synth_K
    from rpe_b200 import _lib as L
    args = ctx.args
    W, H = cfg["W"], cfg["H"]
    K = synth_K(W, H)
    host = ctx.pinned((n_batches, 2 * B, H, W))
    t0 = time.perf_counter()
    # batch b = pairs rank * 100000 + b * B + i: one run of consecutive pair indices, produced by one pool of processes
    hostgen.generate("pairs", host.numpy().reshape(n_batches * 2 * B, H, W), ctx.procs, cfg=cfg["synth_cfg"],
                     first=ctx.rank * 100000)
    gen_s = time.perf_counter() - t0
    dev = host.cuda(non_blocking=False)
    host_np = [host[b].numpy() for b in range(n_batches)]
    engines = [ctx.engine(cfg, B) for _ in range(max(1, args.inflight))]
    for e in engines:
        e.set_pipeline(args.chunk, args.streams)
    eng = engines[0]
    sub_dev = lambda e, i: e.submit_device(dev[i % n_batches].data_ptr(), H * W, B, H, W, W, K)
    sub_host = lambda e, i: e.submit_interleaved(host_np[i % n_batches], K)

    _, outs = pipelined(engines, args.depth, max(warmup, n_batches, len(engines)), sub_dev)
    nkp = float(np.mean([eng.keypoints(i, v)["x"].shape[0] for i in range(min(B, 4)) for v in (0, 1)]))
    ok_frac = float(np.mean([(o["status"] & 0xFF == 0).mean() for o in outs[-n_batches:]]))
    iters = np.concatenate([o["info"][:, 3] for o in outs[-n_batches:]])

    sampler = ClockSampler(ctx.local)
    if ctx.rank == 0:
        sampler.start()
    ms_dev, (launches, _) = ctx.timed(lambda: pipelined(engines, args.depth, steps, sub_dev))     # frames resident in HBM
    pipelined(engines, args.depth, max(2, len(engines)), sub_host)
    ms_e2e, _ = ctx.timed(lambda: pipelined(engines, args.depth, steps, sub_host))                # pinned host frames -> results on host

    def profiled():
        """per-stage CUDA events: one handle, one chunk, stages serial (rpe_set_profiling)"""
        eng.set_profiling(True)
        acc = {k: 0.0 for k in L.RPE_T_NAMES}
        for i in range(steps):
            eng.estimate_batch_device(dev[i % n_batches].data_ptr(), H * W, B, H, W, W, K)
            ms, _ = eng.stage_times()
            for k in acc:
                acc[k] += ms[k]
        eng.set_profiling(False)
        return acc
    ms_prof, stage_ms = ctx.timed(profiled)
    clocks = sampler.stop() if ctx.rank == 0 else None
    for e in engines:
        e.close()
    del dev

    total_pairs = ctx.world * steps * B
    peaks = measured_peaks()
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    model = stage_models(cfg, B, nkp)
    stages = {}
    for k, byts in model.items():
        ms = stage_ms[k] / steps
        gbs = byts / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        stages[k] = {"ms_per_step": ms, "algorithmic_bytes_per_step": int(byts), "gbs": gbs, "frac_hbm": gbs / hbm_peak}
    for k in ("select", "match_sort", "ransac", "pose", "h2d", "d2h"):
        stages[k] = {"ms_per_step": stage_ms[k] / steps}
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
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj[rf_stage]["dram_bytes_per_image"] * 2 * B if cfg["synth_cfg"] == 3 else None
    except Exception:
        pass
    rates = pipe_rates()
    popc_lanes = rates.get("popc", 16.0)
    sm_hz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    popc_peak = 148 * popc_lanes * sm_hz * 1e6
    # the kernel issues 4.25 POPC per 256-bit distance (carry-save adders fold the 8 xor words into 4 words, into 5 on
    # one train descriptor in four), not 8: the pipe fraction counts ISSUED popcounts
    POPC_PER_DISTANCE = 4.25
    popc_issued = POPC_PER_DISTANCE / 8.0 * (stages["match"].get("popc32_per_s") or 0.0)
    xu = {"kernel": "match", "distances_per_s": (stages["match"].get("popc32_per_s") or 0.0) / 8.0,
          "popc_issued_per_distance": POPC_PER_DISTANCE, "achieved_popc32_per_s": popc_issued,
          "plain_form_equivalent_popc32_per_s": stages["match"].get("popc32_per_s"),
          "peak_popc32_per_s": popc_peak, "frac": popc_issued / popc_peak,
          "peak_source": ((f"tools/pipe_rates.cu on a B200 (profiles/r02_pipe_rates.json): {popc_lanes:.2f} POPC / clk / SM"
                           if rates else "16 POPC / clk / SM (unmeasured)") + " x 148 SMs x median SM clock of this run")}
    roofline = {"kernel": rf_stage, "bound": "hbm", "achieved": stages[rf_stage]["gbs"], "peak": hbm_peak,
                "unit": "GB/s", "frac": stages[rf_stage]["frac_hbm"], "traffic": traffic, "peak_source": peak_src,
                "xu_pipe": xu, "dominant_stage": dom, "share_of_step": stages[dom]["ms_per_step"] / (ms_prof / steps),
                "serial_ms_per_step": ms_prof / steps,
                "note": "achieved = algorithmic bytes per step (SURVEY 8d) / CUDA-event time of that stage, measured "
                        "in a separate pass of the same K steps with an event between stages (stages serial, one "
                        "chunk); `value` is the pipelined pass. The matcher is XU-pipe (popc) bound, see "
                        "stages.match.popc32_per_s"}
    cpu = None
    if with_cpu and ctx.world == 1 and ctx.rank == 0 and not args.no_cpu_baseline:
        from bench_support import cpu_arm
        cpu = cpu_arm.cpu_baseline(cfg, name, args.cpu_pairs if cfg["synth_cfg"] == 3 else 2, 8 if cfg["synth_cfg"] == 3 else 1)
    e2e = {"value": total_pairs / (ms_e2e * 1e-3), "unit": "pairs/s", "h2d_bytes_per_step": int(2 * B * H * W + 72),
           "d2h_bytes_per_step": int(B * (15 * 8 + 7 * 4)), "ms_per_step": ms_e2e / steps,
           "h2d_gbs_per_rank": 2 * B * H * W / (ms_e2e / steps * 1e-3) / 1e9}
    ceil = h2d_ceiling(ctx.world)
    if ceil is not None:
        bound = ceil["gbs_per_rank_min"] * 1e9 / (2 * H * W) * ctx.world        # pairs/s the slowest rank's upload allows
        e2e["h2d_ceiling"] = {"gbs_per_rank_min": ceil["gbs_per_rank_min"], "gbs_aggregate": ceil["gbs_aggregate"],
                              "upload_bound_pairs_per_s": bound, "e2e_frac_of_upload_bound": e2e["value"] / bound,
                              "source": "tools/h2d_ceiling.py on the 8xB200 box, copies only, no kernels "
                                        "(profiles/r02_h2d_ceiling.jsonl)"}
    return {"value": total_pairs / (ms_dev * 1e-3), "ms_per_step": ms_dev / steps,
            "e2e": e2e,
            "gpu_launches": int(launches), "roofline": roofline, "cpu_baseline": cpu, "clocks": clocks, "stages": stages,
            "run": {"distinct_batches_cycled": n_batches, "image_bytes_per_step": int(2 * B * H * W),
                    "l2_policy": f"inputs larger than L2 ({2 * B * H * W / 1e6:.0f} MB of frames + "
                                 f"{2 * B * sum(level_pixels(cfg)) / 1e9:.1f} GB of pyramids per step, {n_batches} distinct batches in turn)",
                    "parallelism": f"{ctx.world} independent shards, no collective",
                    "host_numa_binding": f"{len(ctx.numa_cpus)} cpus of the GPU's node" if ctx.numa_cpus else "none",
                    "pipeline": {"handles": len(engines), "batches_pending_per_handle": max(1, min(2, args.depth)),
                                 "chunk_pairs": args.chunk, "streams": args.streams},
                    "mean_keypoints_per_image": nkp, "ok_fraction": ok_frac,
                    "ransac_iterations": {"min": int(iters.min()), "median": float(np.median(iters)), "max": int(iters.max()),
                                          "pairs_with_more_than_one_round": int((iters > 16).sum()), "pairs": int(len(iters))},
                    "input_generation_s": gen_s, "host_processes": ctx.procs}}


# ------------------------------------------------------------------------------------------ line item: config 5
def stream_calls(per_rank, ring_pairs, C):
    """Schedule of chained sequence calls over a ring of ring_pairs + 1 frames: (first frame in the ring, pairs).
    A call starts at the previous call's last frame; at the end of the ring the walk restarts at frame 0 (the pair
    last frame -> frame 0 is not a pair of the stream and is never formed)."""
    calls, done = [], 0
    while done < per_rank:
        pos = done % ring_pairs
        n = min(C, per_rank - done, ring_pairs - pos)
        calls.append((pos, n))
        done += n
    return calls


def run_stream(ctx: Ctx, cfg):
    """BASELINE config 5: `--pairs` consecutive frame pairs of one synthetic 1280x720 video stream (cfg3 parameters),
    sharded over the ranks in contiguous blocks (boundary frame duplicated, no exchange), every rank feeding chained
    rpe_submit_sequence calls (the next call starts at the previous call's last frame; ORB once per frame; two calls
    pending).  The frames come from a ring of distinct pre-generated frames in pinned host memory (a pool of host
    processes fills it before the timed region; the synthetic generator itself is slower than the GPUs, see
    `producer`), re-walked lap after lap until the rank's share of the pairs is done."""
    from bench_support import hostgen
    from bench_support.synth import synth_K
    args, torch = ctx.args, ctx.torch
    W, H = cfg["W"], cfg["H"]
    K = synth_K(W, H)
    per_rank = -(-args.pairs // ctx.world)
    C = min(cfg["batch"], per_rank)                        # pairs per call
    ring_pairs = max(C, min(4096, per_rank) // C * C)      # distinct pairs in the ring (whole calls)
    first = ctx.rank * per_rank                            # global index of this rank's first frame
    ring = ctx.pinned((ring_pairs + 1, H, W))
    prod_fps = hostgen.generate("stream", ring.numpy(), ctx.procs, cfg=5, first=first, stream_idx=0)
    ring_np = ring.numpy()
    dring = ring.cuda(non_blocking=False)
    eng = ctx.engine(cfg, C)
    eng.set_pipeline(args.chunk, args.streams)
    calls = stream_calls(per_rank, ring_pairs, C)
    sub_dev = lambda e, i: e.submit_sequence_device(dring[calls[i][0]].data_ptr(), H * W, calls[i][1] + 1, H, W, W, K)
    sub_host = lambda e, i: e.submit_sequence(ring_np[calls[i][0]:calls[i][0] + calls[i][1] + 1], K)
    lap = min(len(calls), ring_pairs // C)
    _, first_lap = pipelined([eng], args.depth, lap, sub_dev)                 # warm-up = one lap; its results feed the check
    ms_dev, (launches, outs) = ctx.timed(lambda: pipelined([eng], args.depth, len(calls), sub_dev))
    pipelined([eng], args.depth, 2, sub_host)
    ms_e2e, _ = ctx.timed(lambda: pipelined([eng], args.depth, len(calls), sub_host))
    eng.close()
    ok = float(np.mean(np.concatenate([o["status"] & 0xFF == 0 for o in outs])))
    iters = np.concatenate([o["info"][:, 3] for o in outs])
    # repeated laps re-compute the same pairs: the engine is deterministic, so every lap must reproduce the first
    R_first = np.concatenate([o["R"] for o in first_lap])
    R_again = np.concatenate([o["R"] for o in outs[:lap]])
    deterministic = bool(np.array_equal(R_first, R_again))
    total = per_rank * ctx.world
    item = {"workload": "cfg5: synthetic 1280x720 video stream (two-plane scene, generator v2), consecutive frame pairs (stride 1), cfg3 parameters, "
                        "frame-level feature cache (ORB once per frame), chained sequence calls",
            "pairs_total": total, "pairs_per_rank": per_rank, "ranks": ctx.world,
            "value": total / (ms_dev * 1e-3), "e2e": total / (ms_e2e * 1e-3), "unit": "pairs/s",
            "seconds": ms_dev * 1e-3, "e2e_seconds": ms_e2e * 1e-3,
            "pairs_per_call": C, "calls_per_rank": len(calls), "distinct_frames_per_rank": ring_pairs + 1,
            "h2d_bytes_per_pair_e2e": float(sum(n + 1 for _, n in calls) * H * W / per_rank),
            "ok_fraction": ok, "laps_reproduce_first_lap_bit_for_bit": deterministic,
            "ransac_iterations": {"median": float(np.median(iters)), "max": int(iters.max()),
                                  "pairs_with_more_than_one_round": int((iters > 16).sum())},
            "gpu_launches": int(launches),
            "producer": {"frames_per_s": prod_fps * ctx.world, "host_processes": ctx.procs * ctx.world,
                         "consumed_frames_per_s_e2e": total / (ms_e2e * 1e-3),
                         "gpus_would_starve_on_live_generation": bool(prod_fps * ctx.world < total / (ms_e2e * 1e-3)),
                         "note": "frames/s of the synthetic generator (cv2.warpPerspective per frame) over all ranks' "
                                 "host processes while it filled the rings; the timed region reads the pinned rings"}}
    # subsample against the reference's cv2 call sequence (the checker; untimed)
    n_sub = args.subsample // ctx.world
    if n_sub > 0:
        pos = np.unique(np.linspace(0, ring_pairs - 1, n_sub).astype(int))
        path_kw = dict(nfeatures=cfg["nfeatures"], nlevels=cfg["nlevels"], match_mode=cfg["match_mode"], ratio=cfg["ratio"],
                       max_matches=cfg["max_matches"], ransac_max_iters=cfg["ransac_max_iters"])
        t0 = time.perf_counter()
        ref = hostgen.reference_poses(5, 0, [first + int(p) for p in pos], W, H, path_kw, ctx.procs)
        t_first = np.concatenate([o["t"] for o in first_lap])
        st_first = np.concatenate([o["status"] for o in first_lap])
        good, rots, tdirs, failed = 0, [], [], []
        for p in pos:
            Rr, tr = ref[first + int(p)]
            gpu_failed = bool(st_first[p] & 0xFF)
            if Rr is None or gpu_failed:
                good += int(Rr is None and gpu_failed)          # both report a failure: agreement
                continue
            rot, td = rotation_angle_deg(R_first[p], Rr), direction_angle_deg(t_first[p], tr)
            rots.append(rot); tdirs.append(td)
            if rot < 0.1 and td < 0.5:
                good += 1
            else:
                failed.append([int(first + p), round(rot, 4), round(td, 4)])
        counts = np.array([len(pos), good], dtype=np.float64)
        if ctx.world > 1:
            tt = torch.tensor(counts, device="cuda")
            ctx.dist.all_reduce(tt)
            counts = tt.cpu().numpy()
        item["subsample_vs_cv2"] = {"pairs": int(counts[0]), "within_0.1deg_R_and_0.5deg_t": int(counts[1]),
                                    "fraction": float(counts[1] / max(counts[0], 1)),
                                    "rank0_outside_tolerance": failed[:20],
                                    "rank0_median_rotation_deg": float(np.median(rots)) if rots else 0.0,
                                    "cpu_seconds_wall": time.perf_counter() - t0,
                                    "checker": "oracle/reference_path.py (the reference's cv2 call sequence, canonical "
                                               "feature order) on evenly spaced pairs of every rank's ring"}
    return item


# ------------------------------------------------------------------------------------------ line item: hard RANSAC
def run_hard_ransac(ctx: Ctx, cfg):
    """Config-3 matches (the engine's own 500 matched points of 256 synthetic pairs) with a fraction of the
    correspondences replaced by gross outliers, through rpe_estimate_from_matches (findEssentialMat + recoverPose for
    a batch of point sets): the clean synthetic pairs stop RANSAC after one round, these run hundreds of iterations
    (50 %) or to the 2 000 cap (70 %)."""
    from bench_support import hostgen
    from bench_support.synth import synth_K
    from rpe_b200 import _lib as L
    args = ctx.args
    W, H, B = cfg["W"], cfg["H"], cfg["batch"]
    K = synth_K(W, H)
    frames = np.empty((2 * B, H, W), np.uint8)
    hostgen.generate("pairs", frames, ctx.procs, cfg=cfg["synth_cfg"], first=ctx.rank * 100000)
    eng = ctx.engine(cfg, B)
    out = eng.estimate_batch_interleaved(frames, K)
    M = cfg["max_matches"]
    P1 = np.zeros((B, M, 2), np.float32); P2 = np.zeros((B, M, 2), np.float32)
    counts = out["info"][:, 0].astype(np.int32)
    for i in range(B):
        P1[i, :counts[i]] = eng.get_stage(i, 0, L.STAGE_PTS1, 0, np.float32).reshape(-1, 2)
        P2[i, :counts[i]] = eng.get_stage(i, 0, L.STAGE_PTS2, 0, np.float32).reshape(-1, 2)
    res = {"workload": "cfg3 matched points (256 pairs x up to 500), a fraction replaced by uniform random points; "
                       "rpe_estimate_from_matches = findEssentialMat(RANSAC 0.999, 1 px, 2000) + recoverPose"}
    steps = max(3, min(args.steps, 10))
    for frac in (0.0, 0.5, 0.7):
        rng = np.random.default_rng(int(frac * 100))
        Q2 = P2.copy()
        for i in range(B):
            idx = rng.permutation(counts[i])[:int(round(frac * counts[i]))]
            Q2[i, idx] = np.stack([rng.uniform(0, W, len(idx)), rng.uniform(0, H, len(idx))], axis=1).astype(np.float32)
        for _ in range(2):
            o = eng.estimate_from_matches(P1, Q2, counts, K)
        ms, _ = ctx.timed(lambda: [eng.estimate_from_matches(P1, Q2, counts, K) for _ in range(steps)])
        eng.set_profiling(True)
        eng.estimate_from_matches(P1, Q2, counts, K)
        st, _ = eng.stage_times()
        eng.set_profiling(False)
        it = o["info"][:, 3]
        res[f"outliers_{int(frac * 100)}"] = {
            "ms_per_step": ms / steps, "pairs_per_s": B * steps / (ms * 1e-3),
            "stages": {"ransac_ms": st["ransac"], "pose_ms": st["pose"]},
            "ok_fraction": float((o["status"] & 0xFF == 0).mean()),
            "ransac_iterations": {"min": int(it.min()), "median": float(np.median(it)), "max": int(it.max()),
                                  "histogram_edges": [0, 16, 64, 256, 1024, 2000],
                                  "histogram": np.histogram(it, bins=[0, 17, 65, 257, 1025, 2001])[0].tolist()},
            "ransac_inliers_median": float(np.median(o["info"][:, 2]))}
    eng.close()
    return res


# ------------------------------------------------------------------------------------------ line item: config 2
def run_cfg2(ctx: Ctx):
    """BASELINE config 2: the reference's simulator sequence at step 15 (59 frames 640x480 -> 58 pairs, reference
    parameters: 4000 features, cross-check, 500 matches, RANSAC 1000) in one sequence call; frames staged by
    tests/golden/stage_cfg2.py under baseline/_ref (reference DATA, git-ignored)."""
    import cv2
    d = os.path.join(ROOT, "baseline", "_ref", "simulator-data", "images")
    if not os.path.isdir(d):
        return {"unavailable": "config-2 frames not staged (baseline/_ref/simulator-data)"}
    idx = sorted(int(n[:6]) for n in os.listdir(d) if n.endswith(".png"))
    frames = np.stack([cv2.cvtColor(cv2.imread(os.path.join(d, f"{f:06d}.png")), cv2.COLOR_BGR2GRAY) for f in idx])
    sx, sy = 640 / 960.0, 480 / 720.0
    K = np.array([[924.829 * sx, 0, 468.249 * sx], [0, 920.477 * sy, 353.659 * sy], [0, 0, 1.0]])
    pin = ctx.pinned(frames.shape)
    pin.numpy()[...] = frames
    pin_np = pin.numpy()
    cfg = dict(W=640, H=480, nfeatures=4000, nlevels=12, match_mode="crosscheck", ratio=0.75, max_matches=500,
               ransac_max_iters=1000)
    eng = ctx.engine(cfg, len(idx) - 1)
    eng.set_pipeline(ctx.args.chunk, 1)
    for _ in range(3):
        out = eng.estimate_sequence(pin_np, K)
    n = 20
    ms_seq, _ = ctx.timed(lambda: [eng.estimate_sequence(pin_np, K) for _ in range(n)])
    ms_pipe, _ = ctx.timed(lambda: pipelined([eng], 2, n, lambda e, i: e.submit_sequence(pin_np, K)))
    eng.set_profiling(True)
    eng.estimate_sequence(pin_np, K)
    st, _ = eng.stage_times()
    eng.set_profiling(False)
    eng.close()
    it = out["info"][:, 3]
    return {"workload": "cfg2: the reference's simulator sequence, 59 frames 640x480 -> 58 pairs, ORB 4000, cross-check, "
                        "500 matches, RANSAC 1000; pinned host frames in, poses out",
            "ms_per_sequence_blocking": ms_seq / n, "ms_per_sequence_two_pending": ms_pipe / n,
            "pairs_per_s": 58 * n / (ms_pipe * 1e-3), "ok": int((out["status"] == 0).sum()),
            "stages_ms": {k: round(v, 4) for k, v in st.items()},
            "ransac_iterations": {"min": int(it.min()), "median": float(np.median(it)), "max": int(it.max())}}


# ------------------------------------------------------------------------------------------ main
def run_b200(args):
    ctx = Ctx(args)
    main_cfg = CFGS[4 if args.config == 4 else 3]
    B = args.batch or main_cfg["batch"]
    line = {}
    if args.config != 5:
        m = measure_pairs(ctx, main_cfg, B, max(1, args.batches), args.steps, args.warmup, True, f"cfg{main_cfg['synth_cfg']}")
        line = {"metric": metric_name(main_cfg), "value": m["value"], "unit": "pairs/s", "n_gpus": ctx.world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": m["ms_per_step"], "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None,
                "dtype": "u8/i32 (ORB, match), f32 (Harris/angle/blur), f64 (RANSAC, pose)", "data": "synthetic",
                "config": workload_config(main_cfg, B), "run": m["run"], "e2e": m["e2e"], "gpu_launches": m["gpu_launches"],
                "roofline": m["roofline"], "cpu_baseline": m["cpu_baseline"], "clocks": m["clocks"], "stages": m["stages"]}
    items = args.item_set if args.config != 5 else {"stream"}
    if "stream" in items:
        line["stream"] = run_stream(ctx, CFGS[3])
    if ctx.world == 1:
        if "cfg4" in items and args.config == 3:
            c4 = CFGS[4]
            m4 = measure_pairs(ctx, c4, c4["batch"], 2, max(3, min(args.steps, 10)), 3, True, "cfg4")
            line["cfg4"] = {"workload": workload_name(c4), "config": workload_config(c4, c4["batch"]), **m4}
        if "hard_ransac" in items:
            line["hard_ransac"] = run_hard_ransac(ctx, CFGS[3])
        if "cfg2" in items:
            line["cfg2"] = run_cfg2(ctx)
    if args.config == 5:
        s = line["stream"]
        line = {"metric": "pose pairs/sec on a 100k-pair synthetic 1280x720 video stream (config 5)", "value": s["value"],
                "unit": "pairs/s", "n_gpus": ctx.world, "steps": s["calls_per_rank"], "warmup": args.warmup,
                "ms_per_step": 1e3 * s["seconds"] / s["calls_per_rank"], "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "u8/i32, f32, f64", "data": "synthetic",
                "config": {"workload": s["workload"], "pairs_total": s["pairs_total"]},
                "e2e": {"value": s["e2e"], "unit": "pairs/s", "h2d_bytes_per_step": int(s["h2d_bytes_per_pair_e2e"] * s["pairs_per_call"]),
                        "d2h_bytes_per_step": int(s["pairs_per_call"] * (15 * 8 + 7 * 4))},
                "gpu_launches": s["gpu_launches"], "stream": s}
    if ctx.rank == 0:
        print(json.dumps(line), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
