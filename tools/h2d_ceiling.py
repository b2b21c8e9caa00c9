"""What bounds the end-to-end rate at 8 GPUs?  (measure, do not assert.)

Every rank copies the bench's per-step upload (472 MB of frames = 512 x 1280 x 720) from pinned host memory to its
GPU, back to back, with NO kernels running: GB/s per rank and aggregate, for N = 1, 2, 4, 8 ranks on one box
(`--write-combined` allocates the host buffer with cudaHostAllocWriteCombined).  If the aggregate flattens below
N x the single-rank figure, the host side (DRAM / PCIe switch uplinks) is the ceiling of `e2e`, not the engine.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29517 tools/h2d_ceiling.py
"""
import argparse
import ctypes
import json
import os

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--mb", type=int, default=472)
    ap.add_argument("--reps", type=int, default=40)
    ap.add_argument("--write-combined", action="store_true")
    ap.add_argument("--pairs-per-step", type=int, default=256)
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    n = a.mb * 1000 * 1000
    dev = torch.empty(n, dtype=torch.uint8, device="cuda")
    rt = ctypes.CDLL("libcudart.so.12")
    if a.write_combined:
        p = ctypes.c_void_p()
        assert rt.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), ctypes.c_uint(0x04)) == 0      # cudaHostAllocWriteCombined
        ctypes.memset(p, 7, n)
        src_ptr = p.value
    else:
        host = torch.full((n,), 7, dtype=torch.uint8).pin_memory()
        src_ptr = host.data_ptr()
    s = torch.cuda.Stream()
    rt.cudaMemcpyAsync.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int, ctypes.c_void_p]

    def copy():
        assert rt.cudaMemcpyAsync(dev.data_ptr(), src_ptr, n, 1, s.cuda_stream) == 0

    for _ in range(3):
        copy()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(s)
    for _ in range(a.reps):
        copy()
    e1.record(s)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    gbs = a.reps * n / (ms * 1e-3) / 1e9
    t = torch.tensor([gbs, gbs, ms], device="cuda", dtype=torch.float64)
    lo, hi = t.clone(), t.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dist.all_reduce(lo, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if rank == 0:
        slowest_ms = float(hi[2])
        print(json.dumps({"ranks": world, "mb_per_copy": a.mb, "reps": a.reps, "write_combined": a.write_combined,
                          "gbs_per_rank_min": float(lo[0]), "gbs_per_rank_max": float(hi[0]), "gbs_aggregate": float(t[0]),
                          "upload_bound_pairs_per_s": world * a.reps * a.pairs_per_step / (slowest_ms * 1e-3)}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
