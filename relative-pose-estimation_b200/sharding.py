"""Multi-GPU sharding of independent frame pairs.

Pairs are independent in the reference: ``BatchProcessor`` seeds every pair from ground
truth, never from the previous estimate (reference: src/core/batch_processor.py:82-92), so a
sequence shards into contiguous blocks of pair indices, one block per GPU / process, with a
host-side gather of ``R | t | euler | status`` (16 doubles per pair) and NO data-path
collective: NVLink / NVSwitch stay idle by design.  ``torch.distributed`` (NCCL on the GPU
box, gloo in the CPU tests) is used only for that final gather of results.
"""
from __future__ import annotations

import numpy as np

RESULT_WIDTH = 16   # R (9) | t (3) | euler yaw,pitch,roll (3) | status (1)


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


def bind_host_to_gpu(pci_bus_id: str) -> list[int]:
    """Pin the calling process to the CPUs of the NUMA node its GPU hangs off (``/sys/bus/pci/devices/<id>/
    local_cpulist``), so that the pinned frame buffers it allocates next are first-touched in node-local memory and
    every rank's H2D stream reads its own socket's DRAM.  With one process per GPU and ~20 GB/s of uploads per rank
    this is what keeps 8 ranks from sharing one socket's memory bandwidth.  Returns the CPU list ([] = left alone:
    unknown device, single node, or no permission)."""
    import os
    try:
        path = f"/sys/bus/pci/devices/{pci_bus_id.lower()}/local_cpulist"
        cpus = _parse_cpulist(open(path).read()) & set(os.sched_getaffinity(0))
        if not cpus or cpus == set(os.sched_getaffinity(0)):
            return []
        os.sched_setaffinity(0, cpus)
        return sorted(cpus)
    except (OSError, ValueError, AttributeError):
        return []


def gpu_pci_bus_id(device: int) -> str:
    """'dddd:bb:dd.0' of a CUDA device, as sysfs names it."""
    import torch
    p = torch.cuda.get_device_properties(device)
    return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"


def shard_range(n_pairs: int, world: int, rank: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of pair indices owned by ``rank``: [r*n/G, (r+1)*n/G)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    if n_pairs < 0:
        raise ValueError("n_pairs must be >= 0")
    return (rank * n_pairs) // world, ((rank + 1) * n_pairs) // world


def shard_frames(frame_indices, world: int, rank: int):
    """Frames a rank must load for its block of CONSECUTIVE pairs (pair i = frames i, i+1):
    the boundary frame is duplicated on both neighbours rather than exchanged."""
    lo, hi = shard_range(max(len(frame_indices) - 1, 0), world, rank)
    return list(frame_indices[lo:hi + 1]) if hi > lo else []


def pack_results(out: dict) -> np.ndarray:
    """engine.estimate_batch dict -> (B, 16) float64 rows."""
    B = out["R"].shape[0]
    rows = np.empty((B, RESULT_WIDTH), np.float64)
    rows[:, 0:9] = out["R"].reshape(B, 9)
    rows[:, 9:12] = out["t"].reshape(B, 3)
    rows[:, 12:15] = out["euler"].reshape(B, 3)
    rows[:, 15] = out["status"]
    return rows


def unpack_results(rows: np.ndarray) -> dict:
    B = rows.shape[0]
    return {"R": rows[:, 0:9].reshape(B, 3, 3).copy(), "t": rows[:, 9:12].reshape(B, 3, 1).copy(),
            "euler": rows[:, 12:15].copy(), "status": rows[:, 15].astype(np.int32)}


def gather_results(local_rows: np.ndarray, n_pairs: int, group=None):
    """Gather every rank's (hi-lo, 16) rows on rank 0 in pair order.  Returns the (n_pairs, 16)
    array on rank 0 and None elsewhere.  Works without an initialised process group (world 1)."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        if local_rows.shape[0] != n_pairs:
            raise ValueError("single process must hold every pair")
        return local_rows
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    lo, hi = shard_range(n_pairs, world, rank)
    if local_rows.shape != (hi - lo, RESULT_WIDTH):
        raise ValueError(f"rank {rank}: expected {(hi - lo, RESULT_WIDTH)} rows, got {local_rows.shape}")
    cap = max(shard_range(n_pairs, world, r)[1] - shard_range(n_pairs, world, r)[0] for r in range(world))
    dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else torch.device("cpu")
    buf = torch.zeros((cap, RESULT_WIDTH), dtype=torch.float64, device=dev)
    buf[:hi - lo] = torch.from_numpy(np.ascontiguousarray(local_rows)).to(dev)
    parts = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    if rank != 0:
        return None
    out = np.empty((n_pairs, RESULT_WIDTH), np.float64)
    for r in range(world):
        a, b = shard_range(n_pairs, world, r)
        out[a:b] = parts[r][:b - a].cpu().numpy()
    return out
