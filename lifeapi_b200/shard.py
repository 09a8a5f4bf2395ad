"""Batch sharding and a torch.distributed hit-list gather -- the PYTHON-side helper.

Since round 2 the product's multi-GPU path is inside the C library (csrc/multi.cu: lifeb200_multi_*, NCCL gather of
device-resident sorted hit lists; capi.MultiEngine) and bench.py uses that.  What stays here is what a Python caller
with its own process group may still want: the shard split, the optional NUMA binding, and a gather of host-side hit
arrays over whatever backend the group has (gloo in the CPU tests).

Patterns are independent and the 64x64 torus is closed, so the stepping path needs no exchange:
rank r of R owns the contiguous id range [r*n/R, (r+1)*n/R).  The only collective is one small
gather per batch: fixed-size counters (all_gather) and the variable-length hit lists (padded
all_gather), over NCCL on the GPU box and over gloo in the CPU tests.  torch.distributed is
plumbing here; no tensor maths happens in this file.
"""
from __future__ import annotations

import numpy as np

from .capi import HIT_DTYPE


def shard_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous id range of `rank` (same split as the CPU baseline's thread chunks)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world {rank}/{world}")
    return n * rank // world, n * (rank + 1) // world


def bind_host_to_device(device_index: int) -> str:
    """Pin the calling process to the CPU cores next to GPU `device_index` (NVML's ideal affinity), so the
    pinned host buffers of lifeb200_host_alloc are first touched on that GPU's NUMA node and the host
    pipeline's copies of 8 ranks do not all cross the socket link.  One process per GPU only: a single
    process that also runs the host-thread CPU baseline must keep every core.  Returns what was done."""
    import os
    if os.environ.get("LIFEB200_NO_NUMA_BIND"):
        return "off (LIFEB200_NO_NUMA_BIND)"
    try:
        import pynvml
        import torch
        pynvml.nvmlInit()
        try:
            pr = torch.cuda.get_device_properties(device_index)
            bus_id = f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
            handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode())
        except Exception:
            handle = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        pynvml.nvmlDeviceSetCpuAffinity(handle)
        cores = sorted(os.sched_getaffinity(0))
        return f"{len(cores)} cores ({cores[0]}..{cores[-1]})"
    except Exception as exc:  # affinity is an optimisation, never a requirement
        return f"unavailable ({type(exc).__name__}: {exc})"


def merge_hits(per_rank: list[np.ndarray]) -> np.ndarray:
    """Concatenate per-rank hit records and sort by id (deterministic order for comparison)."""
    parts = [np.asarray(h, dtype=HIT_DTYPE) for h in per_rank if len(h)]
    if not parts:
        return np.zeros(0, dtype=HIT_DTYPE)
    allh = np.concatenate(parts)
    ids = allh["id"]
    if len(ids) > 1 and not (ids[1:] >= ids[:-1]).all():  # ranks own ascending id ranges: usually already in order
        allh = allh[np.argsort(ids, kind="stable")]
    return allh


def gather_hits(hits: np.ndarray, counters: list[int], device=None, group=None):
    """All ranks call this with their local sorted hits and a list of integer counters.

    Returns (merged_hits, summed_counters, xor_counters) on every rank.  Two collectives:
    all_gather of [n_hits, counters...] (fixed size), then all_gather of the hit records padded to
    the largest count (hit lists are KBs..MBs; NVLink bandwidth is irrelevant, latency is one
    collective per batch).
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device("cpu")
    hits = np.ascontiguousarray(np.asarray(hits, dtype=HIT_DTYPE))
    head = torch.tensor([len(hits)] + [int(c) & 0x7FFFFFFFFFFFFFFF for c in counters] +
                        [int(c) >> 63 for c in counters], dtype=torch.int64, device=dev)
    heads = [torch.zeros_like(head) for _ in range(world)]
    dist.all_gather(heads, head, group=group)
    heads_np = [h.cpu().numpy() for h in heads]
    k = len(counters)
    counts = [int(h[0]) for h in heads_np]
    per_rank_counters = [[int(h[1 + j]) | (int(h[1 + k + j]) << 63) for j in range(k)] for h in heads_np]
    sums = [sum(c[j] for c in per_rank_counters) & 0xFFFFFFFFFFFFFFFF for j in range(k)]
    xors = [0] * k
    for c in per_rank_counters:
        for j in range(k):
            xors[j] ^= c[j]
    maxc = max(counts) if counts else 0
    if maxc == 0:
        return np.zeros(0, dtype=HIT_DTYPE), sums, xors
    pad = np.zeros(maxc, dtype=HIT_DTYPE)
    pad[: len(hits)] = hits
    send = torch.from_numpy(pad.view(np.int64)).to(dev)  # 2 x int64 per record
    recv = torch.empty(world * send.numel(), dtype=torch.int64, device=dev)
    try:  # one collective into one buffer, one copy back
        dist.all_gather_into_tensor(recv, send, group=group)
    except (RuntimeError, NotImplementedError):  # a backend without the flat form
        parts_t = [torch.empty_like(send) for _ in range(world)]
        dist.all_gather(parts_t, send, group=group)
        recv = torch.cat(parts_t)
    host = recv.cpu().numpy().view(HIT_DTYPE).reshape(world, maxc)
    parts = [host[r, :c] for r, c in enumerate(counts)]
    return merge_hits(parts), sums, xors
