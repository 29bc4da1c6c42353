"""File-parallel sharding of a batch over ranks (BASELINE config 5): no data-path collective, the shards
never talk to each other; the only collectives are the sums / maxima of per-rank scalars for reporting.

One process per GPU (torchrun): rank r decodes the files `assign(...)[r]` with zflac_b200.open_batch on its own device.
"""
from __future__ import annotations

from typing import List, Sequence


def round_robin(n_items: int, world: int) -> List[List[int]]:
    """Item i goes to rank i % world (files of one synthetic batch are all about the same size)."""
    return [list(range(r, n_items, world)) for r in range(world)]


def size_balanced(sizes: Sequence[int], world: int) -> List[List[int]]:
    """Longest-processing-time assignment by compressed size: biggest file first, always to the lightest rank.
    Deterministic (ties: lower index, lower rank), so every rank computes the same plan without communicating."""
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        out[r].append(i)
        load[r] += int(sizes[i])
    for shard in out:
        shard.sort()
    return out


def job_totals(local: Sequence[float], device=None):
    """Sum of per-rank scalars over the job (samples, bytes, files ...). Works without an initialised process group."""
    import torch
    import torch.distributed as dist

    t = torch.tensor([float(x) for x in local], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(x) for x in t.tolist()]


def job_max(value: float, device=None) -> float:
    """Max over ranks (device-timed step durations are reported as the slowest rank's)."""
    import torch
    import torch.distributed as dist

    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())
