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


# ---------------------------------------------------------------- frame-range sharding of ONE stream

def byte_ranges(n_bytes: int, world: int) -> List[tuple]:
    """[lo, hi) of each rank: equal contiguous pieces of the stream's bytes (a frame belongs to the rank its header
    starts in; the library clamps the first range to the audio offset)."""
    cut = [n_bytes * r // world for r in range(world + 1)]
    return [(cut[r], cut[r + 1]) for r in range(world)]


def gather_rows(row: Sequence[int], device=None) -> List[List[int]]:
    """All ranks' rows (small int64 vectors), in rank order. No process group: just this rank's."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1):
        return [[int(x) for x in row]]
    t = torch.tensor([int(x) for x in row], dtype=torch.int64, device=device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    return [[int(x) for x in o.tolist()] for o in out]


def decode_sharded(data, *, rank: int, world: int, device: int = -1, dst=None, gather_device=None):
    """One long stream over `world` GPUs, one process each: this rank decodes the frames that start in its byte range
    (zflac_b200.decode_range), the ranks exchange seven scalars each and every rank checks that the shards link up into
    the chain the reference walks (zflacb200_stitch). Returns (RangeResult of this rank, stitched: bool). When the
    stitch fails every rank knows it (same rows everywhere) and the caller decodes the stream whole on one rank
    (zflac_b200.decode), which also reports the reference's error for a bad stream. No bulk data crosses ranks."""
    import zflac_b200 as z

    n = len(data) if not hasattr(data, "nbytes") else int(data.nbytes)
    lo, hi = byte_ranges(n, world)[rank]
    res = z.decode_range(data, lo, hi, dst, device=device)
    rows = gather_rows(res.as_row(), device=gather_device)
    return res, z.stitch_ranges(rows, int(res.info["audio_offset"]))
