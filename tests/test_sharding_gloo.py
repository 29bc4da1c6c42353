"""Host logic of the N > 1 path on CPU: world_size-2 gloo. Every rank plans the same shards without communicating,
the shards partition the batch, per-rank totals sum to the job totals, timings reduce by max — and the files a rank
would hand to its GPU decode bit-exactly in the oracle (the checker), so no file is lost or duplicated by the plan."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT

N_FILES = 10


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import oracle
        from tools import flacgen as fg
        from zflac_b200 import sharding

        ps = []
        for i in range(N_FILES):
            p = fg.c5_file(i)
            p.n_samples = 2048 + 700 * i
            ps.append(p)
        files = [fg.generate(p, want_pcm=True) for p in ps]
        sizes = [int(f.flac.size) for f in files]
        plan = sharding.size_balanced(sizes, world)
        rr = sharding.round_robin(N_FILES, world)
        mine = plan[rank]
        n_samples = 0
        for i in mine:  # what this rank's GPU would decode
            r = oracle.decode(files[i].flac)
            assert r.error == "OK" and np.array_equal(r.samples, files[i].pcm)
            n_samples += int(r.samples.size)
        tot = sharding.job_totals([n_samples, sum(sizes[i] for i in mine), len(mine)])
        slow = sharding.job_max(10.0 + rank)
        q.put((rank, plan, rr, tot, slow, sum(int(f.pcm.size) for f in files), sum(sizes)))
    finally:
        dist.destroy_process_group()


def test_two_rank_sharding_plan_and_totals():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    res.sort()
    (_, plan0, rr0, tot0, slow0, all_samples, all_bytes), (_, plan1, rr1, tot1, slow1, _, _) = res
    assert plan0 == plan1 and rr0 == rr1                      # same plan on every rank, no communication
    for plan in (plan0, rr0):
        assert sorted(i for shard in plan for i in shard) == list(range(N_FILES))  # a partition
    assert tot0 == tot1 == [float(all_samples), float(all_bytes), float(N_FILES)]
    assert slow0 == slow1 == 11.0


def test_size_balanced_is_balanced():
    from zflac_b200 import sharding
    rng = np.random.default_rng(5)
    sizes = rng.integers(10_000, 2_000_000, size=1000).tolist()
    for world in (2, 4, 8):
        plan = sharding.size_balanced(sizes, world)
        loads = [sum(sizes[i] for i in s) for s in plan]
        assert sorted(i for s in plan for i in s) == list(range(1000))
        assert max(loads) - min(loads) <= max(sizes)


# ---------------------------------------------------------------- frame-range sharding of one stream (host logic)

def _rows_from_oracle(data, world):
    """What zflacb200_decode_range would report per rank, computed from the oracle's frame trace (the checker)."""
    from oracle import oracle
    from zflac_b200 import sharding
    r = oracle.decode(data, trace_frames=True)
    assert r.error == "OK"
    C = r.channels
    fr = r.frames
    names = fr.dtype.names
    off, end, s0, bs = (fr[names[0]].astype(np.int64), fr[names[1]].astype(np.int64), fr[names[2]].astype(np.int64),
                        fr[names[3]].astype(np.int64))
    total = int(r.samples.size) // C
    rows = []
    for lo, hi in sharding.byte_ranges(len(data), world):
        m = (off >= max(lo, r.audio_offset)) & (off < hi)
        if not m.any():
            rows.append([1, max(lo, r.audio_offset), max(lo, r.audio_offset), 0, 0, 0, total])
            continue
        rows.append([1, int(off[m][0]), int(end[m][-1]), int(s0[m][0]) // C, int(bs[m].sum()), int(m.sum()), total])
    return rows, r.audio_offset


def test_byte_ranges_partition():
    from zflac_b200 import sharding
    for n, world in ((0, 3), (5, 8), (1000, 1), (284_294_528, 8), (12345, 7)):
        rg = sharding.byte_ranges(n, world)
        assert len(rg) == world and rg[0][0] == 0 and rg[-1][1] == n
        assert all(a[1] == b[0] for a, b in zip(rg, rg[1:])) and all(lo <= hi for lo, hi in rg)


def test_stitch_rule_against_oracle_frames():
    import zflac_b200 as z
    from tools import flacgen as fg
    s = fg.generate(fg.c2(3.0), want_pcm=False)
    for world in (1, 2, 3, 8, 64):
        rows, audio = _rows_from_oracle(s.flac, world)
        assert z.stitch_ranges(rows, audio)
        if world > 1:
            for k, delta in ((1, 1), (2, -1), (3, 4096), (4, 1)):   # a byte gap, an overlap, a sample gap, a lost frame count
                bad = [list(r) for r in rows]
                j = next(i for i in range(1, world) if bad[i][5])
                bad[j][k] += delta
                if k in (1, 3):
                    assert not z.stitch_ranges(bad, audio)
            bad = [list(r) for r in rows]
            bad[0][0] = 0                                            # a shard that was not accepted
            assert not z.stitch_ranges(bad, audio)
        short = [list(r) for r in rows]
        short[-1][6] += 1                                            # stated total not reached
        short[0][6] += 1
        assert not z.stitch_ranges(short, audio)
        assert not z.stitch_ranges(rows, audio + 1)                  # first frame not at the audio offset


def _range_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import zflac_b200 as z
        from tools import flacgen as fg
        from zflac_b200 import sharding
        s = fg.generate(fg.c2(2.0), want_pcm=False)
        rows, audio = _rows_from_oracle(s.flac, world)
        mine = rows[rank]                        # each rank contributes only its own shard's seven scalars
        gathered = sharding.gather_rows(mine)
        q.put((rank, gathered, rows, z.stitch_ranges(gathered, audio)))
    finally:
        dist.destroy_process_group()


def test_two_rank_range_rows_gather_and_stitch():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 31500 + os.getpid() % 2000
    procs = [ctx.Process(target=_range_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, gathered, rows, ok in res:
        assert gathered == rows and ok           # every rank sees all shards' rows in rank order and reaches the same verdict
