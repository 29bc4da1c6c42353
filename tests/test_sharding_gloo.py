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
