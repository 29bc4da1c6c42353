"""Strong-scaling legs of bench.py: the two ways `north_star` shards the decode path over the GPUs of
one box, measured in the SAME run as the weak-scaling headline and reported as sub-objects of its JSON line.

  strong_c2r   ONE synthetic stream (BASELINE configs[1]) split by frame range: rank r decodes the frames that start in
               its 1/N of the bytes (zflacb200_decode_range) and writes its PCM slice at its own place of ONE pinned host
               buffer shared by all ranks (POSIX shared memory registered with CUDA in every process;
               ZFLACB200_RANGE_INTO_WHOLE) — the gather of SURVEY 8e with no copy afterwards. Seven int64 per rank are
               all-gathered and stitched inside the timed step. No bulk data crosses ranks.
  strong_c5    the batch of 10 000 short files (BASELINE configs[4]) dealt round-robin over the ranks, MD5 gate on
               (one chain per file), host blob in / host PCM out through zflacb200_batch_decode.

Both report `value` (whole job / slowest rank's device time), `e2e` (host to host, max over ranks), the slowest rank's
kernel times, the limiting kernel by name and `efficiency_vs_n1`: value / (N x the same workload decoded by rank 0 alone
in this very run).
"""
from __future__ import annotations

import mmap
import os
import time

import numpy as np

KERNELS = {"ms_scan": "k_sync_scan", "ms_link": "k_hint_chain", "ms_decode": "k_parse_frames(_narrow)",
           "ms_epilogue": "k_predict_epilogue", "ms_md5": "k_md5_lanes"}


def _timed(fn, steps, warmup, stream, barrier):
    import torch
    for _ in range(warmup):
        fn()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    outs = [fn() for _ in range(steps)]
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    return e0.elapsed_time(e1) / steps, outs


def _max_over_ranks(vals, dev, world):
    import torch
    import torch.distributed as dist
    t = torch.tensor(vals, device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return [float(x) for x in t.tolist()]


class SharedPinned:
    """One host buffer every rank of the box maps and registers with CUDA: the gathered PCM of a sharded decode."""

    def __init__(self, name: str, nbytes: int, rank: int, barrier):
        import torch
        self.path = f"/dev/shm/{name}"
        if rank == 0:
            with open(self.path, "wb") as f:
                f.truncate(nbytes)
        barrier()
        self.f = open(self.path, "r+b")
        self.mm = mmap.mmap(self.f.fileno(), nbytes)
        self.arr = np.frombuffer(self.mm, dtype=np.uint8)
        self.rank = rank
        self.registered = int(torch.cuda.cudart().cudaHostRegister(self.arr.ctypes.data, nbytes, 0))  # 0 = success

    def close(self, barrier):
        import torch
        torch.cuda.cudart().cudaHostUnregister(self.arr.ctypes.data)
        barrier()
        self.arr = None
        try:
            self.mm.close()
        except BufferError:
            pass
        self.f.close()
        if self.rank == 0:
            os.unlink(self.path)


def strong_c2r(rank, world, local_rank, dev, stream, barrier, steps=5, warmup=2, seconds=3600.0):
    """ONE stream over the ranks by frame range, slices gathered into one shared pinned buffer."""
    import torch
    import zflac_b200 as z
    from tools import flacgen as fg
    from zflac_b200 import sharding

    p = fg.c2(seconds)  # the same seed on every rank: the same stream (as N processes that mapped the same file would hold)
    p.n_threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    synth = fg.generate_on(local_rank, p, want_pcm=(rank == 0))[0][0]  # built in HBM by tools/flacgen_cuda.cu
    C, cb = int(p.channels), synth.container_bytes
    n_samples = int(p.n_samples) * C
    pcm_bytes, flac_bytes = n_samples * cb, int(synth.flac.size)
    h_flac = torch.from_numpy(synth.flac).pin_memory()
    shared = SharedPinned(f"zflacb200_gather_{os.environ.get('MASTER_PORT', '0')}_{os.getppid() if world > 1 else os.getpid()}",
                          pcm_bytes, rank, barrier)
    sptr = stream.cuda_stream

    def step_ranks(lo, hi, n_ranks):
        res = z.decode_range(h_flac.numpy(), lo, hi, shared.arr, device=local_rank, stream=sptr, into_whole=True)
        rows = sharding.gather_rows(res.as_row(), device=dev) if n_ranks > 1 else [res.as_row()]
        return res, z.stitch_ranges(rows, int(res.info["audio_offset"]))

    lo, hi = sharding.byte_ranges(flac_bytes, world)[rank]
    res, ok = step_ranks(lo, hi, world)
    if not ok or not res.accepted:
        raise SystemExit(f"PARITY FAILURE (strong_c2r): shards do not stitch (rank {rank}: {res.as_row()})")
    barrier()
    if rank == 0 and not np.array_equal(shared.arr.view(synth.pcm.dtype), synth.pcm):  # the GATHERED buffer, all ranks' slices
        raise SystemExit("PARITY FAILURE (strong_c2r): gathered PCM differs from the generator's")
    ms_e2e, outs = _timed(lambda: step_ranks(lo, hi, world), steps, warmup, stream, barrier)
    keys = ("ms_scan", "ms_link", "ms_decode", "ms_epilogue")
    k = [float(np.mean([r.info[f] for r, _ in outs])) for f in keys]
    vals = _max_over_ranks([ms_e2e, sum(k)] + k, dev, world)
    ms_e2e, ms_kernels, kmax = vals[0], vals[1], dict(zip(keys, vals[2:]))
    # the same stream decoded by rank 0 alone, in this run: the denominator of efficiency_vs_n1
    n1 = None
    if world > 1:
        if rank == 0:
            ms1_e2e, o1 = _timed(lambda: step_ranks(0, flac_bytes, 1), steps, warmup, stream, lambda: None)
            n1 = (ms1_e2e, float(np.mean([sum(r.info[f] for f in keys) for r, _ in o1])))
        barrier()
    shared.close(barrier)
    if rank != 0:
        return None
    value, e2e = n_samples / (ms_kernels * 1e-3) / 1e6, n_samples / (ms_e2e * 1e-3) / 1e6
    out = {"workload": f"ONE synthetic {seconds:.0f} s 16-bit stereo stream (BASELINE configs[1]) by frame range over {world} GPU(s); "
                       "slices written into ONE pinned host buffer shared by the ranks (no copy after the decode)",
           "scaling": "strong", "value": round(value, 2), "unit": "Msamples/s", "ms_per_step": round(ms_kernels, 4),
           "value_is": "whole stream / slowest rank's kernel time (scan + hint + K2a + K3p)",
           "e2e": {"value": round(e2e, 2), "ms_per_step": round(ms_e2e, 3), "h2d_bytes_per_step": flac_bytes, "d2h_bytes_per_step": pcm_bytes,
                   "gather": "cudaMemcpyAsync of every rank's slice to its offset in one cudaHostRegister-ed POSIX shared-memory buffer",
                   "registered": shared.registered == 0},
           "kernel_ms_slowest_rank": {f: round(v, 4) for f, v in kmax.items()},
           "limiting_kernel": KERNELS[max(kmax, key=kmax.get)], "md5": "not run (one serial chain over the whole PCM does not shard)",
           "parity": "gathered buffer bit-exact vs the generator's PCM; shards stitched (checked on this run)"}
    if n1:
        v1, e1 = n_samples / (n1[1] * 1e-3) / 1e6, n_samples / (n1[0] * 1e-3) / 1e6
        out["n1_same_run"] = {"value": round(v1, 2), "e2e": round(e1, 2)}
        out["efficiency_vs_n1"] = round(value / (world * v1), 3)
        out["e2e_efficiency_vs_n1"] = round(e2e / (world * e1), 3)
    else:
        out["efficiency_vs_n1"] = 1.0
    return out


def strong_c5(rank, world, local_rank, dev, stream, barrier, steps=5, warmup=2, files=10000):
    """The batch of short files dealt round-robin over the ranks, MD5 gate on, host blob in / host PCM out."""
    import torch
    import torch.distributed as dist
    import zflac_b200 as z
    from tools import flacgen as fg
    from zflac_b200 import sharding

    threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    ids = sharding.round_robin(files, world)[rank]
    mine = fg.generate_on(local_rank, [fg.c5_file(i) for i in ids], want_pcm=True)[0]
    sptr = stream.cuda_stream

    def prepare(fs):
        blob, offs, lens = z.pack_streams([f.flac for f in fs])
        h_blob = torch.from_numpy(blob).pin_memory()
        h_pcm = torch.empty(z.batch_pcm_bound(h_blob.numpy(), offs, lens), dtype=torch.uint8).pin_memory()
        d_blob = h_blob.to(dev)
        return h_blob, d_blob, offs, lens, h_pcm

    def legs(fs, bar):
        h_blob, d_blob, offs, lens, h_pcm = prepare(fs)

        def step_dev():
            b = z.open_batch(h_blob.numpy(), offs, lens, device_blob_ptr=d_blob.data_ptr(), device=local_rank, stream=sptr)
            info = b.infos[0]
            b.close()
            return info

        def step_e2e():
            return z.decode_batch_into(h_blob.numpy(), offs, lens, h_pcm.numpy(), device=local_rank, stream=sptr)

        res = step_e2e()
        if res.statuses != [0] * len(fs) or not all(f.pcm is None or np.array_equal(res.samples(i), f.pcm) for i, f in enumerate(fs)):
            raise SystemExit("PARITY FAILURE (strong_c5): batch PCM differs from the generator's")
        if not all(i["md5_checked"] for i in res.infos):
            raise SystemExit("PARITY FAILURE (strong_c5): MD5 gate did not run")
        ms_dev, infos = _timed(step_dev, steps, warmup, stream, bar)
        ms_e2e, _ = _timed(step_e2e, steps, warmup, stream, bar)
        keys = ("ms_scan", "ms_link", "ms_decode", "ms_epilogue", "ms_md5")
        return ms_dev, ms_e2e, [float(np.mean([i[f] for i in infos])) for f in keys], keys

    ms_dev, ms_e2e, k, keys = legs(mine, barrier)
    vals = _max_over_ranks([ms_dev, ms_e2e] + k, dev, world)
    ms_dev, ms_e2e, kmax = vals[0], vals[1], dict(zip(keys, vals[2:]))
    tot = torch.tensor([sum(int(f.pcm.size) for f in mine), sum(int(f.flac.size) for f in mine), sum(int(f.pcm.nbytes) for f in mine)],
                       device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    job_samples, job_flac, job_pcm = [float(x) for x in tot.tolist()]
    n1 = None
    if world > 1:  # the whole batch on rank 0 alone, in this run
        if rank == 0:
            rest = fg.generate_on(local_rank, [fg.c5_file(i) for i in range(files) if i % world != 0], want_pcm=False)[0]
            allf = sorted(list(zip(ids, mine)) + list(zip([i for i in range(files) if i % world != 0], rest)), key=lambda t: t[0])
            m1_dev, m1_e2e, _, _ = legs([f for _, f in allf], lambda: None)
            n1 = (m1_dev, m1_e2e)
        barrier()
    if rank != 0:
        return None
    value, e2e = job_samples / (ms_dev * 1e-3) / 1e6, job_samples / (ms_e2e * 1e-3) / 1e6
    out = {"workload": f"batch of {files} short synthetic FLAC files (BASELINE configs[4]) dealt round-robin over {world} GPU(s), MD5 gate on",
           "scaling": "strong", "value": round(value, 2), "unit": "Msamples/s", "ms_per_step": round(ms_dev, 4),
           "value_is": "whole batch / slowest rank's device-resident step (CUDA events)",
           "e2e": {"value": round(e2e, 2), "ms_per_step": round(ms_e2e, 3), "h2d_bytes_per_step": int(job_flac), "d2h_bytes_per_step": int(job_pcm),
                   "api": "zflacb200_batch_decode (pinned host blob in, pinned host PCM out) on every rank's shard"},
           "kernel_ms_slowest_rank": {f: round(v, 4) for f, v in kmax.items()}, "limiting_kernel": KERNELS[max(kmax, key=kmax.get)],
           "md5": "included: one chain per file, all files of a rank side by side (k_md5_lanes)",
           "parity": "every file bit-exact vs the generator's PCM, MD5 == STREAMINFO (checked on this run)"}
    if n1:
        v1, e1 = job_samples / (n1[0] * 1e-3) / 1e6, job_samples / (n1[1] * 1e-3) / 1e6
        out["n1_same_run"] = {"value": round(v1, 2), "e2e": round(e1, 2)}
        out["efficiency_vs_n1"] = round(value / (world * v1), 3)
        out["e2e_efficiency_vs_n1"] = round(e2e / (world * e1), 3)
    else:
        out["efficiency_vs_n1"] = 1.0
    return out
