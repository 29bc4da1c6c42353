"""bench.py --workload c2r: BASELINE config 2 as ONE stream over N GPUs by frame range.

Every rank holds the same synthetic 1-hour stream in pinned host memory (as N processes that mapped the same file
would), decodes the frames whose header starts in its 1/N of the bytes with zflacb200_decode_range and keeps its slice
of the PCM in its own pinned buffer. The ranks exchange seven scalars each (all-gather of int64) and check that the
shards link up into the chain the reference walks (zflacb200_stitch) — inside the timed region. No bulk data crosses
ranks; total work is fixed as N grows: "scaling": "strong".

The call is host-in / host-out, so the headline of this workload is `e2e` (host -> host, wall time of the slowest
rank, CUDA events on each rank's stream). `value` is the same step with the two PCIe legs taken out: samples of the
whole stream / max over ranks of the shard's kernel time (scan + hint passes + K2a + K3p, CUDA events inside the
library around the kernels only, inputs already uploaded). The MD5 gate is one serial chain over the whole PCM and is
not part of a sharded decode.
"""
from __future__ import annotations

import json
import os
import time

import numpy as np


def run(args, rank, world, local_rank, metric, ClockSampler, measured_peak):
    import torch
    import torch.distributed as dist

    if args.impl == "reference":
        raise SystemExit("--workload c2r has no reference arm of its own: use --workload c2 --impl reference (the CPU decode of the same stream)")
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: zflac_b200 has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    import zflac_b200 as z
    from tools import flacgen as fg
    from zflac_b200 import _native, sharding
    lib = _native.load()
    if lib.zflacb200_init(local_rank) != 0:
        raise SystemExit(f"zflacb200_init failed: {lib.zflacb200_last_error().decode()}")

    secs = args.seconds if args.seconds else 3600.0
    p = fg.c2(secs)
    p.n_threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    t0 = time.time()
    # same seed on every rank: the same stream
    (synth,), gen_secs = fg.generate_on(None if getattr(args, "generator", "cuda") == "cpu" else local_rank, p, want_pcm=True)
    t_gen = time.time() - t0
    C = int(p.channels)
    n_samples = int(synth.pcm.size)
    cb = synth.container_bytes
    flac_bytes = int(synth.flac.size)
    h_flac = torch.from_numpy(synth.flac).pin_memory()
    lo, hi = sharding.byte_ranges(flac_bytes, world)[rank]
    share = (n_samples * cb) // world
    h_pcm = torch.empty(share + share // 8 + (16 << 20), dtype=torch.uint8).pin_memory()
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()

    def step():
        res = z.decode_range(h_flac.numpy(), lo, hi, h_pcm.numpy(), device=local_rank, stream=sptr)
        rows = sharding.gather_rows(res.as_row(), device=dev)
        ok = z.stitch_ranges(rows, int(res.info["audio_offset"]))
        return res, ok

    # parity of what is timed: this rank's slice bit-exact vs the generator's PCM, and the shards stitch
    res, ok = step()
    if not ok or not res.accepted:
        raise SystemExit(f"PARITY FAILURE: shards do not stitch (rank {rank}: {res.as_row()})")
    if res.n_frames and not np.array_equal(res.samples, synth.pcm[res.first_sample * C:(res.first_sample + res.n_samples) * C]):
        raise SystemExit(f"PARITY FAILURE: rank {rank} slice differs from the generator's PCM")

    sampler = ClockSampler(local_rank)
    for _ in range(args.warmup):
        step()
    barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = lib.zflacb200_kernel_launches()
    t_a = time.time()
    e0.record(stream)
    outs = [step() for _ in range(args.steps)]
    e1.record(stream)
    torch.cuda.synchronize()
    barrier()
    t_b = time.time()
    sampler.stop()
    clocks = ClockSampler.summarise(sampler.window(t_a, t_b))
    launches = (lib.zflacb200_kernel_launches() - l0) // args.steps
    ms_e2e = e0.elapsed_time(e1) / args.steps
    keys = ("ms_scan", "ms_link", "ms_decode", "ms_epilogue", "ms_total")
    k = {f: float(np.mean([r.info[f] for r, _ in outs])) for f in keys}
    ms_kernels = k["ms_scan"] + k["ms_link"] + k["ms_decode"] + k["ms_epilogue"]
    t = torch.tensor([ms_e2e, ms_kernels, k["ms_decode"], k["ms_epilogue"]], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_e2e, ms_kernels, ms_k2a, ms_k3p = [float(x) for x in t.tolist()]
    mine = torch.tensor([res.n_samples * C, int(res.info["pcm_bytes"]), hi - lo], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(mine, op=dist.ReduceOp.SUM)
    job_samples, job_pcm, job_flac = [float(x) for x in mine.tolist()]
    if int(job_samples) != n_samples:
        raise SystemExit(f"PARITY FAILURE: shards cover {int(job_samples)} of {n_samples} samples")
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    algo = (flac_bytes + n_samples * cb) / world  # per launch: one rank's shard
    achieved = algo / (ms_k2a * 1e-3) / 1e9
    line = {
        "metric": metric, "value": round(n_samples / (ms_kernels * 1e-3) / 1e6, 2), "unit": "Msamples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_kernels, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "int32/int64", "data": "synthetic",
        "config": {"workload": f"ONE synthetic {secs:.0f} s 16-bit stereo 44.1 kHz stream (BASELINE configs[1]) split by frame range over "
                               f"{world} GPU(s): each rank decodes the frames starting in its 1/{world} of the bytes",
                   "samples_per_step": n_samples, "compressed_bytes": flac_bytes, "pcm_bytes": n_samples * cb,
                   "sharding": "contiguous byte ranges; 7 int64 per rank all-gathered and stitched inside the timed step; no bulk exchange",
                   "md5": "not run (one serial chain over the whole PCM does not shard)",
                   "l2": "inputs larger than L2 (no flush needed)" if flac_bytes // world > 126 * 2**20 else
                         "a shard's compressed bytes fit L2, but they are uploaded again every step; scratch + PCM written per step exceed L2",
                   "value_is": "whole stream / slowest rank's kernel time (upload and read-back excluded); e2e is host to host",
                   "generator_s": round(t_gen, 2), "generator": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in gen_secs.items()}},
        "kernel_ms": {f: round(v, 4) for f, v in k.items()},
        "parity": "every rank's slice bit-exact vs the generator's PCM; shards stitched by zflacb200_stitch (checked on this run)",
        "roofline": {"bound": "hbm", "kernel": "k_parse_frames", "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 5), "traffic": None, "peak_source": peak_src, "algorithmic_bytes": int(algo),
                     "kernel_ms": round(ms_k2a, 4)},
        "cpu_baseline": None,
        "e2e": {"value": round(n_samples / (ms_e2e * 1e-3) / 1e6, 2), "unit": "Msamples/s", "h2d_bytes_per_step": int(job_flac),
                "d2h_bytes_per_step": int(job_pcm), "ms_per_step": round(ms_e2e, 3),
                "api": "zflac_b200.decode_range(pinned host stream, byte range, pinned host PCM) -> zflacb200_decode_range + "
                       "all_gather(7 x int64) + zflacb200_stitch"},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
