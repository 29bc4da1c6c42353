"""bench.py --workload c5: BASELINE config 5 — a batch of short synthetic FLAC files sharded file-parallel over the
ranks (one process per GPU, no collective on the data path; only the timing barrier / max).

A "step" decodes the WHOLE batch once: every rank runs zflacb200_batch_open over its shard (one sync scan, one
hinted-chain + K2 + K3 launch over all frames of all files, one MD5 chain per file side by side — the MD5 gate is
INSIDE the timed region here, because a batch has enough independent chains). Total work is fixed as N grows:
"scaling": "strong". `--impl reference`: the CPU oracle decodes a bounded sample of the same files on all host cores.
"""
from __future__ import annotations

import json
import os
import threading
import time

import numpy as np


def make_files(args, rank, world, device=None):
    """This rank's files, built in HBM by tools/flacgen_cuda.cu (device given) or by the CPU generator."""
    from tools import flacgen as fg
    ids = list(range(rank, args.files, world))
    ps = [fg.c5_file(i) for i in ids]
    threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    files, secs = fg.generate_on(device, ps, want_pcm=True, n_threads=threads)
    return ids, files, secs


def oracle_files(flacs, n_threads):
    """Decode a list of files with the CPU oracle on n_threads threads (ctypes releases the GIL). Returns seconds."""
    from oracle import oracle
    oracle.lib()
    nxt = [0]
    lock = threading.Lock()

    def work():
        while True:
            with lock:
                i = nxt[0]
                nxt[0] += 1
            if i >= len(flacs):
                return
            r = oracle.decode(flacs[i], copy_samples=False)
            assert r.error == "OK", r.error

    th = [threading.Thread(target=work) for _ in range(n_threads)]
    t = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    return time.perf_counter() - t


def run_reference(args, rank, world, metric):
    if rank != 0:
        return
    from tools import flacgen as fg
    n_sample = min(args.files, 1600)
    files = fg.generate_batch([fg.c5_file(i) for i in range(n_sample)], want_pcm=False, n_threads=os.cpu_count() or 1)
    flacs = [f.flac for f in files]
    n_samples = sum(int(f.params.n_samples) * int(f.params.channels) for f in files)
    cores = os.cpu_count() or 1
    for _ in range(max(args.warmup, 1)):
        oracle_files(flacs[: 4 * cores], cores)
    ts = [oracle_files(flacs, cores) for _ in range(args.steps)]
    ms = 1e3 * float(np.mean(ts))
    value = n_samples / (ms * 1e-3) / 1e6
    sample = f"first {n_sample} of the {args.files} files ({n_samples} samples) per step, file-parallel on {cores} host threads"
    line = {"impl": "reference", "metric": metric, "value": round(value, 3), "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "int32/int64", "data": "synthetic",
            "config": {"workload": workload_name(args), "sample": sample, "md5": "included (one chain per file)"},
            "cpu_baseline": {"value": round(value, 3), "unit": "Msamples/s", "cores": cores, "kind": "port", "sample": sample,
                             "note": "C++ restatement of zflac (oracle/); the Zig reference cannot be compiled on this image"},
            "e2e": {"value": round(value, 3), "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def workload_name(args):
    return (f"batch of {args.files} short synthetic FLAC files (16-bit stereo 44.1 kHz, 1.0 s, 11 frames of 4096, mixed "
            f"FIXED/LPC/CONSTANT subframes, all channel assignments), file-parallel (BASELINE configs[4])")


def run(args, rank, world, local_rank, metric, ClockSampler, measured_peak):
    if args.impl == "reference":
        run_reference(args, rank, world, metric)
        return
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: zflac_b200 has no CPU fallback (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    import zflac_b200 as z
    from zflac_b200 import _native
    lib = _native.load()
    if lib.zflacb200_init(local_rank) != 0:
        raise SystemExit(f"zflacb200_init failed: {lib.zflacb200_last_error().decode()}")

    t0 = time.time()
    ids, files, gen_secs = make_files(args, rank, world, None if getattr(args, "generator", "cuda") == "cpu" else local_rank)
    t_gen = time.time() - t0
    blob, offs, lens = z.pack_streams([f.flac for f in files])
    n_samples = sum(int(f.pcm.size) for f in files)       # this rank's shard
    pcm_bytes = sum(int(f.pcm.nbytes) for f in files)
    flac_bytes = int(lens.sum())
    h_blob = torch.from_numpy(blob).pin_memory()
    d_blob = h_blob.to(dev)
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    b = z.open_batch(h_blob.numpy(), offs, lens, device_blob_ptr=d_blob.data_ptr(), device=local_rank, stream=sptr)
    out_bytes = max(b.pcm_bytes(), z.batch_pcm_bound(h_blob.numpy(), offs, lens))
    h_pcm = torch.empty(out_bytes, dtype=torch.uint8).pin_memory()
    # parity of what is timed: every file bit-exact vs the generator's PCM, MD5 verified on the device
    buf, po = b.read_all(h_pcm.numpy())
    ok = b.statuses == [0] * len(files)
    for f, v in zip(files, b.streams_from(buf, po)):
        ok = ok and v is not None and np.array_equal(v, f.pcm)
    n_fast = sum(i["fast_path"] for i in b.infos)
    b.close()
    if not ok:
        raise SystemExit("PARITY FAILURE: batch PCM differs from the generator's")
    # ... and the same check through the one-call API the e2e figure is measured on
    res = z.decode_batch_into(h_blob.numpy(), offs, lens, h_pcm.numpy(), device=local_rank, stream=sptr)
    if res.statuses != [0] * len(files) or not all(np.array_equal(res.samples(i), f.pcm) for i, f in enumerate(files)):
        raise SystemExit("PARITY FAILURE: zflacb200_batch_decode PCM differs from the generator's")

    def barrier():
        if world > 1:
            dist.barrier()

    def step_device():
        bb = z.open_batch(h_blob.numpy(), offs, lens, device_blob_ptr=d_blob.data_ptr(), device=local_rank, stream=sptr)
        info = bb.infos[0]
        bb.close()
        return info

    def step_e2e():
        if os.environ.get("ZFLACB200_BENCH_TWO_CALL"):  # the r01 shape of the call: open, then read everything back
            bb = z.open_batch(h_blob.numpy(), offs, lens, device=local_rank, stream=sptr)
            bb.read_all(h_pcm.numpy())
            info = bb.infos[0]
            bb.close()
            return info
        res = z.decode_batch_into(h_blob.numpy(), offs, lens, h_pcm.numpy(), device=local_rank, stream=sptr)
        return res.infos[0]

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
        barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        l0 = lib.zflacb200_kernel_launches()
        w0 = time.time()
        e0.record(stream)
        infos = [fn() for _ in range(steps)]
        e1.record(stream)
        torch.cuda.synchronize()
        w1 = time.time()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        launches = (lib.zflacb200_kernel_launches() - l0) // steps
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, infos, launches, (w0, w1)

    sampler = ClockSampler(local_rank)
    time.sleep(0.3)
    ms_dev, infos, launches, win = timed(step_device, args.steps, args.warmup)
    k = {f: float(np.mean([i[f] for i in infos])) for f in ("ms_scan", "ms_link", "ms_decode", "ms_epilogue", "ms_md5", "ms_total")}
    e2e = None
    if not args.no_e2e:
        ms_e2e, _, _, _ = timed(step_e2e, args.steps, args.warmup)
    t_end = time.time()
    sampler.stop()
    clocks = ClockSampler.summarise(sampler.window(win[0], t_end))

    # whole-job totals (strong scaling: the batch is fixed, ranks hold shards)
    tot = torch.tensor([n_samples, flac_bytes, pcm_bytes, len(files), n_fast], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    job_samples, job_flac, job_pcm, job_files, job_fast = [float(x) for x in tot.tolist()]
    if not args.no_e2e:
        e2e = {"value": round(job_samples / (ms_e2e * 1e-3) / 1e6, 2), "unit": "Msamples/s", "h2d_bytes_per_step": int(job_flac),
               "d2h_bytes_per_step": int(job_pcm), "ms_per_step": round(ms_e2e, 3),
               "api": ("zflac_b200.open_batch(pinned host blob) + read_all(pinned host PCM) -> zflacb200_batch_open / _read_all (C ABI)"
                       if os.environ.get("ZFLACB200_BENCH_TWO_CALL") else
                       "zflac_b200.decode_batch_into(pinned host blob, pinned host PCM) -> zflacb200_batch_decode (C ABI), "
                       "sub-batches overlapped over both PCIe directions"),
               "md5": "included"}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = [f.flac for f in files[:400]]
        ns = sum(int(f.pcm.size) for f in files[:400])
        oracle_files(sample[:8], 1)
        dt1 = oracle_files(sample, 1)
        cores = os.cpu_count() or 1
        dtn = oracle_files(sample, cores)
        cpu = {"value": round(ns / dt1 / 1e6, 3), "unit": "Msamples/s", "cores": 1, "kind": "port",
               "sample": f"first 400 files of the batch ({ns} samples), whole decode incl. MD5, 1 thread",
               "value_all_cores": round(ns / dtn / 1e6, 3), "host_cores": cores}
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peak, peak_src = measured_peak()
    dom = max(("ms_decode", "ms_epilogue", "ms_scan", "ms_md5"), key=lambda f: k[f])
    algo_bytes = flac_bytes + pcm_bytes  # per launch = this rank's shard
    achieved = algo_bytes / (k[dom] * 1e-3) / 1e9
    line = {
        "metric": metric, "value": round(job_samples / (ms_dev * 1e-3) / 1e6, 2), "unit": "Msamples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms_dev, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "int32/int64", "data": "synthetic",
        "config": {"workload": workload_name(args), "files": int(job_files), "files_on_fast_path": int(job_fast),
                   "samples_per_step": int(job_samples), "compressed_bytes": int(job_flac), "pcm_bytes": int(job_pcm),
                   "sharding": "files round-robin over ranks, no collective", "md5": "included (one chain per file, side by side)",
                   "l2": "inputs larger than L2 (no flush needed)" if flac_bytes > 126 * 2**20 else "shard smaller than L2: stream re-read from L2 between steps is possible; PCM + scratch written per step exceed L2",
                   "generator_s": round(t_gen, 2), "generator": {k: (round(v, 3) if isinstance(v, float) else v) for k, v in gen_secs.items()}},
        "compressed_gbs": round(job_flac / (ms_dev * 1e-3) / 1e9, 3),
        "kernel_ms": {f: round(v, 4) for f, v in k.items()},
        "parity": "every file bit-exact vs the generator's PCM, MD5 == STREAMINFO on the device (checked on this run)",
        "roofline": {"bound": "hbm", "kernel": {"ms_decode": "k_parse_frames", "ms_epilogue": "k_predict_epilogue", "ms_scan": "k_sync_scan",
                                               "ms_md5": "k_md5"}[dom],
                     "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 5), "traffic": None,
                     "peak_source": peak_src, "algorithmic_bytes": algo_bytes, "kernel_ms": round(k[dom], 4)},
        "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
