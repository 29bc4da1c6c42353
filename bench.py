#!/usr/bin/env python3
"""bench.py — headline benchmark of the B200 FLAC decode path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload c2|c3|c5|c2r] [--seconds S]

One "step" = one pass of the whole hot path (K1 sync scan + CRC-16 frame link, K2 subframe decode,
K3 decorrelate/shift/interleave; K4 MD5 gate reported separately) over one synthetic stream.
Default workload = BASELINE config 2: 1-hour 16-bit stereo 44.1 kHz, block size 4096, LPC order 8,
mid/side (317.52 M interleaved samples, 38 760 frames), produced in-process by tools/flacgen.

Prints ONE JSON line (rank 0). `value` is device-timed (CUDA events on the launch stream) with the
compressed stream already resident in HBM; `e2e` goes through the public API with pinned HOST
buffers (H2D of the stream and D2H of the PCM inside the timed region). N > 1: one process per GPU,
each rank decodes its own stream (file-parallel shards, no collective on the data path): weak scaling.

`--impl reference` times the CPU oracle (C++ restatement of zflac; the Zig reference cannot be built
here) on the box's host cores on a bounded sample of the same workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "decoded Msamples/s (interleaved PCM samples out; compressed GB/s alongside)"


def shared_config(name: str) -> dict:
    """config of BOTH arms, byte-identical (the driver compares them): what is measured, not how long it took."""
    return {"workload": name,
            "frame_crc16": "the library's per-frame CRC-16 diagnostic is off in value and e2e (the reference and the CPU arm ignore the field); value_with_frame_crc16 has it on",
            "md5": "value excludes the MD5 pass on BOTH arms (one serial chain per stream); `gated` / value_with_md5 include it, as the reference's decode() does"}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c2", choices=["c2", "c3", "c5", "c2r"])
    ap.add_argument("--seconds", type=float, default=None, help="stream length (default: c2 3600 s, c3 600 s)")
    ap.add_argument("--md5-steps", type=int, default=2, help="timed steps of the MD5-gated variant (one serial chain on a host thread, ~1 s per step on c2)")
    ap.add_argument("--md5-device-steps", type=int, default=0, help="timed steps with the chain on the GPU (k_md5: ~5 s per step on c2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--files", type=int, default=10000, help="c5: number of files in the batch (whole job)")
    ap.add_argument("--no-strong", action="store_true", help="skip the strong-scaling legs (strong_c2r, strong_c5) of the default line")
    ap.add_argument("--generator", default="cuda", choices=["cuda", "cpu"],
                    help="who builds the synthetic input: tools/flacgen_cuda.cu in HBM (default) or tools/flacgen.cpp on the host (same bytes)")
    ap.add_argument("--no-pin", action="store_true", help="do not pin the rank's threads to the GPU's NUMA node")
    return ap.parse_args()


def workload_params(args, rank: int):
    from tools import flacgen as fg
    if args.workload == "c2":
        secs = args.seconds or 3600.0
        p = fg.c2(secs)
        name = f"synthetic {secs:.0f} s 16-bit stereo 44.1 kHz, blocksize 4096, LPC order 8, mid/side (BASELINE configs[1])"
    elif args.workload == "c3":
        secs = args.seconds or 600.0
        p = fg.c3(secs)
        name = f"synthetic {secs:.0f} s 24-bit 96 kHz 8-channel, blocksize 4608, LPC order 32, partition order 7 (BASELINE configs[2])"
    else:
        raise SystemExit("c5 is handled by tools/bench_batch.py")
    p.seed += rank  # every rank decodes its own stream (file-parallel shard)
    return p, name


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows = []
        self.proc = None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
                                          "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def window(self, t0, t1):
        return [r for (t, r) in self.rows if t0 - 0.05 <= t <= t1 + 0.15]

    def stop(self):
        if self.proc:
            self.proc.terminate()

    @staticmethod
    def summarise(rows):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            f = [x.strip() for x in r.split(",")]
            try:
                sm.append(float(f[0]))
                mx = max(mx, float(f[1]))
            except (ValueError, IndexError):
                continue
            for n, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}


def pin_to_gpu_numa(index: int):
    """Threads of this rank (and the pinned buffers it allocates afterwards: first touch) onto the CPUs NVML lists as local
    to the GPU. Returns what was done, for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cpus = {64 * w + b for w, word in enumerate(words) for b in range(64) if (word >> b) & 1 and 64 * w + b < n_cpu}
        allowed = os.sched_getaffinity(0)
        cpus &= allowed
        if cpus and cpus != allowed:
            os.sched_setaffinity(0, cpus)
            return {"pinned": True, "cpus": len(cpus), "of": len(allowed)}
        return {"pinned": False, "cpus": len(allowed), "why": "the GPU's CPU set is every CPU this process may use (one NUMA node, or a cpuset)"}
    except Exception as e:  # noqa: BLE001
        return {"pinned": False, "why": f"{type(e).__name__}: {e}"}


def pcie_floor_ms(torch, dev, h_in, n_out_bytes, reps=5):
    """Concurrent H2D of the stream and D2H of the PCM, nothing else: what the wire alone costs at this N (every rank runs
    it at the same time; the caller takes the max over ranks). Returns (best, mean): the best single repetition, and the
    mean of `reps` repetitions run back to back with the ranks free-running — the regime of the end-to-end loop, where
    the ranks' transfers overlap each other differently from step to step (N = 8 on this pool: calls of 38 ms when a
    rank's group of four has the host path to itself, 89-95 ms when both groups move data)."""
    d_in = torch.empty(h_in.numel(), dtype=torch.uint8, device=dev)
    d_out = torch.empty(n_out_bytes, dtype=torch.uint8, device=dev)
    h_out = torch.empty(n_out_bytes, dtype=torch.uint8).pin_memory()
    s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    best, total = None, 0.0
    for r in range(reps + 1):
        torch.cuda.synchronize()
        e0, e1, e2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s1.wait_event(e0)
        s2.wait_event(e0)
        with torch.cuda.stream(s1):
            d_in.copy_(h_in, non_blocking=True)
            e1.record()
        with torch.cuda.stream(s2):
            h_out.copy_(d_out, non_blocking=True)
            e2.record()
        torch.cuda.synchronize()
        ms = max(e0.elapsed_time(e1), e0.elapsed_time(e2))
        if r > 0:  # the first repetition warms the path up
            best = ms if best is None else min(best, ms)
            total += ms
    return best, total / reps


PASS_A_VARIANTS = ("k_parse_frames<StagedSrc>", "k_parse_frames<QuadSrc>", "k_parse_frames_narrow<64>", "k_parse_frames_narrow<32>",
                   "k_parse_frames_narrow<64, WIDE>", "k_parse_frames_narrow<32, WIDE>")


def pass_a_counts():
    """Launches of every pass-A variant since the library was loaded (zflacb200_pass_a_launches)."""
    from zflac_b200 import _native
    lib = _native.load()
    return [int(lib.zflacb200_pass_a_launches(v)) for v in range(len(PASS_A_VARIANTS))]


def kernel_names(pass_a_before=None):
    """Which kernels the ms_decode / ms_epilogue phases are (two-pass decoder unless ZFLACB200_SINGLE_PASS is forced).
    Pass A has variants (int32 scratch, or int16 "narrow" scratch for streams of <= 16 bits): the one that ran since
    `pass_a_before` (a pass_a_counts() snapshot) is named, with its template argument."""
    single = int(os.environ.get("ZFLACB200_EXTRA_FLAGS", "0"), 16) & 0x20
    pass_a = "k_parse_frames"
    if pass_a_before is not None:
        delta = [b - a for a, b in zip(pass_a_before, pass_a_counts())]
        if any(delta):
            pass_a = PASS_A_VARIANTS[max(range(len(delta)), key=lambda v: delta[v])]
    return {"ms_decode": "k_decode_frames" if single else pass_a,
            "ms_epilogue": "k_epilogue" if single else "k_predict_epilogue",
            "ms_scan": "k_sync_scan", "ms_link": "k_hint_chain", "ms_md5": "k_md5"}


def measured_traffic(workload: str, kernel: str):
    """DRAM bytes per launch of `kernel` from the committed ncu --set full capture (profiles/traffic.json), or None."""
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        table = json.load(open(path)).get(workload, {})
        return table.get(kernel, table.get(kernel.split("<")[0]))
    except (OSError, ValueError):
        return None


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except (KeyError, ValueError):
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def time_oracle(flac: np.ndarray, n_threads: int = 1, copies: int = 1, timed: bool = True):
    """Decode `copies` streams with the CPU oracle, one thread per stream. Returns (seconds, OracleResult).
    timed: the -DZO_TIMED build (no overflow bookkeeping, LPC loop instantiated per order like the reference's comptime
    unroll) — the fair stand-in for the reference in ReleaseFast; timed=False is the instrumented build the tests use."""
    from oracle import oracle
    oracle.lib_timed() if timed else oracle.lib()
    if copies == 1:
        t = time.perf_counter()
        r = oracle.decode(flac, copy_samples=False, timed=timed)
        return time.perf_counter() - t, r
    res = [None] * copies

    def work(i):
        res[i] = oracle.decode(flac, copy_samples=False, timed=timed)  # ctypes releases the GIL

    th = [threading.Thread(target=work, args=(i,)) for i in range(copies)]
    t = time.perf_counter()
    for x in th:
        x.start()
    for x in th:
        x.join()
    return time.perf_counter() - t, res[0]


def run_reference(args, rank, world):
    """The reference arm: the CPU implementation of the same path on the host cores."""
    if rank != 0:
        return
    from tools import flacgen as fg
    full, name = workload_params(args, 0)
    per_ch_rate = full.sample_rate
    # bounded sample: the first 600 s (c2) / 60 s (c3) of the same stream, same seed => same frames
    sample_secs = min(full.n_samples / per_ch_rate, 600.0 if args.workload == "c2" else 60.0)
    p = fg.c2(sample_secs) if args.workload == "c2" else fg.c3(sample_secs)
    s = fg.generate(p, want_pcm=False)
    n_samples = int(p.n_samples) * int(p.channels)
    cores = os.cpu_count() or 1
    # the reference is single-threaded per stream; with N shards (weak scaling) it runs N streams on N host threads
    copies = max(1, min(world, cores))
    for _ in range(max(args.warmup, 1)):
        time_oracle(s.flac, copies=copies)
    ts, ts_nomd5 = [], []
    for _ in range(args.steps):
        dt, r = time_oracle(s.flac, copies=copies)
        assert r.error == "OK", r.error
        ts.append(dt)
        ts_nomd5.append(max(dt - r.t_md5, 1e-9))  # the oracle times its MD5 pass separately
    dt_instr, r_instr = time_oracle(s.flac, copies=copies, timed=False)  # once, for the record: the build the tests use
    ms_with = 1e3 * float(np.mean(ts))
    ms = 1e3 * float(np.mean(ts_nomd5))
    value = copies * n_samples / (ms * 1e-3) / 1e6
    value_with = copies * n_samples / (ms_with * 1e-3) / 1e6
    sample = f"first {sample_secs:.0f} s of the stream ({n_samples} samples, {s.flac.size} bytes) per step, {copies} stream(s) on {copies} thread(s)"
    line = {
        "impl": "reference", "metric": METRIC, "value": round(value, 3), "unit": "Msamples/s", "n_gpus": world,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32/int64", "data": "synthetic",
        "config": shared_config(name), "sample": sample,
        "value_with_md5": round(value_with, 3), "ms_per_step_with_md5": round(ms_with, 3),
        "compressed_gbs": round(copies * s.flac.size / (ms * 1e-3) / 1e9, 4),
        "cpu_baseline": {"value": round(value, 3), "unit": "Msamples/s", "cores": copies, "kind": "port", "sample": sample,
                         "note": "C++ restatement of zflac (oracle/, -DZO_TIMED build: overflow bookkeeping compiled out, LPC loop "
                                 "instantiated per order like the reference's comptime unroll); the Zig reference cannot be compiled on this image",
                         "value_with_md5": round(value_with, 3),
                         "instrumented_build_value": round(copies * n_samples / max(dt_instr - r_instr.t_md5, 1e-9) / 1e6, 3),
                         "t_frames_s": round(r.t_frames, 4), "t_md5_s": round(r.t_md5, 4), "t_shift_s": round(r.t_shift, 4)},
        "e2e": {"value": round(value, 3), "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0,
                "value_with_md5": round(value_with, 3)},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "c5":
        from tools import bench_batch
        bench_batch.run(args, rank, world, local_rank, METRIC, ClockSampler, measured_peak)
        return
    if args.workload == "c2r":  # ONE c2 stream over the ranks by frame range (strong scaling)
        from tools import bench_range
        bench_range.run(args, rank, world, local_rank, METRIC, ClockSampler, measured_peak)
        return
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: zflac_b200 has no CPU fallback (use --impl reference for the CPU oracle)")
    torch.cuda.set_device(local_rank)
    numa = {"pinned": False, "why": "--no-pin"} if args.no_pin else pin_to_gpu_numa(local_rank)  # before any pinned allocation
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)

    import zflac_b200 as z
    from zflac_b200 import _native
    from tools import flacgen as fg

    lib = _native.load()
    st = lib.zflacb200_init(local_rank)
    if st != 0:
        raise SystemExit(f"zflacb200_init failed: {lib.zflacb200_last_error().decode()}")

    # ---------------------------------------------------------------- synthetic input
    p, wl_name = workload_params(args, rank)
    p.n_threads = max(1, (os.cpu_count() or 1) // max(1, min(world, 8)))
    t0 = time.time()
    if args.generator == "cuda":
        # the input is BUILT in HBM (tools/flacgen_cuda.cu, same bytes as tools/flacgen.cpp); the stream and
        # the expected PCM stay on the device, host copies are made for the end-to-end legs only
        dgen = fg.generate_device(p, device=local_rank)
        gen_secs = {"generator": "cuda", **{k: round(v, 3) for k, v in dgen.seconds.items()}}
        d_flac = torch.empty(dgen.flac_bytes, dtype=torch.uint8, device=dev)
        d_exp = torch.empty(dgen.pcm_bytes, dtype=torch.uint8, device=dev)
        dgen.copy_to(d_flac.data_ptr(), True, d_exp.data_ptr(), True)
        st0 = dgen.streams[0]
        dgen.close()
        t_gen = time.time() - t0
        h_flac = torch.empty(d_flac.numel(), dtype=torch.uint8).pin_memory()     # e2e input: pinned host
        h_flac.copy_(d_flac)
        synth = fg.Synth(h_flac.numpy(), d_exp.cpu().numpy().view(fg._DT[st0["container_bytes"]]), st0["md5"], st0["n_frames"], 42,
                         st0["container_bytes"], p)
        gen_secs["host_copies_s"] = round(time.time() - t0 - t_gen, 3)
    else:
        synth = fg.generate(p, want_pcm=True)
        t_gen = time.time() - t0
        gen_secs = {"generator": "cpu", "total": round(t_gen, 3)}
        d_flac = torch.from_numpy(synth.flac).to(dev)                       # resident in HBM before the timed region
        d_exp = torch.from_numpy(synth.pcm.view(np.uint8)).to(dev)
        h_flac = torch.from_numpy(synth.flac).pin_memory()                   # e2e input: pinned host
    n_samples = int(synth.pcm.size)
    cb = synth.container_bytes
    pcm_bytes = n_samples * cb
    flac_bytes = int(synth.flac.size)
    audio_bytes = flac_bytes - synth.audio_offset

    h_pcm = torch.empty(pcm_bytes, dtype=torch.uint8).pin_memory()       # e2e output: pinned host
    stream = torch.cuda.current_stream()
    sptr = stream.cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()

    # The per-frame CRC-16 is a diagnostic of this library (info.n_crc16_mismatch): the reference reads the field and
    # ignores it (src/zflac.zig:551) and so does the CPU arm, so the timed steps of both arms leave it out; the same
    # step with the diagnostic on is reported next to it (value_with_frame_crc16).
    def step_device(verify_md5, frame_crc16=False, md5_on_device=False):
        s = z.open_device(d_flac.data_ptr(), flac_bytes, device=local_rank, stream=sptr, verify_md5=verify_md5, frame_crc16=frame_crc16,
                          md5_on_device=md5_on_device)
        info = s.info
        s.close()
        return info

    def step_e2e(verify_md5):
        return z.decode_into(h_flac.numpy(), h_pcm.numpy(), device=local_rank, verify_md5=verify_md5, stream=sptr, frame_crc16=False)

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

    # ---------------------------------------------------------------- device-resident: value
    pa0 = pass_a_counts()
    ms_dev, infos, launches, win_dev = timed(lambda: step_device(False), args.steps, args.warmup)
    names = kernel_names(pa0)
    k = {f: float(np.mean([i[f] for i in infos])) for f in ("ms_scan", "ms_link", "ms_decode", "ms_epilogue", "ms_md5", "ms_total")}
    info0 = infos[-1]

    # parity of what was just timed: PCM bit-exact vs the generator's PCM, MD5 == STREAMINFO MD5
    s = z.open_device(d_flac.data_ptr(), flac_bytes, device=local_rank, stream=sptr, verify_md5=True)
    exp = d_exp
    import ctypes as C
    got = torch.empty(pcm_bytes, dtype=torch.uint8, device=dev)
    C.cdll.LoadLibrary("libcudart.so.12").cudaMemcpy(C.c_void_p(got.data_ptr()), C.c_void_p(s.device_ptr()), C.c_size_t(pcm_bytes), 3)
    parity_ok = bool(torch.equal(got, exp)) and s.info["md5_computed"] == synth.md5 and s.n_samples == n_samples
    ms_md5_kernel = s.info["ms_md5"]
    s.close()
    del exp, got, d_exp
    if not parity_ok:
        raise SystemExit("PARITY FAILURE: device PCM / MD5 differs from the generator's")

    ms_dev_crc, _, _, _ = timed(lambda: step_device(False, True), max(3, args.steps // 2), 1)

    # MD5-gated variant (K1->K4): a single stream's MD5 is one serial chain
    ms_dev_md5 = ms_dev_md5_gpu = None
    if args.md5_steps > 0:
        ms_dev_md5, _, _, _ = timed(lambda: step_device(True), args.md5_steps, 1)
        if args.md5_device_steps > 0:  # the device-side chain (k_md5), for the record: ~5 s per step on C2
            ms_dev_md5_gpu, _, _, _ = timed(lambda: step_device(True, md5_on_device=True), args.md5_device_steps, 0)

    # ---------------------------------------------------------------- end to end through the public API
    e2e = None
    if not args.no_e2e:
        ms_e2e, _, _, win_e2e = timed(lambda: step_e2e(False), args.steps, args.warmup)
        ok = np.array_equal(h_pcm.numpy().view(synth.pcm.dtype), synth.pcm)
        if not ok:
            raise SystemExit("PARITY FAILURE: e2e PCM differs from the generator's")
        e2e = {"value": round(world * n_samples / (ms_e2e * 1e-3) / 1e6, 2), "unit": "Msamples/s",
               "h2d_bytes_per_step": flac_bytes, "d2h_bytes_per_step": pcm_bytes, "ms_per_step": round(ms_e2e, 3),
               "api": "zflac_b200.decode_into(pinned host stream, pinned host PCM) -> zflacb200_decode (C ABI)",
               "md5": "skipped in value; see value_with_md5"}
        if args.md5_steps > 0:
            ms_e2e_md5, _, _, _ = timed(lambda: step_e2e(True), args.md5_steps, 0)
            e2e["value_with_md5"] = round(world * n_samples / (ms_e2e_md5 * 1e-3) / 1e6, 2)
            e2e["ms_per_step_with_md5"] = round(ms_e2e_md5, 3)
        # what the wire alone costs at this N: every rank moves its bytes both ways at the same time, nothing else runs
        barrier()
        floor, floor_mean = pcie_floor_ms(torch, dev, h_flac, pcm_bytes, reps=max(5, args.steps))
        if world > 1:
            tf = torch.tensor([floor, floor_mean], device=dev, dtype=torch.float64)
            dist.all_reduce(tf, op=dist.ReduceOp.MAX)
            floor, floor_mean = (float(x) for x in tf.tolist())
        e2e["pcie_floor_ms"] = round(floor, 3)             # best repetition (max over ranks)
        e2e["pcie_floor_mean_ms"] = round(floor_mean, 3)   # mean of free-running repetitions: the regime of ms_per_step
        e2e["vs_pcie_floor"] = round(ms_e2e / floor, 3)
        e2e["vs_pcie_floor_mean"] = round(ms_e2e / floor_mean, 3)
        e2e["numa"] = numa
    t_end = time.time()
    sampler.stop()
    # samples cover every timed region of this run (device-resident, MD5-gated, end-to-end)
    clocks = ClockSampler.summarise(sampler.window(win_dev[0], t_end))

    # ---------------------------------------------------------------- CPU baseline (rank 0, N = 1)
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        secs = min(p.n_samples / p.sample_rate, 600.0 if args.workload == "c2" else 60.0)
        ps = fg.c2(secs) if args.workload == "c2" else fg.c3(secs)
        ss = fg.generate(ps, want_pcm=False)
        time_oracle(ss.flac)
        dt, r = time_oracle(ss.flac)
        dt_i, r_i = time_oracle(ss.flac, timed=False)
        ns = int(ps.n_samples) * int(ps.channels)
        cpu = {"value": round(ns / dt / 1e6, 3), "unit": "Msamples/s", "cores": 1, "kind": "port",
               "sample": f"first {secs:.0f} s of the stream ({ns} samples), whole decode incl. MD5, 1 thread (the reference is single-threaded)",
               "value_without_md5": round(ns / max(dt - r.t_md5, 1e-9) / 1e6, 3),
               "build": "-DZO_TIMED (no overflow bookkeeping, LPC per order); instrumented_build_value: the build the tests use",
               "instrumented_build_value": round(ns / dt_i / 1e6, 3),
               "host_cores": os.cpu_count()}

    # ---------------------------------------------------------------- strong scaling: the two shardings north_star names
    strong = {}
    if not args.no_strong and args.workload == "c2":
        from tools import bench_strong
        del d_flac, h_flac, h_pcm
        torch.cuda.empty_cache()
        strong["strong_c2r"] = bench_strong.strong_c2r(rank, world, local_rank, dev, stream, barrier, steps=max(3, args.steps // 2), warmup=2,
                                                       seconds=args.seconds or 3600.0)
        strong["strong_c5"] = bench_strong.strong_c5(rank, world, local_rank, dev, stream, barrier, steps=max(3, args.steps // 2), warmup=2,
                                                     files=args.files)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # the drop-in decode() contract: MD5 gate ON (the library's default and the only mode the reference has)
    gated = None
    if ms_dev_md5:
        gated = {"value": round(world * n_samples / (ms_dev_md5 * 1e-3) / 1e6, 2), "ms_per_step": round(ms_dev_md5, 3),
                 "e2e": e2e.get("value_with_md5") if e2e else None,
                 "e2e_ms_per_step": e2e.get("ms_per_step_with_md5") if e2e else None,
                 "vs_cpu": round(e2e["value_with_md5"] / cpu["value"], 2) if e2e and cpu and e2e.get("value_with_md5") else None,
                 "unit": "Msamples/s",
                 "md5": "one stream's MD5 is a serial chain: it runs on ONE host thread over the PCM as it lands in host memory "
                        "(verification only, no sample is produced on the CPU); the device chain (ZFLACB200_MD5_ON_DEVICE) is in md5_on_device",
                 "md5_on_device_ms_per_step": round(ms_dev_md5_gpu, 3) if ms_dev_md5_gpu else None}
    peak, peak_src = measured_peak()
    algo_bytes = audio_bytes + pcm_bytes  # compressed read once + PCM written once
    dom = max(("ms_decode", "ms_epilogue", "ms_scan", "ms_link"), key=lambda f: k[f])
    dom_ms = k[dom]
    achieved = algo_bytes / (dom_ms * 1e-3) / 1e9
    value = world * n_samples / (ms_dev * 1e-3) / 1e6
    line = {
        "metric": METRIC, "value": round(value, 2), "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": round(ms_dev, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int32/int64", "data": "synthetic",
        "config": shared_config(wl_name),
        "workload_detail": {"samples_per_step_per_gpu": n_samples, "frames": int(info0["n_frames"]),
                            "compressed_bytes": flac_bytes, "pcm_bytes": pcm_bytes, "sharding": "one stream per GPU, no collective",
                            "l2": "inputs larger than L2 (no flush needed)" if flac_bytes > 126 * 2**20 else "input smaller than L2",
                            "generator_s": round(t_gen, 2), "generator": gen_secs},
        "gated": gated,
        **strong,
        "compressed_gbs": round(world * flac_bytes / (ms_dev * 1e-3) / 1e9, 3),
        "value_with_frame_crc16": round(world * n_samples / (ms_dev_crc * 1e-3) / 1e6, 2), "ms_per_step_with_frame_crc16": round(ms_dev_crc, 4),
        "value_with_md5": round(world * n_samples / (ms_dev_md5 * 1e-3) / 1e6, 2) if ms_dev_md5 else None,
        "ms_per_step_with_md5": round(ms_dev_md5, 3) if ms_dev_md5 else None,
        "kernel_ms": {**{f: round(v, 4) for f, v in k.items()}, "ms_md5_when_on": round(ms_md5_kernel, 3)},
        "parity": "PCM bit-exact vs generator and MD5 == STREAMINFO (checked on this run)",
        "roofline": {"bound": "hbm", "kernel": names[dom],
                     "achieved": round(achieved, 2), "peak": peak, "unit": "GB/s", "frac": round(achieved / peak, 5),
                     "traffic": measured_traffic(args.workload, names[dom]), "peak_source": peak_src,
                     "algorithmic_bytes": algo_bytes, "kernel_ms": round(dom_ms, 4),
                     "whole_step": {"achieved": round(algo_bytes / (ms_dev * 1e-3) / 1e9, 2),
                                    "frac": round(algo_bytes / (ms_dev * 1e-3) / 1e9 / peak, 5)},
                     "kernels": {f: names[f] for f in ("ms_scan", "ms_link", "ms_decode", "ms_epilogue")},
                     "note": "pass A walks one serial bit chain per frame: bound by instruction issue / latency at ~2 warps per scheduler, not by HBM; see DESIGN.md"},
        "cpu_baseline": cpu,
        "e2e": e2e,
        "gpu_launches": int(launches),
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
