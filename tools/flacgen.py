"""ctypes front end of tools/flacgen.cpp — the synthetic FLAC generator (TEST / BENCH INFRASTRUCTURE).

Presets follow BASELINE.json `configs[1..4]` / SURVEY.md §8(d):

    c2(seconds)      16-bit stereo 44.1 kHz, block size 4096, LPC order 8, mid/side on every frame
    c3(seconds)      24-bit 96 kHz 8 channels, block size 4608, LPC order 32, Rice partition order 7 (max valid)
    c4_variable(), c4_32bit(), c4_wasted(), c4_bs65535()     std_uncommon-style edge cases
    c5_file(i)       one short 16-bit stereo file of the 10k-file batch (mixed subframes, all channel assignments)
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "_build", "libflacgen.so")
_SRC = os.path.join(_HERE, "flacgen.cpp")
_HDR = os.path.join(_HERE, "flacgen_common.h")


class Params(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("sample_rate", C.c_uint32), ("channels", C.c_uint32), ("bps", C.c_uint32),
                ("block_size", C.c_uint32), ("lpc_order", C.c_uint32), ("qlp_precision", C.c_uint32),
                ("rice_method", C.c_uint32), ("min_partition_order", C.c_uint32), ("max_partition_order", C.c_uint32),
                ("stereo_mode", C.c_uint32), ("subframe_mix", C.c_uint32), ("wasted_bits", C.c_uint32),
                ("variable_blocksize", C.c_uint32), ("escape_every", C.c_uint32), ("signal", C.c_uint32),
                ("set_total", C.c_uint32), ("n_threads", C.c_uint32), ("n_samples", C.c_uint64), ("seed", C.c_uint64),
                ("amplitude", C.c_double), ("noise", C.c_double)]


class _Result(C.Structure):
    _fields_ = [("flac", C.c_void_p), ("flac_bytes", C.c_uint64), ("pcm", C.c_void_p), ("pcm_bytes", C.c_uint64),
                ("n_frames", C.c_uint64), ("audio_offset", C.c_uint64), ("container_bytes", C.c_uint32),
                ("pad", C.c_uint32), ("md5", C.c_uint8 * 16)]


def build(force: bool = False) -> str:
    stale = (not os.path.exists(_LIB)) or max(os.path.getmtime(_SRC), os.path.getmtime(_HDR)) > os.path.getmtime(_LIB)
    if force or stale:
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        # x86-64-v3, not -march=native: the .so built here also runs on the GPU box's host CPU
        # -ffp-contract=off: the CUDA generator must reproduce every double of this one (flacgen_common.h)
        subprocess.run(["g++", "-O3", "-march=x86-64-v3", "-ffp-contract=off", "-std=c++17", "-shared", "-fPIC", "-pthread", "-Wall",
                        "-o", _LIB, _SRC], check=True, capture_output=True)
    return _LIB


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB)
        _lib.fg_generate.argtypes = [C.POINTER(Params), C.c_int, C.POINTER(_Result)]
        _lib.fg_generate.restype = C.c_int
        _lib.fg_generate_batch.argtypes = [C.POINTER(Params), C.c_uint32, C.c_int, C.POINTER(_Result), C.c_uint32]
        _lib.fg_generate_batch.restype = C.c_int
        _lib.fg_free.argtypes = [C.POINTER(_Result)]
        _lib.fg_free.restype = None
    return _lib


def params(**kw) -> Params:
    d = dict(sample_rate=44100, channels=2, bps=16, block_size=4096, lpc_order=8, qlp_precision=12, rice_method=0,
             min_partition_order=0, max_partition_order=4, stereo_mode=0, subframe_mix=0, wasted_bits=0,
             variable_blocksize=0, escape_every=0, signal=0, set_total=1, n_threads=0, n_samples=44100, seed=1,
             amplitude=0.30, noise=0.0012)
    d.update(kw)
    p = Params(**d)
    p.struct_size = C.sizeof(Params)
    return p


@dataclass
class Synth:
    flac: np.ndarray          # uint8, the whole stream
    pcm: np.ndarray | None    # int8/int16/int32 interleaved, what zflac.decode returns (MSB aligned)
    md5: bytes
    n_frames: int
    audio_offset: int
    container_bytes: int
    params: Params


_DT = {1: np.int8, 2: np.int16, 4: np.int32}


def _take(res: _Result, p: Params, want_pcm: bool) -> Synth:
    flac = np.frombuffer((C.c_uint8 * res.flac_bytes).from_address(res.flac), dtype=np.uint8).copy()
    pcm = None
    if want_pcm and res.pcm:
        n = res.pcm_bytes // res.container_bytes
        pcm = np.frombuffer((C.c_uint8 * res.pcm_bytes).from_address(res.pcm), dtype=_DT[res.container_bytes], count=n).copy()
    out = Synth(flac, pcm, bytes(res.md5), int(res.n_frames), int(res.audio_offset), int(res.container_bytes), p)
    lib().fg_free(C.byref(res))
    return out


def generate(p: Params, want_pcm: bool = True) -> Synth:
    res = _Result()
    rc = lib().fg_generate(C.byref(p), 1 if want_pcm else 0, C.byref(res))
    if rc != 0:
        raise RuntimeError(f"flacgen failed with code {rc}")
    return _take(res, p, want_pcm)


def generate_batch(ps: list[Params], want_pcm: bool = True, n_threads: int = 0) -> list[Synth]:
    n = len(ps)
    arr = (Params * n)(*ps)
    res = (_Result * n)()
    rc = lib().fg_generate_batch(arr, n, 1 if want_pcm else 0, res, n_threads)
    if rc != 0:
        raise RuntimeError(f"flacgen batch failed with code {rc}")
    return [_take(res[i], ps[i], want_pcm) for i in range(n)]


# ---------------------------------------------------------------- presets
def c2(seconds: float = 3600.0, **kw) -> Params:
    """BASELINE config 2: 1-hour 16-bit stereo 44.1 kHz, fixed block size 4096, LPC order 8, mid/side."""
    return params(sample_rate=44100, channels=2, bps=16, block_size=4096, lpc_order=8, qlp_precision=12,
                  rice_method=0, max_partition_order=4, stereo_mode=3, n_samples=int(round(seconds * 44100)),
                  seed=0x5EED0002, **kw)


def c3(seconds: float = 600.0, **kw) -> Params:
    """BASELINE config 3: 24-bit 96 kHz 8 channels, block size 4608, LPC order 32, max Rice partition order
    (7: 4608 >> 7 = 36 >= 32; order 8 would leave a negative first partition)."""
    return params(sample_rate=96000, channels=8, bps=24, block_size=4608, lpc_order=32, qlp_precision=15,
                  rice_method=1, min_partition_order=7, max_partition_order=7, stereo_mode=0,
                  n_samples=int(round(seconds * 96000)), seed=0x5EED0003, amplitude=0.2, noise=0.0004, **kw)


def c4_variable(seconds: float = 30.0, **kw) -> Params:
    return params(variable_blocksize=1, stereo_mode=4, subframe_mix=1, n_samples=int(seconds * 44100), seed=0x5EED0004, **kw)


def c4_32bit(seconds: float = 30.0, **kw) -> Params:
    return params(bps=32, channels=2, sample_rate=48000, lpc_order=12, qlp_precision=15, rice_method=1, max_partition_order=6,
                  escape_every=7, signal=1, n_samples=int(seconds * 48000), seed=0x5EED0005, amplitude=0.25, **kw)


def c4_wasted(seconds: float = 30.0, **kw) -> Params:
    return params(wasted_bits=1, subframe_mix=1, stereo_mode=0, n_samples=int(seconds * 44100), seed=0x5EED0006, **kw)


def c4_bs65535(seconds: float = 30.0, **kw) -> Params:
    return params(block_size=65535, min_partition_order=0, max_partition_order=0, stereo_mode=3,
                  n_samples=int(seconds * 44100), seed=0x5EED0007, **kw)


def c5_file(i: int, seconds: float = 1.0, **kw) -> Params:
    """BASELINE config 5: file i of the batch of short 16-bit stereo files (11 frames of 4096 at 1.0 s)."""
    return params(stereo_mode=4, subframe_mix=1, n_samples=int(seconds * 44100), seed=0x5EED1000 + i, n_threads=1,
                  amplitude=0.25 + 0.01 * (i % 7), noise=0.0008 + 0.0002 * (i % 5), **kw)


# ---------------------------------------------------------------- the CUDA generator (tools/flacgen_cuda.cu)
_CU_LIB = os.path.join(_HERE, "_build", "libflacgen_cuda.so")
_CU_SRC = os.path.join(_HERE, "flacgen_cuda.cu")


def build_cuda(force: bool = False) -> str:
    """nvcc cross-compiles for sm_100a without a GPU. -fmad=false / -ffp-contract=off: see flacgen_common.h."""
    stale = (not os.path.exists(_CU_LIB)) or max(os.path.getmtime(_CU_SRC), os.path.getmtime(_HDR)) > os.path.getmtime(_CU_LIB)
    if force or stale:
        os.makedirs(os.path.dirname(_CU_LIB), exist_ok=True)
        subprocess.run(["nvcc", "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-fmad=false", "-lineinfo",
                        "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared", "-o", _CU_LIB, _CU_SRC], check=True, capture_output=True)
    return _CU_LIB


class _DevStream(C.Structure):
    _fields_ = [("flac_off", C.c_uint64), ("flac_bytes", C.c_uint64), ("pcm_off", C.c_uint64), ("pcm_bytes", C.c_uint64),
                ("n_frames", C.c_uint64), ("container_bytes", C.c_uint32), ("pad", C.c_uint32), ("md5", C.c_uint8 * 16)]


_cu = None


def lib_cuda():
    global _cu
    if _cu is None:
        if not os.path.exists(_CU_LIB):
            build_cuda()
        _cu = C.CDLL(_CU_LIB)
        _cu.fgc_generate.argtypes = [C.POINTER(Params), C.c_uint32, C.c_int, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p)]
        _cu.fgc_generate.restype = C.c_int
        _cu.fgc_info.argtypes = [C.c_void_p, C.POINTER(_DevStream), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        _cu.fgc_info.restype = C.c_int
        _cu.fgc_copy.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_void_p]
        _cu.fgc_copy.restype = C.c_int
        _cu.fgc_trace.argtypes = [C.c_void_p, C.c_void_p]
        _cu.fgc_trace.restype = C.c_int
        _cu.fgc_free.argtypes = [C.c_void_p]
        _cu.fgc_free.restype = None
        _cu.fgc_last_error.restype = C.c_char_p
    return _cu


class DeviceBatch:
    """Streams generated in device memory by tools/flacgen_cuda.cu: one FLAC blob (stream i at streams[i]['flac_off'], a
    multiple of `align`) and one PCM blob (what zflac.decode returns for every stream, at streams[i]['pcm_off']).
    Same bytes as generate() / generate_batch() for the same parameters (tests/test_gpu_flacgen.py)."""

    def __init__(self, ps: list[Params], device: int = 0, align: int = 16, keep_trace: bool = False):
        self.params = list(ps)
        n = len(self.params)
        arr = (Params * n)(*self.params)
        self._h = C.c_void_p()
        rc = lib_cuda().fgc_generate(arr, n, device, align, 1 if keep_trace else 0, C.byref(self._h))
        if rc != 0:
            raise RuntimeError(f"flacgen_cuda failed with code {rc}: {lib_cuda().fgc_last_error().decode()}")
        outs = (_DevStream * n)()
        sizes = (C.c_uint64 * 4)()
        secs = (C.c_double * 6)()
        lib_cuda().fgc_info(self._h, outs, sizes, secs)
        self.flac_bytes, self.pcm_bytes, self.n_frames, self.n_subframes = (int(x) for x in sizes)
        self.seconds = dict(zip(("host_plan", "alloc_upload", "encode_kernels", "write_kernels", "md5", "total"), (float(x) for x in secs)))
        self.streams = [dict(flac_off=int(o.flac_off), flac_bytes=int(o.flac_bytes), pcm_off=int(o.pcm_off), pcm_bytes=int(o.pcm_bytes),
                             n_frames=int(o.n_frames), container_bytes=int(o.container_bytes), md5=bytes(o.md5)) for o in outs]
        self.device = device

    def copy_to(self, flac_ptr: int = 0, flac_on_device: bool = True, pcm_ptr: int = 0, pcm_on_device: bool = True, stream: int = 0):
        """Copies the blobs to caller memory (raw pointers: a torch tensor's data_ptr() or a numpy array's ctypes.data)."""
        rc = lib_cuda().fgc_copy(self._h, flac_ptr or None, int(flac_on_device), pcm_ptr or None, int(pcm_on_device), stream or None)
        if rc != 0:
            raise RuntimeError(f"fgc_copy failed with code {rc}: {lib_cuda().fgc_last_error().decode()}")

    def to_host(self, want_pcm: bool = True):
        flac = np.empty(self.flac_bytes, dtype=np.uint8)
        pcm = np.empty(self.pcm_bytes, dtype=np.uint8) if want_pcm else None
        self.copy_to(flac.ctypes.data, False, pcm.ctypes.data if want_pcm else 0, False)
        return flac, pcm

    def synths(self, want_pcm: bool = True) -> list[Synth]:
        """The streams as host-side Synth objects (what generate() returns)."""
        flac, pcm = self.to_host(want_pcm)
        out = []
        for p, s in zip(self.params, self.streams):
            # views of the two host blobs (PCM offsets are multiples of 16): no per-stream copies
            f = flac[s["flac_off"]: s["flac_off"] + s["flac_bytes"]]
            x = pcm[s["pcm_off"]: s["pcm_off"] + s["pcm_bytes"]].view(_DT[s["container_bytes"]]) if want_pcm else None
            out.append(Synth(f, x, s["md5"], s["n_frames"], 42, s["container_bytes"], p))
        return out

    def trace(self) -> np.ndarray:
        t = np.zeros((self.n_subframes, 8), dtype=np.uint32)
        rc = lib_cuda().fgc_trace(self._h, t.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"fgc_trace failed with code {rc} (keep_trace=True?): {lib_cuda().fgc_last_error().decode()}")
        return t

    def close(self):
        if self._h:
            lib_cuda().fgc_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def generate_device(ps, device: int = 0, align: int = 16, keep_trace: bool = False) -> DeviceBatch:
    return DeviceBatch([ps] if isinstance(ps, Params) else list(ps), device, align, keep_trace)


def generate_on(device, ps, want_pcm: bool = True, n_threads: int = 0):
    """The streams ps (one Params or a list) as host Synth objects, built by the CUDA generator on `device` (an int) or,
    with device=None, by the CPU generator. Returns (list of Synth, seconds dict) — same bytes either way."""
    import time
    single = isinstance(ps, Params)
    lst = [ps] if single else list(ps)
    t0 = time.time()
    if device is None:
        out = [generate(lst[0], want_pcm)] if single else generate_batch(lst, want_pcm, n_threads)
        return out, {"generator": "cpu", "total": time.time() - t0}
    d = DeviceBatch(lst, device)
    try:
        out = d.synths(want_pcm)
        secs = dict(d.seconds)
    finally:
        d.close()
    secs.update(generator="cuda", in_hbm=secs["total"], total=time.time() - t0)  # total includes the copy to the host
    return out, secs


def cpu_trace(p: Params) -> tuple[Synth, np.ndarray]:
    """generate(p) plus the per-subframe debug records of the CPU generator (fg_set_trace)."""
    bs = 16 if p.variable_blocksize else max(1, int(p.block_size))
    n_rec = (int(p.n_samples) // bs + 8) * int(p.channels) * (8 if p.variable_blocksize else 1)
    t = np.zeros((n_rec, 8), dtype=np.uint32)
    lib().fg_set_trace.argtypes = [C.c_void_p, C.c_uint64]
    lib().fg_set_trace(t.ctypes.data, n_rec)
    try:
        s = generate(p)
    finally:
        lib().fg_set_trace(None, 0)
    return s, t[: s.n_frames * int(p.channels)]


if __name__ == "__main__":
    import sys
    import time
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    t = time.time()
    s = generate(c2(secs), want_pcm=False)
    dt = time.time() - t
    n = int(secs * 44100) * 2
    print(f"c2 {secs:.0f}s: {s.flac.size} bytes, {s.n_frames} frames, {s.flac.size / n:.3f} B/sample, {n / dt / 1e6:.1f} Msamples/s gen")
