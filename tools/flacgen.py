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
    stale = (not os.path.exists(_LIB)) or os.path.getmtime(_SRC) > os.path.getmtime(_LIB)
    if force or stale:
        os.makedirs(os.path.dirname(_LIB), exist_ok=True)
        # x86-64-v3, not -march=native: the .so built here also runs on the GPU box's host CPU
        subprocess.run(["g++", "-O3", "-march=x86-64-v3", "-std=c++17", "-shared", "-fPIC", "-pthread", "-Wall",
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


if __name__ == "__main__":
    import sys
    import time
    secs = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    t = time.time()
    s = generate(c2(secs), want_pcm=False)
    dt = time.time() - t
    n = int(secs * 44100) * 2
    print(f"c2 {secs:.0f}s: {s.flac.size} bytes, {s.n_frames} frames, {s.flac.size / n:.3f} B/sample, {n / dt / 1e6:.1f} Msamples/s gen")
