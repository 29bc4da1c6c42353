"""ctypes binding of the CPU oracle (oracle/zflac_oracle.cpp).

TEST INFRASTRUCTURE, NOT PRODUCT CODE. Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this module. The
product package (zflac_b200/) never does.

The oracle restates /root/reference/src/zflac.zig:217-666 (decode,
decode_frames, decode_residuals, ...) and src/bit_reader.zig; see the header of
zflac_oracle.h for how its parity is pinned.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_build", "libzflac_oracle.so")

STATUS_NAMES = {
    0: "OK", 1: "InvalidSignature", 2: "InvalidMetadataHeader", 3: "Unimplemented",
    4: "InvalidChecksum", 5: "MissingStreaminfo", 6: "InvalidFrameHeader",
    7: "InvalidCodedNumber", 8: "InconsistentParameters", 9: "InvalidSubframeHeader",
    10: "InvalidResidualCodingMethod", 11: "EndOfStream", 12: "OutOfMemory",
    90: "UndefinedInReference",
}

FLAG_TRACE_FRAMES = 1
FLAG_KEEP_ON_MD5_MISMATCH = 2


class _Frame(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("end_offset", C.c_uint64), ("sample_offset", C.c_uint64),
                ("block_size", C.c_uint32), ("header_len", C.c_uint8), ("channel_code", C.c_uint8),
                ("bit_depth_code", C.c_uint8), ("sample_rate_code", C.c_uint8)]


class _Result(C.Structure):
    _fields_ = [("status", C.c_int32), ("channels", C.c_uint8), ("bits_per_sample", C.c_uint8),
                ("container_bytes", C.c_uint8), ("overflow_flag", C.c_uint8), ("sample_rate", C.c_uint32),
                ("n_samples", C.c_uint64), ("samples", C.c_void_p), ("md5_computed", C.c_uint8 * 16),
                ("md5_streaminfo", C.c_uint8 * 16), ("n_frames", C.c_uint64), ("frames", C.POINTER(_Frame)),
                ("audio_offset", C.c_uint64), ("t_frames", C.c_double), ("t_md5", C.c_double),
                ("t_shift", C.c_double)]


def build(force: bool = False) -> str:
    """Compile the oracle with the committed Makefile (g++ -O3, no -march=native so the
    .so built in the build container also runs on the GPU box's host CPU)."""
    src = [os.path.join(_HERE, f) for f in ("zflac_oracle.cpp", "zflac_oracle.h", "Makefile")]
    stale = (not os.path.exists(_LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in src)
    if force or stale:
        subprocess.run(["make", "-C", _HERE] + (["-B"] if force else []), check=True, capture_output=True)
    return _LIB_PATH


_lib = None
_lib_timed = None
_LIB_TIMED_PATH = os.path.join(_HERE, "_build", "libzflac_oracle_timed.so")


def lib_timed():
    """The build bench.py TIMES as the CPU arm (-DZO_TIMED: no overflow bookkeeping, LPC loop instantiated per order like
    the reference's comptime unroll). Same entry points and results as lib(); tests/test_oracle_golden.py pins it too."""
    global _lib_timed
    if _lib_timed is None:
        build()
        if not os.path.exists(_LIB_TIMED_PATH) or os.path.getmtime(_LIB_TIMED_PATH) < os.path.getmtime(os.path.join(_HERE, "zflac_oracle.cpp")):
            subprocess.run(["make", "-C", _HERE, "-B"], check=True, capture_output=True)
        _lib_timed = C.CDLL(_LIB_TIMED_PATH)
        _lib_timed.zo_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.POINTER(_Result)]
        _lib_timed.zo_decode.restype = C.c_int
        _lib_timed.zo_free.argtypes = [C.POINTER(_Result)]
        _lib_timed.zo_free.restype = None
    return _lib_timed


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.zo_decode.argtypes = [C.c_void_p, C.c_size_t, C.c_uint32, C.POINTER(_Result)]
        _lib.zo_decode.restype = C.c_int
        _lib.zo_free.argtypes = [C.POINTER(_Result)]
        _lib.zo_free.restype = None
        _lib.zo_crc8.argtypes = [C.c_void_p, C.c_size_t]
        _lib.zo_crc8.restype = C.c_uint8
        _lib.zo_crc16.argtypes = [C.c_void_p, C.c_size_t]
        _lib.zo_crc16.restype = C.c_uint16
        _lib.zo_md5.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p]
        _lib.zo_md5.restype = None
    return _lib


@dataclass
class OracleResult:
    status: int
    error: str  # Zig error name, "OK" on success
    channels: int = 0
    sample_rate: int = 0
    bits_per_sample: int = 0
    container_bytes: int = 0
    samples: np.ndarray | None = None  # int8/int16/int32, interleaved, post MSB-align shift
    overflow_flag: int = 0
    md5_computed: bytes = b""
    md5_streaminfo: bytes = b""
    audio_offset: int = 0
    frames: np.ndarray | None = None  # structured array of the frame trace
    t_frames: float = 0.0
    t_md5: float = 0.0
    t_shift: float = 0.0
    extra: dict = field(default_factory=dict)


_FRAME_DTYPE = np.dtype([("offset", "<u8"), ("end_offset", "<u8"), ("sample_offset", "<u8"),
                         ("block_size", "<u4"), ("header_len", "u1"), ("channel_code", "u1"),
                         ("bit_depth_code", "u1"), ("sample_rate_code", "u1")])


def _as_buffer(data):
    if isinstance(data, np.ndarray):
        arr = np.ascontiguousarray(data, dtype=np.uint8)
    else:
        arr = np.frombuffer(bytes(data) if not isinstance(data, (bytes, bytearray, memoryview)) else data, dtype=np.uint8)
    return arr


def decode(data, trace_frames: bool = False, keep_on_md5_mismatch: bool = False, copy_samples: bool = True,
           timed: bool = False) -> OracleResult:
    """Oracle twin of zflac.decode(allocator, reader) (zflac.zig:217). timed: the -DZO_TIMED build (lib_timed())."""
    arr = _as_buffer(data)
    res = _Result()
    flags = (FLAG_TRACE_FRAMES if trace_frames else 0) | (FLAG_KEEP_ON_MD5_MISMATCH if keep_on_md5_mismatch else 0)
    the_lib = lib_timed() if timed else lib()
    status = the_lib.zo_decode(arr.ctypes.data, arr.size, flags, C.byref(res))
    out = OracleResult(status=status, error=STATUS_NAMES.get(status, f"Unknown({status})"),
                       overflow_flag=res.overflow_flag, md5_computed=bytes(res.md5_computed),
                       md5_streaminfo=bytes(res.md5_streaminfo), audio_offset=res.audio_offset,
                       t_frames=res.t_frames, t_md5=res.t_md5, t_shift=res.t_shift)
    if res.samples:
        out.channels = res.channels
        out.sample_rate = res.sample_rate
        out.bits_per_sample = res.bits_per_sample
        out.container_bytes = res.container_bytes
        dt = {1: np.int8, 2: np.int16, 4: np.int32}[res.container_bytes]
        n = int(res.n_samples)
        if copy_samples:
            buf = (C.c_uint8 * (n * res.container_bytes)).from_address(res.samples) if n else b""
            out.samples = np.frombuffer(buf, dtype=dt).copy() if n else np.zeros(0, dtype=dt)
        else:
            out.extra["n_samples"] = n
        if trace_frames and res.frames:
            nf = int(res.n_frames)
            fb = (C.c_uint8 * (nf * _FRAME_DTYPE.itemsize)).from_address(C.addressof(res.frames.contents)) if nf else b""
            out.frames = np.frombuffer(fb, dtype=_FRAME_DTYPE).copy() if nf else np.zeros(0, dtype=_FRAME_DTYPE)
    the_lib.zo_free(C.byref(res))
    return out


def crc8(data) -> int:
    arr = _as_buffer(data)
    return int(lib().zo_crc8(arr.ctypes.data, arr.size))


def crc16(data) -> int:
    arr = _as_buffer(data)
    return int(lib().zo_crc16(arr.ctypes.data, arr.size))


def md5(data) -> bytes:
    arr = _as_buffer(data)
    out = (C.c_uint8 * 16)()
    lib().zo_md5(arr.ctypes.data, arr.size, out)
    return bytes(out)
