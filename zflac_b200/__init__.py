"""zflac_b200 — B200-native FLAC decode path behind Senryoku/zflac's public API.

Host-side mirror of the reference interface (the reference is Zig; there is no Zig toolchain
in this image, so the host side above the C ABI is Python, see INTEGRATION.md for the Zig shim):

    reference                                   here
    ---------                                   ----
    zflac.decode(allocator, reader)             zflac_b200.decode(data)            (src/zflac.zig:217)
    DecodedFLAC{channels, sample_rate,          DecodedFLAC(channels, sample_rate, (src/zflac.zig:18-24)
      bits_per_sample, samples: s8|s16|s32}       bits_per_sample, samples: ndarray int8|int16|int32)
    error.InvalidSignature, ...                 zflac_b200.InvalidSignature, ...   (one class per Zig error)

All decoding runs in libzflacb200.so (hand-written sm_100a kernels, include/zflacb200.h). If the
library or a CUDA device is missing, decode() raises; nothing falls back to the CPU.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np

from . import _native
from ._native import INPUT_ON_DEVICE, MD5_ON_DEVICE, NO_CRC_ACCEL, NO_FAST_PATH, SKIP_FRAME_CRC16, SKIP_MD5, VERIFY_MD5  # noqa: F401

__all__ = ["decode", "decode_into", "decode_batch", "open_batch", "DeviceBatch", "pack_streams", "DecodedFLAC", "ZflacError",
           "CudaError", "open_device", "DeviceStream", "decode_frames", "metadata"]


class ZflacError(Exception):
    """Base of the Zig error set decode() can return."""
    status = -1


def _mk(name, status):
    cls = type(name, (ZflacError,), {"status": status, "__doc__": f"error.{name}"})
    globals()[name] = cls
    __all__.append(name)
    return cls


_ERRORS = {
    1: _mk("InvalidSignature", 1), 2: _mk("InvalidMetadataHeader", 2), 3: _mk("Unimplemented", 3),
    4: _mk("InvalidChecksum", 4), 5: _mk("MissingStreaminfo", 5), 6: _mk("InvalidFrameHeader", 6),
    7: _mk("InvalidCodedNumber", 7), 8: _mk("InconsistentParameters", 8), 9: _mk("InvalidSubframeHeader", 9),
    10: _mk("InvalidResidualCodingMethod", 10), 11: _mk("EndOfStream", 11), 12: _mk("OutOfMemory", 12),
    90: _mk("UndefinedInReference", 90), 100: _mk("CudaError", 100), 101: _mk("InvalidArgument", 101),
}


def _raise(status: int):
    lib = _native.load()
    detail = (lib.zflacb200_last_error() or b"").decode()
    cls = _ERRORS.get(status, ZflacError)
    raise cls(f"{lib.zflacb200_status_name(status).decode()}" + (f": {detail}" if detail else ""))


@dataclass
class DecodedFLAC:
    """Mirror of `pub const DecodedFLAC` (src/zflac.zig:18-28). `samples` plays the role of the
    s8/s16/s32 union: its dtype is the union tag; values are interleaved, frame-major /
    channel-minor, MSB-aligned for 9-15 and 17-31 bit streams (zflac.zig:287-306)."""
    channels: int
    sample_rate: int
    bits_per_sample: int
    samples: np.ndarray
    info: dict = field(default_factory=dict)

    @property
    def tag(self) -> str:
        return {1: "s8", 2: "s16", 4: "s32"}[self.samples.dtype.itemsize]

    def deinit(self):  # the reference frees _samples_backing here; numpy owns the buffer
        self.samples = None


def _info_dict(info: _native.Info) -> dict:
    d = {k: getattr(info, k) for k, _ in _native.Info._fields_ if k not in ("md5_computed", "md5_streaminfo")}
    d["md5_computed"] = bytes(info.md5_computed)
    d["md5_streaminfo"] = bytes(info.md5_streaminfo)
    return d


def _options(flags=0, device=-1, stream=None) -> _native.Options:
    return _native.Options(C.sizeof(_native.Options), flags, device, stream)


def _as_u8(data) -> np.ndarray:
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8)
    if hasattr(data, "read"):
        data = data.read()  # the reference takes a reader; slurp it
    return np.frombuffer(data, dtype=np.uint8)


_DTYPES = {1: np.int8, 2: np.int16, 4: np.int32}
FORMATS = {"native": 0, "f32": 1, "s24": 2, "u8": 3}  # zflacb200_format


def decode(data, *, device: int = -1, verify_md5: bool = True, use_crc_links: bool = True,
           use_fast_path: bool = True, md5_on_device: bool = False, word_pipeline: bool = False,
           single_pass: bool = False) -> DecodedFLAC:
    """Drop-in for `zflac.decode(allocator, reader)`: `data` is bytes / ndarray / a readable object
    holding the whole stream. Raises the ZflacError subclass named like the Zig error.
    word_pipeline (diagnostics, ZFLACB200_WORD_PIPELINE): pass A of 17..31-bit streams by the first bit reader.
    single_pass (diagnostics, ZFLACB200_SINGLE_PASS): the fused subframe decoder instead of parse + predict."""
    lib = _native.load()
    arr = _as_u8(data)
    flags = (0 if verify_md5 else SKIP_MD5) | (0 if use_crc_links else NO_CRC_ACCEL) | (0 if use_fast_path else NO_FAST_PATH)
    flags |= MD5_ON_DEVICE if md5_on_device else 0
    flags |= _native.WORD_PIPELINE if word_pipeline else 0
    flags |= _native.SINGLE_PASS if single_pass else 0
    opt = _options(flags, device)
    handle = C.c_void_p()
    info = _native.Info()
    st = lib.zflacb200_open(arr.ctypes.data, arr.size, C.byref(opt), C.byref(handle), C.byref(info))
    if st != 0:
        _raise(st)
    try:
        # the caller-side allocation of the reference (allocator.allocWithOptions(u8, n, 32), zflac.zig:331)
        out = np.empty(int(info.n_samples), dtype=_DTYPES[info.container_bytes])
        st = lib.zflacb200_read(handle, out.ctypes.data, out.nbytes)
        if st != 0:
            _raise(st)
    finally:
        lib.zflacb200_close(handle)
    return DecodedFLAC(info.channels, info.sample_rate, info.bits_per_sample, out, _info_dict(info))


def decode_frames(data, byte_offset: int = 0, *, device: int = -1, use_crc_links: bool = True) -> DecodedFLAC:
    """Additive (not in the reference, whose Readme.md:54 lists "sync to next frame" as a TODO): decode a raw frame
    stream, or resync inside any stream from `byte_offset` on (zflacb200_open_frames). No signature, no metadata, no MD5
    gate; the first frame's header supplies channels / bits per sample / sample rate. `info["audio_offset"]` is the byte
    offset of the frame decoding began at."""
    lib = _native.load()
    arr = _as_u8(data)
    opt = _options(0 if use_crc_links else NO_CRC_ACCEL, device)
    handle = C.c_void_p()
    info = _native.Info()
    st = lib.zflacb200_open_frames(arr.ctypes.data, arr.size, C.byref(opt), int(byte_offset), C.byref(handle), C.byref(info))
    if st != 0:
        _raise(st)
    try:
        out = np.empty(int(info.n_samples), dtype=_DTYPES[info.container_bytes])
        st = lib.zflacb200_read(handle, out.ctypes.data, out.nbytes)
        if st != 0:
            _raise(st)
    finally:
        lib.zflacb200_close(handle)
    return DecodedFLAC(info.channels, info.sample_rate, info.bits_per_sample, out, _info_dict(info))


def decode_into(data, dst: np.ndarray, *, device: int = -1, verify_md5: bool = True, stream: int | None = None,
                frame_crc16: bool = True, md5_on_device: bool = False) -> dict:
    """One-call variant (zflacb200_decode): decodes into a caller-owned buffer (pinned host memory
    makes both copies run at full PCIe speed). frame_crc16=False skips the per-frame CRC-16 diagnostic."""
    lib = _native.load()
    arr = _as_u8(data)
    opt = _options((0 if verify_md5 else SKIP_MD5) | (0 if frame_crc16 else SKIP_FRAME_CRC16) |
                   (MD5_ON_DEVICE if md5_on_device else 0), device, stream)
    info = _native.Info()
    st = lib.zflacb200_decode(arr.ctypes.data, arr.size, C.byref(opt), dst.ctypes.data, dst.nbytes, C.byref(info))
    if st != 0:
        _raise(st)
    return _info_dict(info)


@dataclass
class RangeResult:
    """One shard of a frame-range decode (zflacb200_decode_range): the frames whose header starts in the byte range."""
    accepted: bool
    first_frame_offset: int
    end_offset: int
    first_sample: int
    n_samples: int          # per channel
    n_frames: int
    total_samples: int      # STREAMINFO total per channel
    samples: np.ndarray | None = None   # interleaved, n_samples * channels elements (None if not accepted / empty)
    info: dict = field(default_factory=dict)

    def as_row(self):
        """The seven scalars the stitch needs, as int64 (what ranks all-gather)."""
        return [int(self.accepted), self.first_frame_offset, self.end_offset, self.first_sample, self.n_samples, self.n_frames,
                self.total_samples]


def decode_range(data, byte_lo: int, byte_hi: int, dst: np.ndarray | None = None, *, device: int = -1,
                 stream: int | None = None, into_whole: bool = False) -> RangeResult:
    """Decodes the frames of ONE stream whose header starts in [byte_lo, byte_hi) (zflacb200_decode_range). `dst`, if
    given, receives the shard's PCM (bytes; pinned memory for full PCIe speed); else a buffer sized from the byte range's
    share of the stream's PCM (+ two maximum blocks) is allocated."""
    lib = _native.load()
    arr = _as_u8(data)
    if into_whole and dst is None:
        raise ValueError("into_whole needs the whole stream's PCM buffer as dst")
    opt = _options(SKIP_MD5 | (_native.RANGE_INTO_WHOLE if into_whole else 0), device, stream)
    info, rng = _native.Info(), _native.Range()
    rng.struct_size = C.sizeof(_native.Range)
    own = dst is None
    if own:  # the byte range's share of the PCM STREAMINFO promises, plus two maximum blocks; if that is short the call says how much
        dst = np.empty(0, dtype=np.uint8)
        try:
            if arr.size >= 42 and bytes(arr[:4]) == b"fLaC" and (int(arr[4]) & 0x7F) == 0:  # STREAMINFO, the first block
                si = arr[8:42]
                max_block = (int(si[2]) << 8) | int(si[3])
                packed = int.from_bytes(bytes(si[10:18]), "big")
                channels, depth, total = ((packed >> 41) & 7) + 1, ((packed >> 36) & 31) + 1, packed & 0xFFFFFFFFF
                cb = 1 if depth <= 8 else 2 if depth <= 16 else 4
                if total and total <= 65535 * (arr.size // 10 + 1):  # a total the bytes could hold
                    share = min(1.0, max(0, int(byte_hi) - int(byte_lo)) / max(1, arr.size - 42))
                    dst = np.empty((int(total * share) + 2 * max_block + 16) * channels * cb, dtype=np.uint8)
        except Exception:  # noqa: BLE001 - no stated total, a broken head: the first call reports the size (or the error)
            pass
    for attempt in range(2):
        d8 = dst.view(np.uint8).reshape(-1)
        st = lib.zflacb200_decode_range(arr.ctypes.data, arr.size, C.byref(opt), int(byte_lo), int(byte_hi),
                                        d8.ctypes.data if d8.size else None, d8.nbytes, C.byref(rng), C.byref(info))
        if st == 12 and own and attempt == 0:  # ZFLACB200_OUT_OF_MEMORY: now the size is known
            dst = np.empty(int(info.pcm_bytes), dtype=np.uint8)
            continue
        break
    if st != 0:
        _raise(st)
    samples = None
    if rng.accepted and rng.n_frames:
        off = int(rng.first_sample) * int(info.channels or 0) * int(info.container_bytes) if into_whole else 0
        samples = dst.view(np.uint8).reshape(-1)[off: off + int(info.pcm_bytes)].view(_DTYPES[info.container_bytes])
    return RangeResult(bool(rng.accepted), int(rng.first_frame_offset), int(rng.end_offset), int(rng.first_sample),
                       int(rng.n_samples), int(rng.n_frames), int(rng.total_samples), samples, _info_dict(info))


def stitch_ranges(rows, audio_offset: int) -> bool:
    """zflacb200_stitch over RangeResult objects or their as_row() lists, in stream order."""
    rows = [r.as_row() if isinstance(r, RangeResult) else [int(x) for x in r] for r in rows]
    buf = (_native.Range * max(len(rows), 1))()
    for b, r in zip(buf, rows):
        b.struct_size = C.sizeof(_native.Range)
        b.accepted, b.first_frame_offset, b.end_offset, b.first_sample, b.n_samples, b.n_frames, b.total_samples = r
    return bool(_native.load().zflacb200_stitch(buf, len(rows), int(audio_offset)))


def trim() -> None:
    """Hands the library's idle device blocks back to the driver (zflacb200_trim); open handles stay valid."""
    _native.load().zflacb200_trim()


BLOCK_TYPES = ("STREAMINFO", "PADDING", "APPLICATION", "SEEKTABLE", "VORBIS_COMMENT", "CUESHEET", "PICTURE")


def metadata(data) -> dict:
    """Additive, host only: what the reference's metadata walk steps over (src/zflac.zig:243-247) — the block list, the
    SEEKTABLE and the VORBIS_COMMENT fields. Raises the same errors decode() raises for a broken head."""
    lib = _native.load()
    arr = _as_u8(data)
    n, ao = C.c_size_t(), C.c_uint64()
    st = lib.zflacb200_metadata(arr.ctypes.data, arr.size, None, 0, C.byref(n), C.byref(ao))
    if st != 0:
        _raise(st)
    blocks = (_native.MetadataBlock * max(n.value, 1))()
    lib.zflacb200_metadata(arr.ctypes.data, arr.size, blocks, n.value, C.byref(n), C.byref(ao))
    out = {"audio_offset": ao.value,
           "blocks": [dict(type=BLOCK_TYPES[b.type], offset=b.offset, length=b.length, stored_bytes=b.stored_bytes,
                           is_last=bool(b.is_last)) for b in blocks[: n.value]]}
    st = lib.zflacb200_seektable(arr.ctypes.data, arr.size, None, 0, C.byref(n))
    if st != 0:
        _raise(st)
    pts = (_native.SeekPoint * max(n.value, 1))()
    lib.zflacb200_seektable(arr.ctypes.data, arr.size, pts, n.value, C.byref(n))
    out["seektable"] = [dict(sample_number=p.sample_number, stream_offset=p.stream_offset, frame_samples=p.frame_samples,
                             is_placeholder=bool(p.is_placeholder)) for p in pts[: n.value]]
    vendor = _native.Span()
    st = lib.zflacb200_vorbis_comments(arr.ctypes.data, arr.size, C.byref(vendor), None, 0, C.byref(n))
    if st != 0:  # a malformed comment block (flac-test-files faulty/10) does not stop the reference, which skips the block
        out.update(vendor="", comments=[], comments_error=lib.zflacb200_status_name(st).decode())
        return out
    spans = (_native.Span * max(n.value, 1))()
    lib.zflacb200_vorbis_comments(arr.ctypes.data, arr.size, C.byref(vendor), spans, n.value, C.byref(n))
    raw = arr.tobytes() if n.value or vendor.length else b""
    out["vendor"] = raw[vendor.offset: vendor.offset + vendor.length].decode("utf-8", "replace")
    out["comments"] = [raw[s.offset: s.offset + s.length].decode("utf-8", "replace") for s in spans[: n.value]]
    return out


def pictures(data) -> list:
    """Additive, host only: the PICTURE blocks (RFC 9639 8.8) the reference skips (src/zflac.zig:243-247). `image` is a
    (offset, length) span of the picture's bytes inside `data`."""
    lib = _native.load()
    arr = _as_u8(data)
    n = C.c_size_t()
    st = lib.zflacb200_pictures(arr.ctypes.data, arr.size, None, 0, C.byref(n))
    if st != 0:
        _raise(st)
    pics = (_native.Picture * max(n.value, 1))()
    lib.zflacb200_pictures(arr.ctypes.data, arr.size, pics, n.value, C.byref(n))
    raw = arr.tobytes() if n.value else b""
    return [dict(picture_type=p.picture_type, width=p.width, height=p.height, depth=p.depth, colors=p.colors,
                 mime=raw[p.mime.offset: p.mime.offset + p.mime.length].decode("ascii", "replace"),
                 description=raw[p.description.offset: p.description.offset + p.description.length].decode("utf-8", "replace"),
                 image=(p.image.offset, p.image.length)) for p in pics[: n.value]]


def cuesheet(data):
    """Additive, host only: the CUESHEET block (RFC 9639 8.7), or None when the stream has none."""
    lib = _native.load()
    arr = _as_u8(data)
    sheet, found = _native.CueSheet(), C.c_int()
    st = lib.zflacb200_cuesheet_read(arr.ctypes.data, arr.size, C.byref(sheet), None, 0, None, 0, C.byref(found))
    if st != 0:
        _raise(st)
    if not found.value:
        return None
    tracks = (_native.CueTrack * max(sheet.n_tracks, 1))()
    idx = (_native.CueIndex * max(sheet.n_indices, 1))()
    lib.zflacb200_cuesheet_read(arr.ctypes.data, arr.size, C.byref(sheet), tracks, sheet.n_tracks, idx, sheet.n_indices, C.byref(found))
    return dict(media_catalog_number=sheet.media_catalog_number.decode("ascii", "replace"), is_cd=bool(sheet.is_cd),
                lead_in_samples=sheet.lead_in_samples,
                tracks=[dict(offset_samples=t.offset_samples, number=t.number, isrc=t.isrc.decode("ascii", "replace"),
                             is_audio=bool(t.is_audio), pre_emphasis=bool(t.pre_emphasis),
                             indices=[dict(offset_samples=i.offset_samples, number=i.number)
                                      for i in idx[t.first_index: t.first_index + t.n_indices]])
                        for t in tracks[: sheet.n_tracks]])


class DeviceStream:
    """A decoded stream whose PCM stays in HBM (zflacb200_open with ZFLACB200_INPUT_ON_DEVICE)."""

    def __init__(self, handle, info):
        self._h = handle
        self.info = _info_dict(info)
        self.channels, self.sample_rate, self.bits_per_sample = info.channels, info.sample_rate, info.bits_per_sample
        self.container_bytes, self.n_samples, self.pcm_bytes = info.container_bytes, int(info.n_samples), int(info.pcm_bytes)

    def device_ptr(self) -> int:
        p, n = C.c_void_p(), C.c_size_t()
        _native.load().zflacb200_device_pcm(self._h, C.byref(p), C.byref(n))
        return p.value or 0

    def to_host(self) -> np.ndarray:
        out = np.empty(self.n_samples, dtype=_DTYPES[self.container_bytes])
        st = _native.load().zflacb200_read(self._h, out.ctypes.data, out.nbytes)
        if st != 0:
            _raise(st)
        return out

    def to_host_as(self, fmt: str) -> np.ndarray:
        """Read the PCM back in another output convention (zflacb200_read_as): "native", "f32" (normalised float),
        "s24" (packed 3-byte little-endian, s32 streams; returned as a uint8 array of 3 * n bytes), "u8" (s8 streams)."""
        lib = _native.load()
        code = FORMATS[fmt]
        need = C.c_uint64()
        st = lib.zflacb200_format_bytes(self._h, code, C.byref(need))
        if st != 0:
            _raise(st)
        dt = {0: _DTYPES[self.container_bytes], 1: np.float32, 2: np.uint8, 3: np.uint8}[code]
        out = np.empty(need.value // np.dtype(dt).itemsize, dtype=dt)
        st = lib.zflacb200_read_as(self._h, code, out.ctypes.data, out.nbytes)
        if st != 0:
            _raise(st)
        return out

    def frames(self) -> np.ndarray:
        lib = _native.load()
        n = C.c_size_t()
        lib.zflacb200_frames(self._h, None, 0, C.byref(n))
        buf = (_native.Frame * max(n.value, 1))()
        lib.zflacb200_frames(self._h, buf, n.value, C.byref(n))
        dt = np.dtype([("offset", "<u8"), ("end_offset", "<u8"), ("sample_offset", "<u8"), ("block_size", "<u4"),
                       ("header_len", "u1"), ("channel_code", "u1"), ("crc16_ok", "u1"), ("probed", "u1")])
        return np.frombuffer(bytes(buf), dtype=dt)[: n.value].copy()

    def close(self):
        if self._h:
            _native.load().zflacb200_close(self._h)
            self._h = None

    def __del__(self):
        self.close()


def open_device(dev_ptr: int, nbytes: int, *, device: int = -1, stream: int | None = None, verify_md5: bool = True,
                use_crc_links: bool = True, use_fast_path: bool = True, host_data=None, frame_crc16: bool = True,
                md5_on_device: bool = False) -> DeviceStream:
    """Decode a stream that already sits in device memory (or, with host_data, upload it first)
    and keep the PCM on the device."""
    lib = _native.load()
    flags = (0 if verify_md5 else SKIP_MD5) | (0 if use_crc_links else NO_CRC_ACCEL) | (0 if use_fast_path else NO_FAST_PATH)
    flags |= 0 if frame_crc16 else SKIP_FRAME_CRC16
    flags |= MD5_ON_DEVICE if md5_on_device else 0
    if host_data is not None:
        arr = _as_u8(host_data)
        ptr, n = arr.ctypes.data, arr.size
    else:
        flags |= INPUT_ON_DEVICE
        ptr, n = dev_ptr, nbytes
    opt = _options(flags, device, stream)
    handle = C.c_void_p()
    info = _native.Info()
    st = lib.zflacb200_open(ptr, n, C.byref(opt), C.byref(handle), C.byref(info))
    if st != 0:
        _raise(st)
    return DeviceStream(handle, info)


# --------------------------------------------------------------------------------------------------
# File-parallel batch (BASELINE config 5): many independent streams, one pipeline pass per device.
# --------------------------------------------------------------------------------------------------
def pack_streams(streams, align: int = 16):
    """Packs byte strings / uint8 arrays back to back (each start `align`-byte aligned) into one blob.
    Returns (blob uint8 ndarray, offsets uint64 ndarray, lens uint64 ndarray)."""
    arrs = [_as_u8(x) for x in streams]
    lens = np.array([a.size for a in arrs], dtype=np.uint64)
    padded = (lens + np.uint64(align - 1)) // np.uint64(align) * np.uint64(align)
    offsets = np.zeros(len(arrs), dtype=np.uint64)
    if len(arrs) > 1:
        offsets[1:] = np.cumsum(padded[:-1])
    blob = np.zeros(int(padded.sum()) if len(arrs) else 0, dtype=np.uint8)
    for a, o in zip(arrs, offsets):
        blob[int(o):int(o) + a.size] = a
    return blob, offsets, lens


class _LazyInfos:
    """Sequence view over the C info array; element i becomes a dict when it is read."""

    def __init__(self, raw, n):
        self._raw, self._n = raw, n

    def __len__(self):
        return self._n

    def __getitem__(self, i):
        if isinstance(i, slice):
            return [_info_dict(self._raw[k]) for k in range(*i.indices(self._n))]
        if i < 0:
            i += self._n
        if not 0 <= i < self._n:
            raise IndexError(i)
        return _info_dict(self._raw[i])

    def __iter__(self):
        return (_info_dict(self._raw[k]) for k in range(self._n))


class DeviceBatch:
    """Result of zflacb200_batch_open: per-stream statuses / infos, PCM resident in HBM."""

    def __init__(self, handle, infos, statuses, n):
        self._h = handle
        self.n = n
        self.statuses = np.frombuffer(statuses, dtype=np.int32, count=n).tolist() if n else []
        self.errors = [None if st == 0 else _native.load().zflacb200_status_name(st).decode() for st in self.statuses]
        self._raw_infos = infos
        self.infos = _LazyInfos(infos, n)  # dicts are built on access: 10k eager conversions cost more than the decode

    def pcm_bytes(self) -> int:
        v = C.c_uint64()
        _native.load().zflacb200_batch_pcm_bytes(self._h, C.byref(v))
        return int(v.value)

    def read_all(self, dst: np.ndarray | None = None):
        """One bulk transfer of every decoded stream's PCM. Returns (uint8 buffer, byte offsets; 2**64-1 = no data)."""
        if dst is None:
            dst = np.empty(self.pcm_bytes(), dtype=np.uint8)
        offs = np.zeros(self.n, dtype=np.uint64)
        st = _native.load().zflacb200_batch_read_all(self._h, dst.ctypes.data, dst.nbytes, offs.ctypes.data_as(C.POINTER(C.c_uint64)))
        if st != 0:
            _raise(st)
        return dst, offs

    def stream(self, i: int) -> DecodedFLAC:
        """Stream i as the reference's decode() would have returned it (raises its error otherwise)."""
        if self.statuses[i] != 0:
            _raise(self.statuses[i])
        inf = self._raw_infos[i]
        out = np.empty(int(inf.n_samples), dtype=_DTYPES[inf.container_bytes])
        st = _native.load().zflacb200_batch_read_stream(self._h, i, out.ctypes.data, out.nbytes)
        if st != 0:
            _raise(st)
        return DecodedFLAC(inf.channels, inf.sample_rate, inf.bits_per_sample, out, self.infos[i])

    def streams_from(self, buf: np.ndarray, offs: np.ndarray):
        """Views of the per-stream sample arrays inside a read_all() buffer (None where a stream failed)."""
        out = []
        for i in range(self.n):
            if self.statuses[i] != 0 or int(offs[i]) == 2**64 - 1:
                out.append(None)
                continue
            inf = self._raw_infos[i]
            o = int(offs[i])
            out.append(buf[o:o + int(inf.pcm_bytes)].view(_DTYPES[inf.container_bytes]))
        return out

    def close(self):
        if self._h:
            _native.load().zflacb200_batch_close(self._h)
            self._h = None

    def __del__(self):
        self.close()


def open_batch(blob, offsets, lens, *, device_blob_ptr: int = 0, device: int = -1, stream: int | None = None,
               verify_md5: bool = True, use_fast_path: bool = True) -> DeviceBatch:
    """zflacb200_batch_open over a packed blob (see pack_streams). `device_blob_ptr`: device copy of the blob, if any."""
    lib = _native.load()
    blob = _as_u8(blob)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    lens = np.ascontiguousarray(lens, dtype=np.uint64)
    n = int(offsets.size)
    flags = (0 if verify_md5 else SKIP_MD5) | (0 if use_fast_path else NO_FAST_PATH)
    opt = _options(flags, device, stream)
    infos = (_native.Info * max(n, 1))()
    statuses = (C.c_int * max(n, 1))()
    handle = C.c_void_p()
    st = lib.zflacb200_batch_open(blob.ctypes.data, device_blob_ptr or None, offsets.ctypes.data_as(C.POINTER(C.c_uint64)),
                                  lens.ctypes.data_as(C.POINTER(C.c_uint64)), n, C.byref(opt), C.byref(handle), infos, statuses)
    if st != 0:
        _raise(st)
    return DeviceBatch(handle, infos, statuses, n)


def batch_pcm_bound(blob, offsets, lens) -> int:
    """Bytes zflacb200_batch_decode needs in its destination (from the STREAMINFO totals, host only)."""
    blob = _as_u8(blob)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    lens = np.ascontiguousarray(lens, dtype=np.uint64)
    v = C.c_uint64()
    st = _native.load().zflacb200_batch_pcm_bound(blob.ctypes.data, offsets.ctypes.data_as(C.POINTER(C.c_uint64)),
                                                  lens.ctypes.data_as(C.POINTER(C.c_uint64)), int(offsets.size), C.byref(v))
    if st != 0:
        _raise(st)
    return int(v.value)


class BatchResult:
    """What zflacb200_batch_decode returned: statuses, infos, and where each stream's PCM sits in the caller's buffer."""

    def __init__(self, infos, statuses, offs, dst, n):
        self.n = n
        self._raw_infos = infos
        self.infos = _LazyInfos(infos, n)
        self.statuses = np.frombuffer(statuses, dtype=np.int32, count=n).tolist() if n else []
        self.pcm_offsets = offs
        self.dst = dst

    def samples(self, i: int):
        """View of stream i's samples inside the destination buffer (None if it did not decode)."""
        if self.statuses[i] != 0 or int(self.pcm_offsets[i]) == 2**64 - 1:
            return None
        inf = self._raw_infos[i]
        o = int(self.pcm_offsets[i])
        return self.dst[o:o + int(inf.pcm_bytes)].view(_DTYPES[inf.container_bytes])


def decode_batch_into(blob, offsets, lens, dst: np.ndarray, *, device: int = -1, stream: int | None = None,
                      verify_md5: bool = True, use_fast_path: bool = True) -> BatchResult:
    """zflacb200_batch_decode: host blob in, host PCM out, upload / decode / read-back overlapped over sub-batches."""
    lib = _native.load()
    blob = _as_u8(blob)
    offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
    lens = np.ascontiguousarray(lens, dtype=np.uint64)
    n = int(offsets.size)
    flags = (0 if verify_md5 else SKIP_MD5) | (0 if use_fast_path else NO_FAST_PATH)
    opt = _options(flags, device, stream)
    infos = (_native.Info * max(n, 1))()
    statuses = (C.c_int * max(n, 1))()
    offs = np.zeros(max(n, 1), dtype=np.uint64)
    dst8 = dst.view(np.uint8).reshape(-1)
    st = lib.zflacb200_batch_decode(blob.ctypes.data, offsets.ctypes.data_as(C.POINTER(C.c_uint64)),
                                    lens.ctypes.data_as(C.POINTER(C.c_uint64)), n, C.byref(opt), dst8.ctypes.data, dst8.nbytes,
                                    offs.ctypes.data_as(C.POINTER(C.c_uint64)), infos, statuses)
    if st != 0:
        _raise(st)
    return BatchResult(infos, statuses, offs, dst8, n)


def decode_batch(streams, *, device: int = -1, verify_md5: bool = True, use_fast_path: bool = True):
    """Decodes a list of independent streams in one pipeline pass. Returns a list with, per stream, a DecodedFLAC
    or the ZflacError instance the reference's decode() would have raised."""
    blob, offsets, lens = pack_streams(streams)
    b = open_batch(blob, offsets, lens, device=device, verify_md5=verify_md5, use_fast_path=use_fast_path)
    try:
        buf, offs = b.read_all()
        views = b.streams_from(buf, offs)
        out = []
        for i in range(b.n):
            if b.statuses[i] != 0:
                out.append(_ERRORS.get(b.statuses[i], ZflacError)(b.errors[i]))
            else:
                inf = b._raw_infos[i]
                out.append(DecodedFLAC(inf.channels, inf.sample_rate, inf.bits_per_sample, views[i].copy(), b.infos[i]))
        return out
    finally:
        b.close()
