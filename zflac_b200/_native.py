"""ctypes binding of libzflacb200.so (include/zflacb200.h). Fails loudly when the library is
missing or cannot be loaded: there is no Python or CPU decode fallback."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "_lib", "libzflacb200.so")

# zflacb200_options.flags
VERIFY_MD5 = 0x1
SKIP_MD5 = 0x2
SKIP_FRAME_CRC16 = 0x80
INPUT_ON_DEVICE = 0x4
NO_CRC_ACCEL = 0x8
NO_FAST_PATH = 0x10
MD5_ON_DEVICE = 0x100
RANGE_INTO_WHOLE = 0x200
WORD_PIPELINE = 0x400
SINGLE_PASS = 0x20

EXPORTS = [
    "zflacb200_abi_version", "zflacb200_init", "zflacb200_device_count", "zflacb200_shutdown", "zflacb200_trim", "zflacb200_scan_no_alloc",
    "zflacb200_status_name", "zflacb200_last_error", "zflacb200_kernel_launches", "zflacb200_pass_a_launches", "zflacb200_open", "zflacb200_read",
    "zflacb200_device_pcm", "zflacb200_close", "zflacb200_decode", "zflacb200_pcm_bound", "zflacb200_frames", "zflacb200_batch_open",
    "zflacb200_batch_pcm_bytes", "zflacb200_batch_read_all", "zflacb200_batch_read_stream", "zflacb200_batch_device_pcm",
    "zflacb200_batch_close", "zflacb200_batch_pcm_bound", "zflacb200_batch_decode",
    "zflacb200_decode_range", "zflacb200_stitch", "zflacb200_open_frames",
    "zflacb200_format_bytes", "zflacb200_read_as", "zflacb200_metadata", "zflacb200_seektable", "zflacb200_vorbis_comments", "zflacb200_pictures", "zflacb200_cuesheet_read",
]


class Options(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("flags", C.c_uint32), ("device", C.c_int32), ("stream", C.c_void_p)]


class Info(C.Structure):
    _fields_ = [("channels", C.c_uint8), ("bits_per_sample", C.c_uint8), ("container_bytes", C.c_uint8),
                ("md5_checked", C.c_uint8), ("sample_rate", C.c_uint32), ("n_samples", C.c_uint64),
                ("pcm_bytes", C.c_uint64), ("n_frames", C.c_uint64), ("audio_offset", C.c_uint64),
                ("audio_bytes", C.c_uint64), ("n_candidates", C.c_uint64), ("n_probed_frames", C.c_uint64),
                ("n_crc16_mismatch", C.c_uint64), ("fast_path", C.c_uint64), ("md5_computed", C.c_uint8 * 16),
                ("md5_streaminfo", C.c_uint8 * 16), ("ms_scan", C.c_float), ("ms_link", C.c_float),
                ("ms_decode", C.c_float), ("ms_epilogue", C.c_float), ("ms_md5", C.c_float), ("ms_total", C.c_float)]


class Range(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("accepted", C.c_uint32), ("first_frame_offset", C.c_uint64),
                ("end_offset", C.c_uint64), ("first_sample", C.c_uint64), ("n_samples", C.c_uint64),
                ("n_frames", C.c_uint64), ("total_samples", C.c_uint64)]


class MetadataBlock(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("length", C.c_uint32), ("stored_bytes", C.c_uint32), ("type", C.c_uint8),
                ("is_last", C.c_uint8), ("pad", C.c_uint8 * 6)]


class SeekPoint(C.Structure):
    _fields_ = [("sample_number", C.c_uint64), ("stream_offset", C.c_uint64), ("frame_samples", C.c_uint32),
                ("is_placeholder", C.c_uint32)]


class Span(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("length", C.c_uint64)]


class Picture(C.Structure):
    _fields_ = [("picture_type", C.c_uint32), ("width", C.c_uint32), ("height", C.c_uint32), ("depth", C.c_uint32),
                ("colors", C.c_uint32), ("pad", C.c_uint32), ("mime", Span), ("description", Span), ("image", Span)]


class CueSheet(C.Structure):
    _fields_ = [("media_catalog_number", C.c_char * 129), ("is_cd", C.c_uint8), ("pad", C.c_uint8 * 6),
                ("lead_in_samples", C.c_uint64), ("n_tracks", C.c_uint32), ("n_indices", C.c_uint32)]


class CueTrack(C.Structure):
    _fields_ = [("offset_samples", C.c_uint64), ("isrc", C.c_char * 13), ("number", C.c_uint8), ("is_audio", C.c_uint8),
                ("pre_emphasis", C.c_uint8), ("first_index", C.c_uint32), ("n_indices", C.c_uint32)]


class CueIndex(C.Structure):
    _fields_ = [("offset_samples", C.c_uint64), ("number", C.c_uint32), ("pad", C.c_uint32)]


class Frame(C.Structure):
    _fields_ = [("offset", C.c_uint64), ("end_offset", C.c_uint64), ("sample_offset", C.c_uint64),
                ("block_size", C.c_uint32), ("header_len", C.c_uint8), ("channel_code", C.c_uint8),
                ("crc16_ok", C.c_uint8), ("probed", C.c_uint8)]


_lib = None


def load():
    """Load libzflacb200.so. Raises if it has not been built (run __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: the CUDA library has not been built. Run "
            "`python -c 'import __graft_entry__ as g; g.build()'` (needs nvcc). zflac_b200 has no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    lib.zflacb200_abi_version.restype = C.c_int
    lib.zflacb200_init.argtypes = [C.c_int]
    lib.zflacb200_init.restype = C.c_int
    lib.zflacb200_device_count.restype = C.c_int
    lib.zflacb200_shutdown.restype = None
    lib.zflacb200_trim.restype = None
    lib.zflacb200_scan_no_alloc.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_void_p]
    lib.zflacb200_scan_no_alloc.restype = C.c_int
    lib.zflacb200_status_name.argtypes = [C.c_int]
    lib.zflacb200_status_name.restype = C.c_char_p
    lib.zflacb200_last_error.restype = C.c_char_p
    lib.zflacb200_kernel_launches.restype = C.c_uint64
    lib.zflacb200_pass_a_launches.restype = C.c_uint64
    lib.zflacb200_pass_a_launches.argtypes = [C.c_int]
    lib.zflacb200_open.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Options), C.POINTER(C.c_void_p), C.POINTER(Info)]
    lib.zflacb200_open.restype = C.c_int
    lib.zflacb200_open_frames.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Options), C.c_uint64, C.POINTER(C.c_void_p),
                                          C.POINTER(Info)]
    lib.zflacb200_open_frames.restype = C.c_int
    lib.zflacb200_read.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
    lib.zflacb200_read.restype = C.c_int
    lib.zflacb200_format_bytes.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]
    lib.zflacb200_format_bytes.restype = C.c_int
    lib.zflacb200_read_as.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_size_t]
    lib.zflacb200_read_as.restype = C.c_int
    lib.zflacb200_metadata.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(MetadataBlock), C.c_size_t, C.POINTER(C.c_size_t),
                                       C.POINTER(C.c_uint64)]
    lib.zflacb200_metadata.restype = C.c_int
    lib.zflacb200_seektable.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(SeekPoint), C.c_size_t, C.POINTER(C.c_size_t)]
    lib.zflacb200_seektable.restype = C.c_int
    lib.zflacb200_vorbis_comments.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Span), C.POINTER(Span), C.c_size_t,
                                              C.POINTER(C.c_size_t)]
    lib.zflacb200_vorbis_comments.restype = C.c_int
    lib.zflacb200_pictures.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Picture), C.c_size_t, C.POINTER(C.c_size_t)]
    lib.zflacb200_pictures.restype = C.c_int
    lib.zflacb200_cuesheet_read.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(CueSheet), C.POINTER(CueTrack), C.c_size_t,
                                            C.POINTER(CueIndex), C.c_size_t, C.POINTER(C.c_int)]
    lib.zflacb200_cuesheet_read.restype = C.c_int
    lib.zflacb200_pcm_bound.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]
    lib.zflacb200_pcm_bound.restype = C.c_int
    lib.zflacb200_device_pcm.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.zflacb200_device_pcm.restype = C.c_int
    lib.zflacb200_close.argtypes = [C.c_void_p]
    lib.zflacb200_close.restype = None
    lib.zflacb200_decode.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Options), C.c_void_p, C.c_size_t, C.POINTER(Info)]
    lib.zflacb200_decode.restype = C.c_int
    lib.zflacb200_batch_open.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_size_t,
                                         C.POINTER(Options), C.POINTER(C.c_void_p), C.POINTER(Info), C.POINTER(C.c_int)]
    lib.zflacb200_batch_open.restype = C.c_int
    lib.zflacb200_batch_pcm_bytes.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
    lib.zflacb200_batch_pcm_bytes.restype = C.c_int
    lib.zflacb200_batch_read_all.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64)]
    lib.zflacb200_batch_read_all.restype = C.c_int
    lib.zflacb200_batch_read_stream.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    lib.zflacb200_batch_read_stream.restype = C.c_int
    lib.zflacb200_batch_device_pcm.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(C.c_void_p), C.POINTER(C.c_size_t)]
    lib.zflacb200_batch_device_pcm.restype = C.c_int
    lib.zflacb200_batch_close.argtypes = [C.c_void_p]
    lib.zflacb200_batch_close.restype = None
    lib.zflacb200_decode_range.argtypes = [C.c_void_p, C.c_size_t, C.POINTER(Options), C.c_uint64, C.c_uint64, C.c_void_p,
                                           C.c_size_t, C.POINTER(Range), C.POINTER(Info)]
    lib.zflacb200_decode_range.restype = C.c_int
    lib.zflacb200_stitch.argtypes = [C.POINTER(Range), C.c_size_t, C.c_uint64]
    lib.zflacb200_stitch.restype = C.c_int
    lib.zflacb200_batch_pcm_bound.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(C.c_uint64)]
    lib.zflacb200_batch_pcm_bound.restype = C.c_int
    lib.zflacb200_batch_decode.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_size_t, C.POINTER(Options),
                                           C.c_void_p, C.c_size_t, C.POINTER(C.c_uint64), C.POINTER(Info), C.POINTER(C.c_int)]
    lib.zflacb200_batch_decode.restype = C.c_int
    lib.zflacb200_frames.argtypes = [C.c_void_p, C.POINTER(Frame), C.c_size_t, C.POINTER(C.c_size_t)]
    lib.zflacb200_frames.restype = C.c_int
    _lib = lib
    return lib
