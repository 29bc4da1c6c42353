"""CPU unit test of the K2 per-frame decoder logic (zflac_b200/csrc/decode_core.cuh).

tests/host_core/host_core.cu instantiates the SAME header the CUDA kernel uses with a host word source;
here every frame of every corpus file / synthetic stream is decoded through it and compared with the
oracle: frame end offsets, and the PCM after a numpy restatement of K3 (decorrelation, MSB-align shift,
interleave). This pins the bit reader, Rice fast/slow paths, partition walk and predictor windows on CPU;
the `-m gpu` tests pin the device word ring on top. Test infrastructure only — nothing in zflac_b200/
loads this library.
"""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from conftest import ROOT, corpus_bytes, corpus_entries, corpus_ids
from oracle import oracle
import synth_cases

HERE = os.path.join(ROOT, "tests", "host_core")
LIB = os.path.join(HERE, "_build", "libzf_host_core.so")
SRC = os.path.join(HERE, "host_core.cu")
DEPS = [SRC] + [os.path.join(ROOT, "zflac_b200", "csrc", f) for f in
                ("decode_core.cuh", "parse_core.cuh", "common.cuh", "pos_reader.cuh", "parse_narrow.cuh")]

FRAME_REC = np.dtype([("offset", "<u8"), ("stream_end", "<u8"), ("sample_base", "<u8"), ("scratch_base", "<u8"),
                      ("block_size", "<u4"), ("scratch_pitch", "<u4"), ("header_len", "u1"), ("ch_code", "u1"),
                      ("channels", "u1"), ("bps", "u1"), ("stream_id", "<u4")])
FRAME_OUT = np.dtype([("end_offset", "<u8"), ("status", "<u4"), ("stored_crc16", "<u4")])


@pytest.fixture(scope="module")
def core():
    stale = (not os.path.exists(LIB)) or any(os.path.getmtime(d) > os.path.getmtime(LIB) for d in DEPS)
    if stale:
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc, "-O2", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
                        "--expt-relaxed-constexpr", "-o", LIB, SRC], check=True, capture_output=True)
    lib = C.CDLL(LIB)
    lib.zf_host_decode_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int]
    lib.zf_host_decode_frame.restype = C.c_int
    assert lib.zf_host_framerec_size() == FRAME_REC.itemsize
    lib.zf_host_parse_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.zf_host_parse_frame.restype = C.c_int
    lib.zf_host_predict_plane.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]
    lib.zf_host_predict_plane.restype = None
    assert lib.zf_host_subrec_size() == 80
    lib.zf_host_parse_frame_narrow.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                               C.c_int, C.c_int, C.c_int]
    lib.zf_host_parse_frame_narrow.restype = C.c_int
    return lib


def host_decode(lib, data: bytes, ref, two_pass: bool = False, narrow: bool = False, stats=None, pos_wide: bool = False):
    """Decode all frames of the oracle's frame trace with the host core; return interleaved PCM like the reference's.
    two_pass: pass A (parse_core.cuh) + a scalar restatement of pass B instead of the single-pass decoder.
    narrow: pass A is parse_narrow.cuh (PosReader, folded 16-bit scratch); a frame that reports overflow is redone with
    parse_core.cuh, the way the host library falls back (stats["overflow_frames"] counts them).
    pos_wide: pass A is parse_narrow.cuh with int32 rows (WIDE), both layouts alternating; nothing can overflow."""
    arr = np.frombuffer(data, dtype=np.uint8)
    ch, bps = ref.channels, ref.bits_per_sample
    wide = bps == 32
    cb = ref.container_bytes
    n_total = int(ref.samples.size)
    pcm = np.zeros(n_total, dtype=np.int64)
    depth_si = bps  # corpus/synth: STREAMINFO depth == first-frame depth whenever decode succeeds with a valid MD5
    for fr in ref.frames:
        bs = int(fr["block_size"])
        pitch = (bs + 7) // 8 * 8
        rec = np.zeros(1, dtype=FRAME_REC)
        rec["offset"], rec["stream_end"] = fr["offset"], len(data)
        rec["block_size"], rec["scratch_pitch"] = bs, pitch
        rec["header_len"], rec["ch_code"], rec["channels"], rec["bps"] = fr["header_len"], fr["channel_code"], ch, bps
        scratch = np.zeros(pitch * ch, dtype=np.int64 if wide else np.int32)
        out = np.zeros(1, dtype=FRAME_OUT)
        if two_pass:
            assert not wide
            subs = np.zeros(8 * 80, dtype=np.uint8)
            done = False
            if pos_wide:
                nb = pitch // 8
                warp_layout = (int(fr["offset"]) // 7) % 2 == 1
                lane = int(fr["offset"]) % 32
                s32 = np.zeros(ch * nb * 2 * 32 * 4 if warp_layout else pitch * ch, dtype=np.int32)
                ovf = np.zeros(1, dtype=np.uint32)
                viol = lib.zf_host_parse_frame_narrow(arr.ctypes.data, rec.ctypes.data, s32.ctypes.data, subs.ctypes.data, 1,
                                                      out.ctypes.data, ovf.ctypes.data, int(warp_layout), lane, 1)
                assert viol == 0, "unchecked ring entry read outside the top_up guarantee"
                assert int(ovf[0]) == 0
                if warp_layout:  # half h of row t of channel c of lane L at (((c * nb + t) * 2 + h) * 32 + L) * 16 bytes
                    rows = s32.reshape(ch, nb, 2, 32, 4)
                    assert not rows[:, :, :, np.arange(32) != lane, :].any(), "wrote into another lane's rows"
                    planes32 = rows[:, :, :, lane, :].reshape(ch, pitch)
                    scratch = planes32.reshape(ch, pitch // 8, 8).transpose(1, 0, 2).reshape(-1).copy()  # the sub_idx() layout
                else:
                    scratch = s32  # the per-frame layout IS sub_idx()
                done = True
            elif narrow:
                # both row layouts of parse_narrow.cuh, alternating from frame to frame: the frame's own region (groups of
                # 16 steps x channels x 16 values) and the warp-interleaved one (row t of channel c of lane L at
                # ((c * nb + t) * 32 + L) * 16 bytes), here as a lane that changes with the frame
                nb = pitch // 8
                warp_layout = (int(fr["offset"]) // 7) % 2 == 1
                lane = int(fr["offset"]) % 32
                s16 = np.zeros(ch * nb * 32 * 8 if warp_layout else (bs + 15) // 16 * 16 * ch, dtype=np.uint16)
                ovf = np.zeros(1, dtype=np.uint32)
                viol = lib.zf_host_parse_frame_narrow(arr.ctypes.data, rec.ctypes.data, s16.ctypes.data, subs.ctypes.data, 1,
                                                      out.ctypes.data, ovf.ctypes.data, int(warp_layout), lane, 0)
                assert viol == 0, "unchecked ring entry read outside the top_up guarantee"
                if int(ovf[0]) == 0:
                    if warp_layout:
                        rows = s16.view(np.int16).astype(np.int32).reshape(ch, nb, 32, 8)
                        assert not rows[:, :, np.arange(32) != lane, :].any(), "wrote into another lane's rows"
                        planes16 = rows[:, :, lane, :].reshape(ch, pitch)
                    else:
                        val = s16.view(np.int16).astype(np.int32).reshape(-1, ch, 16)
                        planes16 = val.transpose(1, 0, 2).reshape(ch, -1)[:, :pitch]   # [channel][time step]
                    scratch = planes16.reshape(ch, pitch // 8, 8).transpose(1, 0, 2).reshape(-1).copy()  # the int32 sub_idx() layout
                    done = True
                elif stats is not None:
                    stats["overflow_frames"] = stats.get("overflow_frames", 0) + 1
            if not done:
                viol = lib.zf_host_parse_frame(arr.ctypes.data, rec.ctypes.data, scratch.ctypes.data, subs.ctypes.data, 1, out.ctypes.data)
            if int(out["status"][0]) == 0:
                planar = np.zeros(pitch * ch, dtype=np.int32)
                for c in range(ch):
                    lib.zf_host_predict_plane(subs.ctypes.data + 80 * c, scratch.ctypes.data, c, ch, bs, planar.ctypes.data + 4 * pitch * c)
                scratch = planar
        else:
            viol = lib.zf_host_decode_frame(arr.ctypes.data, rec.ctypes.data, scratch.ctypes.data, int(wide), 1, out.ctypes.data, int(cb == 4))
        assert viol == 0, "unchecked word read beyond the top_up guarantee"
        assert int(out["status"][0]) == 0, (int(fr["offset"]), int(out["status"][0]))
        assert int(out["end_offset"][0]) == int(fr["end_offset"]), int(fr["offset"])
        # the reference narrows every subframe value into the container and shifts the wasted bits back in THERE (:494, :537:
        # `<<=` drops what leaves the type), so the decorrelation sees container values
        planes = scratch.reshape(ch, pitch)[:, :bs].astype({1: np.int8, 2: np.int16, 4: np.int32}[cb]).astype(np.int64)
        code = int(fr["channel_code"])
        if code == 8:
            planes[1] = planes[0] - planes[1]
        elif code == 9:
            planes[0] = planes[0] + planes[1]
        elif code == 10:
            mid = (planes[0] << 1) + (planes[1] & 1)
            side = planes[1].copy()
            planes[0] = (mid + side) >> 1
            planes[1] = (mid - side) >> 1
        base = int(fr["sample_offset"])
        pcm[base:base + bs * ch] = planes.T.reshape(-1)
    shift = 0
    if 9 <= depth_si <= 15:
        shift = 16 - depth_si
    elif 17 <= depth_si <= 31:
        shift = 32 - depth_si
    dt = {1: np.int8, 2: np.int16, 4: np.int32}[cb]
    return (pcm << shift).astype(dt)


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]],
                         ids=[i for i, e in zip(corpus_ids(), corpus_entries()) if not e["expect_error"]])
def test_host_core_corpus(core, entry):
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    if ref.error != "OK":
        pytest.skip(f"oracle: {ref.error}")
    got = host_decode(core, data, ref)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]],
                         ids=[i for i, e in zip(corpus_ids(), corpus_entries()) if not e["expect_error"]])
def test_host_parse_core_corpus(core, entry):
    """Pass A (parse_core.cuh: residuals + SubRec) followed by the recurrence == the oracle, frame ends included."""
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    if ref.error != "OK":
        pytest.skip(f"oracle: {ref.error}")
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    got = host_decode(core, data, ref, two_pass=True)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("name", sorted(synth_cases.CASES))
def test_host_parse_core_synth(core, name):
    s = synth_cases.make(name)
    data = s.flac.tobytes()
    ref = oracle.decode(data, trace_frames=True)
    assert ref.error == "OK"
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    got = host_decode(core, data, ref, two_pass=True)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("name", sorted(synth_cases.CASES))
def test_host_core_synth(core, name):
    s = synth_cases.make(name)
    data = s.flac.tobytes()
    ref = oracle.decode(data, trace_frames=True)
    assert ref.error == "OK"
    got = host_decode(core, data, ref)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]],
                         ids=[i for i, e in zip(corpus_ids(), corpus_entries()) if not e["expect_error"]])
def test_host_parse_narrow_corpus(core, entry):
    """Narrow pass A (parse_narrow.cuh: PosReader, float-exponent unary count, folded 16-bit scratch) + the recurrence ==
    the oracle, frame ends included. Streams deeper than 16 bits run too: frames whose values do not fit report overflow."""
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    if ref.error != "OK":
        pytest.skip(f"oracle: {ref.error}")
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    stats = {}
    got = host_decode(core, data, ref, two_pass=True, narrow=True, stats=stats)
    assert np.array_equal(got, ref.samples)
    if ref.bits_per_sample <= 16 and entry["id"] not in ("61",):  # 61: "predictor overflow check", full-scale residuals
        assert stats.get("overflow_frames", 0) == 0


@pytest.mark.parametrize("name", sorted(synth_cases.CASES))
def test_host_parse_narrow_synth(core, name):
    s = synth_cases.make(name)
    data = s.flac.tobytes()
    ref = oracle.decode(data, trace_frames=True)
    assert ref.error == "OK"
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    got = host_decode(core, data, ref, two_pass=True, narrow=True)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]],
                         ids=[i for i, e in zip(corpus_ids(), corpus_entries()) if not e["expect_error"]])
def test_host_parse_pos_wide_corpus(core, entry):
    """Pass A with the position reader and int32 rows (parse_narrow.cuh, WIDE: what 17..31-bit streams take, and the
    rerun of a narrower launch group that overflowed 16 bits) + the recurrence == the oracle, frame ends included, on
    every stream of the corpus up to 31 bits per sample, both row layouts."""
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    if ref.error != "OK":
        pytest.skip(f"oracle: {ref.error}")
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    got = host_decode(core, data, ref, two_pass=True, pos_wide=True)
    assert np.array_equal(got, ref.samples)


@pytest.mark.parametrize("name", sorted(synth_cases.CASES))
def test_host_parse_pos_wide_synth(core, name):
    s = synth_cases.make(name)
    data = s.flac.tobytes()
    ref = oracle.decode(data, trace_frames=True)
    assert ref.error == "OK"
    if ref.bits_per_sample == 32:
        pytest.skip("32 bps streams stay on the single-pass int64 decoder")
    got = host_decode(core, data, ref, two_pass=True, pos_wide=True)
    assert np.array_equal(got, ref.samples)


def test_host_core_samples_wider_than_the_frame_depth(core):
    """decode_core.cuh's `inter64` and the scalar pass B against the oracle's i64 sums (bitcraft.wide_sample_stream)."""
    from bitcraft import wide_sample_stream
    data = wide_sample_stream()
    ref = oracle.decode(data, trace_frames=True)
    assert ref.error == "OK" and len(ref.frames) == 40
    s1 = (1 << 29) + 12345 + ((127 * 100) >> 15)
    assert int(ref.samples[2]) == (((7 + ((127 * s1) >> 15)) << 8) & 0xFFFFFFFF)  # the i64 sum, MSB-aligned (:287-306)
    for two_pass in (False, True):
        assert np.array_equal(host_decode(core, data.tobytes(), ref, two_pass=two_pass), ref.samples)


FUZZ_SEEDS = list(range(250))


def test_host_core_random_streams(core):
    """Any legal field combination (tests/bitcraft.py: every subframe type and order, wasted bits, both Rice methods, all
    partition orders, parameters far from the ideal one, escapes, all channel assignments, depths 8..24, odd block sizes):
    the single-pass decoder and the three forms of pass A against the oracle."""
    from bitcraft import fuzz_case
    done = 0
    for seed in FUZZ_SEEDS:
        case = fuzz_case(seed)
        if case is None:
            continue
        data, ref, _ = case
        raw = data.tobytes()
        assert np.array_equal(host_decode(core, raw, ref), ref.samples), seed
        assert np.array_equal(host_decode(core, raw, ref, two_pass=True), ref.samples), seed
        assert np.array_equal(host_decode(core, raw, ref, two_pass=True, pos_wide=True), ref.samples), seed
        if ref.bits_per_sample <= 16:
            assert np.array_equal(host_decode(core, raw, ref, two_pass=True, narrow=True), ref.samples), seed
        done += 1
    assert done >= len(FUZZ_SEEDS) * 2 // 3


def test_host_core_random_streams_32bit(core):
    """32 bits per sample (int64 scratch, decode_predicted64): a full-scale signal has residuals of 33 bits — Rice codewords
    beyond 2^32, which the reference reads into its i64 (src/zflac.zig:614-666) whatever RFC 9639 allows an encoder."""
    from bitcraft import fuzz_case
    done = big = 0
    for seed in range(40000, 40120):
        case = fuzz_case(seed, (32,))
        if case is None:
            continue
        data, ref, _ = case
        assert np.array_equal(host_decode(core, data.tobytes(), ref), ref.samples), seed
        done += 1
        big += int(np.abs(ref.samples.astype(np.int64)).max() >= (1 << 31) - 1)
    assert done >= 60 and big >= 10
