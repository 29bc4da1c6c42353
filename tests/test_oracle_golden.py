"""Pins the CPU oracle (oracle/) to the reference's own tests — CPU only.

Mirrors /root/reference/tests/basic.zig (3 RFC 9639 known-answer vectors),
tests/std_subset.zig (whole-stream goldens), tests/std_uncommon.zig and
tests/std_faulty.zig (expected error names / MD5-gated success), from the
fixtures committed under tests/golden/ by tools/make_golden.py.
"""
import hashlib

import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries, corpus_ids, load_kats
from oracle import oracle


@pytest.mark.parametrize("kat", load_kats(), ids=lambda k: k["name"])
def test_rfc9639_known_answer(kat):
    r = oracle.decode(bytes.fromhex(kat["flac_hex"]))
    assert r.error == "OK"
    assert r.channels == kat["channels"]
    assert r.container_bytes == kat["container_bytes"]
    assert r.samples.tolist() == kat["samples"]


@pytest.mark.parametrize("entry", corpus_entries(), ids=corpus_ids())
def test_corpus(entry):
    r = oracle.decode(corpus_bytes(entry), trace_frames=True)
    if entry["expect_error"]:
        assert r.error == entry["expect_error"]
        assert r.samples is None
        return
    assert r.error == "OK"
    assert r.overflow_flag == 0
    assert (r.channels, r.sample_rate, r.bits_per_sample, r.container_bytes) == (
        entry["channels"], entry["sample_rate"], entry["bits_per_sample"], entry["container_bytes"])
    assert r.samples.size == entry["n_samples"]
    assert hashlib.sha256(r.samples.tobytes()).hexdigest() == entry["pcm_sha256"]
    assert r.md5_computed == r.md5_streaminfo == bytes.fromhex(entry["md5_streaminfo"])
    assert r.frames.size == entry["n_frames"]
    # the frame trace is a chain: each frame starts where the previous one ended
    assert int(r.frames["offset"][0]) == entry["audio_offset"]
    assert np.array_equal(r.frames["offset"][1:], r.frames["end_offset"][:-1])


def test_corpus_crcs_are_valid():
    """The reference reads CRC-8/CRC-16 and ignores them (zflac.zig:407-410, :548-551); on the
    corpus they are all valid, which is what lets the GPU scan use them as accelerators."""
    entry = next(e for e in corpus_entries() if e["group"] == "subset" and e["id"] == "01")
    data = corpus_bytes(entry)
    r = oracle.decode(data, trace_frames=True)
    for f in r.frames:
        o, e, h = int(f["offset"]), int(f["end_offset"]), int(f["header_len"])
        assert oracle.crc8(data[o:o + h - 1]) == data[o + h - 1]
        assert oracle.crc16(data[o:e - 2]) == int.from_bytes(data[e - 2:e], "big")


def test_md5_known_answers():
    assert oracle.md5(b"").hex() == "d41d8cd98f00b204e9800998ecf8427e"
    assert oracle.md5(b"abc").hex() == "900150983cd24fb0d6963f7d28e17f72"
    msg = bytes(range(256)) * 5
    assert oracle.md5(msg) == hashlib.md5(msg).digest()


def test_truncated_stream_errors():
    entry = next(e for e in corpus_entries() if e["group"] == "subset" and e["id"] == "60")
    data = corpus_bytes(entry)
    assert oracle.decode(data[:3]).error == "EndOfStream"
    assert oracle.decode(b"fLaX" + data[4:]).error == "InvalidSignature"
    # total sample count is known, so running out of bytes mid-stream is EndOfStream (zflac.zig:344)
    assert oracle.decode(data[: len(data) // 2]).error == "EndOfStream"


@pytest.mark.parametrize("entry", corpus_entries(), ids=corpus_ids())
def test_timed_build_is_the_same_decoder(entry):
    """The -DZO_TIMED build (what bench.py times as the CPU arm: no overflow bookkeeping, LPC loop instantiated per order
    like the reference's comptime unroll) against the same golden vectors and the instrumented build."""
    data = corpus_bytes(entry)
    a, b = oracle.decode(data), oracle.decode(data, timed=True)
    assert a.error == b.error == (entry["expect_error"] or "OK")
    if a.error == "OK":
        assert hashlib.sha256(b.samples.tobytes()).hexdigest() == entry["pcm_sha256"]
        assert np.array_equal(a.samples, b.samples) and a.md5_computed == b.md5_computed
        assert (a.channels, a.sample_rate, a.bits_per_sample) == (b.channels, b.sample_rate, b.bits_per_sample)


def test_timed_build_on_the_bench_workloads():
    from tools import flacgen as fg
    for p in (fg.c2(4.0), fg.c3(0.5), fg.c5_file(3)):
        s = fg.generate(p, want_pcm=True)
        r = oracle.decode(s.flac, timed=True)
        assert r.error == "OK" and np.array_equal(r.samples, s.pcm)
