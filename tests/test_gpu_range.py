"""GPU parity of the frame-range decode (one long stream over several GPUs). The shards are decoded one
after the other on the one GPU of the test box — nothing links them on the device, so that is exactly what N ranks do —
and the pieces, laid end to end in the order the stitch accepts, must be the oracle's whole-stream samples bit for bit.
Streams the hints cannot place (no stated total, damage) must say so (accepted = 0 / stitch fails): the caller then
decodes the stream whole, which is the already-tested path."""
import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries
from synth_cases import make
from tools import flacgen as fg

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


def sharded(z, data, world):
    from zflac_b200 import sharding
    parts = [z.decode_range(data, lo, hi) for lo, hi in sharding.byte_ranges(len(data), world)]
    audio_offset = parts[0].info["audio_offset"]
    return parts, z.stitch_ranges(parts, audio_offset)


def check_stitched(z, data, world, must_stitch=True):
    from oracle import oracle
    ref = oracle.decode(data)
    assert ref.error == "OK"
    parts, ok = sharded(z, data, world)
    if not ok:
        assert not must_stitch, [p.as_row() for p in parts]
        return False
    pcm = np.concatenate([p.samples for p in parts if p.n_frames])
    assert pcm.dtype == ref.samples.dtype and np.array_equal(pcm, ref.samples)
    C = ref.channels
    assert sum(p.n_samples for p in parts) * C == ref.samples.size
    for p in parts:  # header fields every shard reports are the first frame's (zflac.zig:376-388)
        assert (p.info["channels"], p.info["sample_rate"], p.info["bits_per_sample"]) == (C, ref.sample_rate, ref.bits_per_sample)
        if p.n_frames:
            assert np.array_equal(p.samples, ref.samples[p.first_sample * C:(p.first_sample + p.n_samples) * C])
    return True


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_long_stereo_stream_in_shards(z, world):
    s = fg.generate(fg.c2(20.0), want_pcm=False)  # 216 frames of 4096, mid/side, LPC 8
    assert check_stitched(z, s.flac, world)


@pytest.mark.parametrize("world", [2, 5])
def test_hires_8ch_stream_in_shards(z, world):
    s = fg.generate(fg.c3(1.5), want_pcm=False)   # 24-bit (MSB-align shift), 8 channels, order 32, big frames
    assert check_stitched(z, s.flac, world)


def test_more_shards_than_frames(z):
    s = fg.generate(fg.c2(0.5), want_pcm=False)   # 6 frames over 16 ranks: most shards hold no frame start
    parts, ok = sharded(z, s.flac, 16)
    assert ok and any(p.n_frames == 0 for p in parts)
    assert check_stitched(z, s.flac, 16)


def test_synthetic_edge_streams_in_shards(z):
    stitched = {}
    for name in ("c4_variable_blocksize", "c4_32bit_rice2_escapes", "c4_wasted_bits_mixed", "ch6_mixed", "bps12_midside", "unknown_total"):
        data = make(name).flac
        stitched[name] = check_stitched(z, data, 3, must_stitch=False)
    assert not stitched["unknown_total"]              # hints need the stated total: the whole-stream path takes it
    assert stitched["ch6_mixed"] and stitched["bps12_midside"], stitched


def test_corpus_in_shards(z):
    n_ok = 0
    for e in corpus_entries():
        if e.get("expect_error"):
            continue
        data = corpus_bytes(e)
        n_ok += bool(check_stitched(z, data, 4, must_stitch=False))
    assert n_ok >= 20, n_ok  # the subset files with a fixed block size and a stated total all stitch


def test_damage_and_errors(z):
    s = fg.generate(fg.c2(10.0), want_pcm=False)
    bad = s.flac.copy()
    bad[bad.size * 3 // 4] ^= 0x10                    # payload damage in the middle of shard 1 of 2
    parts, ok = sharded(z, bad, 2)
    # residual damage still decodes to (different) samples unless it breaks the frame's bit chain; either way the
    # untouched shard is exact and a broken chain is reported, never papered over
    from oracle import oracle
    ref = oracle.decode(s.flac)
    assert parts[0].accepted and np.array_equal(parts[0].samples, ref.samples[:parts[0].samples.size])
    hdr = s.flac.copy()
    first = parts[1].first_frame_offset
    hdr[first + 1] ^= 0x04                            # sync code of shard 1's first frame: the frame vanishes
    parts, ok = sharded(z, hdr, 2)
    assert not ok
    with pytest.raises(z.ZflacError) as ei:
        z.decode_range(np.frombuffer(b"fLaX" + bytes(100), dtype=np.uint8), 0, 104)
    assert type(ei.value).__name__ == "InvalidSignature"
    small = np.zeros(16, dtype=np.uint8)
    with pytest.raises(z.ZflacError) as ei:
        z.decode_range(s.flac, 0, len(s.flac), small)
    assert type(ei.value).__name__ == "OutOfMemory"
