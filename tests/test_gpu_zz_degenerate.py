"""GPU parity on degenerate inputs of decode(): empty, signature only, metadata only, cuts at frame boundaries.
Last GPU file in collection order on purpose (first captured r03b)."""
import numpy as np
import pytest

from synth_cases import make

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


def _same_as_oracle(z, data):
    from oracle import oracle
    ref = oracle.decode(data)
    if ref.error == "OK":
        got = z.decode(data)
        assert got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples)
        assert (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
    else:
        with pytest.raises(z.ZflacError) as ei:
            z.decode(data)
        assert type(ei.value).__name__ == ref.error, (data.size, ref.error)
    return ref.error


def test_empty_and_metadata_only_streams(z):
    """Degenerate inputs of decode() (src/zflac.zig:217-310): no bytes at all, the signature alone, a stream that ends right
    after its metadata (stated total: the frame loop's first read fails; total 0: the loop ends at EOF and the MD5 gate
    answers), and a cut at / just after a frame boundary of a stream without a stated total. Same answer as the oracle,
    alone and as members of one batch."""
    from oracle import oracle
    cases = [np.zeros(0, dtype=np.uint8), np.frombuffer(b"fLaC", dtype=np.uint8).copy()]
    for name in ("c5_file_0", "unknown_total"):
        s = make(name)
        r = oracle.decode(s.flac, trace_frames=True)
        a = int(r.frames["offset"][0])
        e2 = int(r.frames["end_offset"][2])
        cases += [s.flac[:c].copy() for c in (a, a + 1, a + 4, e2, e2 + 1, e2 + 4)]
    errors = {_same_as_oracle(z, d) for d in cases}
    assert "EndOfStream" in errors
    from test_gpu_batch import check_against_oracle
    check_against_oracle(z, cases, z.decode_batch(cases))
    # an empty slice may arrive as (NULL, 0): still the reference's EndOfStream, not an argument error
    import ctypes as C
    from zflac_b200 import _native
    lib = _native.load()
    handle, info = C.c_void_p(), _native.Info()
    st = lib.zflacb200_open(None, 0, None, C.byref(handle), C.byref(info))
    assert lib.zflacb200_status_name(st).decode() == "EndOfStream" and not handle.value


def _overstate_total(flac: np.ndarray, total: int) -> np.ndarray:
    """STREAMINFO total (36 bits at bit 108 of the block body) replaced: a claim the bytes cannot back up."""
    out = flac.copy()
    body = 8  # signature + block header
    v = int.from_bytes(out[body + 13: body + 18].tobytes(), "big")
    v = (v & ~((1 << 36) - 1)) | (total & ((1 << 36) - 1))
    out[body + 13: body + 18] = np.frombuffer(v.to_bytes(5, "big"), dtype=np.uint8)
    return out


@pytest.mark.parametrize("total", [(1 << 36) - 1, 50_000_000], ids=["2^36-1", "5e7"])
def test_overstated_total_solo(z, total):
    """the fast path must not size its buffers from a total the bytes cannot hold; the stream goes the robust
    way and ends like the reference's loop does (EndOfStream when the bytes run out before the stated total)."""
    from oracle import oracle
    from tools import flacgen as fg
    s = fg.generate(fg.c2(2.0), want_pcm=False)
    bad = _overstate_total(s.flac, total)
    # the reference allocates total * channels samples up front (:329-334): OutOfMemory where the allocator refuses
    # (the oracle's malloc here: 275 GB), else it decodes until the bytes run out: EndOfStream (:343-350). Never a CUDA error.
    assert oracle.decode(bad).error in ("EndOfStream", "OutOfMemory")
    with pytest.raises((z.EndOfStream, z.OutOfMemory)):
        z.decode(bad)
    with pytest.raises((z.EndOfStream, z.OutOfMemory)):
        z.decode_into(bad, np.empty(1 << 20, dtype=np.int16))
    assert np.array_equal(z.decode(s.flac).samples, oracle.decode(s.flac).samples)  # the library is fine afterwards


def test_overstated_total_in_a_batch(z):
    """One member with an absurd total must not take the batch (or its launch group) down with it."""
    from tools import flacgen as fg
    files = [fg.generate(fg.c5_file(i), want_pcm=True) for i in range(6)]
    streams = [f.flac for f in files]
    streams[2] = _overstate_total(streams[2], (1 << 36) - 1)
    streams[4] = _overstate_total(streams[4], 3_000_000_000)
    res = z.decode_batch(streams)
    for i, (f, r) in enumerate(zip(files, res)):
        if i in (2, 4):
            assert isinstance(r, (z.EndOfStream, z.OutOfMemory)), r
        else:
            assert np.array_equal(r.samples, f.pcm)


def test_samples_wider_than_the_frame_depth(z):
    """A malformed but defined 24-bit stream: samples of 30 bits, so that the LPC sum leaves int32 where the frame's depth
    says it cannot (bitcraft.wide_sample_stream). The reference sums in i64 in a 32-bit container (src/zflac.zig:314-319);
    pass B and the single-pass decoder must too — every path against the oracle, alone and in a batch."""
    from oracle import oracle
    from bitcraft import wide_sample_stream
    data = wide_sample_stream()
    ref = oracle.decode(data)
    assert ref.error == "OK" and ref.container_bytes == 4
    assert _same_as_oracle(z, data) == "OK"
    for kw in (dict(use_fast_path=False), dict(verify_md5=False), dict(word_pipeline=True), dict(md5_on_device=True),
               dict(single_pass=True), dict(single_pass=True, use_fast_path=False)):
        got = z.decode(data, **kw)
        assert np.array_equal(got.samples, ref.samples), kw
    from test_gpu_batch import check_against_oracle
    cases = [data, make("c5_file_0").flac, data]
    check_against_oracle(z, cases, z.decode_batch(cases))


def test_random_streams(z):
    """tests/bitcraft.py's random encoder (any legal field combination, PCM known by construction): every stream the
    reference is defined on — one batch (mixed depths and channel counts: several launch groups), then each path alone on
    a sample of them. Where a quirk of the reference changes the PCM (fuzz_case: not exact) the gate must say
    InvalidChecksum, and without the gate the samples are the oracle's."""
    from bitcraft import fuzz_case
    from test_gpu_batch import check_against_oracle
    cases = [c for c in (fuzz_case(seed) for seed in range(1000, 1400)) if c is not None]
    assert len(cases) >= 250
    streams = [c[0] for c in cases]
    check_against_oracle(z, streams, z.decode_batch(streams))
    paths = [dict(), dict(use_fast_path=False), dict(word_pipeline=True), dict(single_pass=True), dict(md5_on_device=True)]
    for k, (data, ref, exact) in enumerate(cases[:150]):
        kw = dict(paths[k % 5])
        if not exact:
            with pytest.raises(z.ZflacError) as ei:
                z.decode(data, **kw)
            assert type(ei.value).__name__ == "InvalidChecksum"
            kw["verify_md5"] = False
        got = z.decode(data, **kw)
        assert got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples), (k, kw)
        assert (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)


def test_random_streams_32bit(z):
    """The same at 32 bits per sample (int64 scratch, the single-pass decoder): full-scale signals, whose residuals need 33
    bits and whose Rice codewords pass 2^32 (the reference reads them into an i64, src/zflac.zig:614-666)."""
    from bitcraft import fuzz_case
    from test_gpu_batch import check_against_oracle
    cases = [c for c in (fuzz_case(seed, (32,)) for seed in range(40000, 40120)) if c is not None]
    assert len(cases) >= 60
    streams = [c[0] for c in cases]
    check_against_oracle(z, streams, z.decode_batch(streams))
    for k, (data, ref, exact) in enumerate(cases):
        kw = [dict(), dict(use_fast_path=False), dict(md5_on_device=True)][k % 3]
        if not exact:
            kw["verify_md5"] = False
        got = z.decode(data, **kw)
        assert got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples), (k, kw)
