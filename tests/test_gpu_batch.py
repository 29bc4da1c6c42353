"""GPU parity of the file-parallel batch path (BASELINE config 5): every stream of a mixed batch must come back
exactly as a one-by-one decode by the oracle would return it — samples, header fields, MD5, or the same error."""
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


def check_against_oracle(z, streams, results):
    from oracle import oracle
    assert len(results) == len(streams)
    for data, got in zip(streams, results):
        ref = oracle.decode(data)
        if ref.error != "OK":
            assert isinstance(got, z.ZflacError) and type(got).__name__ == ref.error
            continue
        assert not isinstance(got, Exception), got
        assert (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
        assert got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples)
        assert got.info["md5_computed"] == ref.md5_computed


def test_batch_of_short_files(z):
    streams = [fg.generate(fg.c5_file(i), want_pcm=False).flac for i in range(96)]
    res = z.decode_batch(streams)
    check_against_oracle(z, streams, res)
    assert all(r.info["fast_path"] == 1 for r in res)


def test_batch_large_enough_for_lane_md5(z):
    """>= 256 streams: the MD5 gate runs one chain per LANE (k_md5_lanes); odd lengths exercise its padding blocks."""
    ps = []
    for i in range(300):
        p = fg.c5_file(i)
        p.n_samples = 4096 + 257 * i  # a frame or a few, every length different
        ps.append(p)
    streams = [fg.generate(p, want_pcm=False).flac for p in ps]
    res = z.decode_batch(streams)
    check_against_oracle(z, streams, res)
    bad = [s.copy() for s in streams]
    bad[7][bad[7].size - 9] ^= 0x01  # flips a residual bit: only the MD5 gate can notice
    res = z.decode_batch(bad)
    check_against_oracle(z, bad, res)


def test_mixed_batch_containers_errors_and_fallbacks(z):
    names = ["c5_file_0", "bps8_fixed_lowprec", "bps24_stereo_leftside", "unknown_total", "c4_32bit_rice2_escapes",
             "bps12_midside", "ch6_mixed", "c4_variable_blocksize", "c5_file_1", "bps20_mono_rice2", "tiny_blocks_16"]
    streams = [make(n).flac for n in names]
    streams.append(streams[0][: streams[0].size // 2].copy())     # truncated
    streams.append(np.frombuffer(b"fLaX" + bytes(100), dtype=np.uint8))  # bad signature
    damaged = streams[8].copy()
    damaged[damaged.size - 50] ^= 0x40                              # payload damage in the last frame
    streams.append(damaged)
    res = z.decode_batch(streams)
    check_against_oracle(z, streams, res)
    # same results when every stream is walked on the host
    res2 = z.decode_batch(streams, use_fast_path=False)
    check_against_oracle(z, streams, res2)


def test_batch_of_corpus_files(z):
    entries = [e for e in corpus_entries()][:40]
    streams = [np.frombuffer(corpus_bytes(e), dtype=np.uint8) for e in entries]
    res = z.decode_batch(streams)
    check_against_oracle(z, streams, res)


def test_batch_device_resident_and_read_all(z):
    import torch
    streams = [fg.generate(fg.c5_file(100 + i), want_pcm=True) for i in range(32)]
    blob, offs, lens = z.pack_streams([s.flac for s in streams])
    d_blob = torch.from_numpy(blob).cuda()
    b = z.open_batch(blob, offs, lens, device_blob_ptr=d_blob.data_ptr())
    assert b.statuses == [0] * 32
    buf, po = b.read_all()
    for s, v in zip(streams, b.streams_from(buf, po)):
        assert np.array_equal(v, s.pcm)
    assert np.array_equal(b.stream(5).samples, streams[5].pcm)
    b.close()


def test_batch_decode_one_call_pipelined(z):
    """zflacb200_batch_decode (sub-batches overlapped over both PCIe directions) == the oracle, stream by stream; a stream
    the device chain refuses and a bad signature ride along; a destination that is too small says OutOfMemory."""
    from oracle import oracle
    streams = [fg.generate(fg.c5_file(200 + i), want_pcm=False).flac for i in range(80)]     # ~9.6 MB: one sub-batch
    streams.insert(17, make("unknown_total").flac)
    streams.insert(40, np.frombuffer(b"fLaX" + bytes(60), dtype=np.uint8))
    blob, offs, lens = z.pack_streams(streams)
    blob = np.concatenate([blob, np.zeros(24 << 20, dtype=np.uint8)])  # trailing bytes: >= 3 sub-batches by size
    bound = z.batch_pcm_bound(blob, offs, lens)
    dst = np.zeros(bound + (4 << 20), dtype=np.uint8)
    res = z.decode_batch_into(blob, offs, lens, dst)
    for i, data in enumerate(streams):
        ref = oracle.decode(data)
        if ref.error != "OK":
            assert res.statuses[i] != 0 and res.samples(i) is None
            continue
        assert res.statuses[i] == 0, (i, res.statuses[i])
        assert np.array_equal(res.samples(i), ref.samples), i
        assert res.infos[i]["md5_computed"] == ref.md5_computed
    small = np.zeros(1 << 20, dtype=np.uint8)
    res = z.decode_batch_into(blob, offs, lens, small)
    assert 12 in res.statuses  # ZFLACB200_OUT_OF_MEMORY for the streams that no longer fitted
    ok = [i for i in range(len(streams)) if res.statuses[i] == 0]
    assert ok and all(np.array_equal(res.samples(i), oracle.decode(streams[i]).samples) for i in ok[:4])


def test_batch_decode_in_eight_sub_batches(z, monkeypatch):
    """The one-call batch really cut into sub-batches (ZFLACB200_BATCH_PART_MIN_BYTES lowers the size below which a batch
    stays whole): 8 parts of ~12 files, every file == the generator's PCM, offsets in blob order, MD5 gate on."""
    files = fg.generate_batch([fg.c5_file(300 + i) for i in range(96)], want_pcm=True)
    blob, offs, lens = z.pack_streams([f.flac for f in files])
    dst = np.zeros(z.batch_pcm_bound(blob, offs, lens) + 4096, dtype=np.uint8)
    monkeypatch.setenv("ZFLACB200_BATCH_PART_MIN_BYTES", str(1 << 20))
    res = z.decode_batch_into(blob, offs, lens, dst)
    monkeypatch.delenv("ZFLACB200_BATCH_PART_MIN_BYTES")
    whole = z.decode_batch_into(blob, offs, lens, np.zeros_like(dst))
    assert res.statuses == [0] * len(files) == whole.statuses
    for i, f in enumerate(files):
        assert np.array_equal(res.samples(i), f.pcm), i
        assert res.infos[i]["md5_checked"] == 1 and res.infos[i]["md5_computed"] == f.md5
        assert np.array_equal(whole.samples(i), f.pcm), i
