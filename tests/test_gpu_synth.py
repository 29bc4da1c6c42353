"""GPU parity on synthetic streams (BASELINE configs 2-5 shapes + edge cases): CUDA path through the C ABI ==
oracle == the PCM the generator encoded, bit for bit; MD5 == STREAMINFO."""
import numpy as np
import pytest

from synth_cases import CASES, make

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


@pytest.mark.parametrize("name", sorted(CASES))
def test_cuda_matches_oracle_and_generator(z, name):
    from oracle import oracle
    s = make(name)
    ref = oracle.decode(s.flac, trace_frames=True)
    got = z.decode(s.flac)
    assert (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
    assert got.samples.dtype == ref.samples.dtype
    assert np.array_equal(got.samples, s.pcm)
    assert np.array_equal(got.samples, ref.samples)
    assert got.info["md5_computed"] == s.md5 == ref.md5_computed
    assert got.info["n_frames"] == s.n_frames
    if name != "unknown_total":  # regular streams must take the device-side chain (fast path)
        assert got.info["fast_path"] == 1, name
    # same answer when the host walks the chain (CRC-16 links), and when every frame end is found by decoding
    got1 = z.decode(s.flac, use_fast_path=False)
    assert got1.info["fast_path"] == 0 and np.array_equal(got1.samples, s.pcm)
    got2 = z.decode(s.flac, use_crc_links=False)
    assert np.array_equal(got2.samples, s.pcm)


def test_crc_damage_is_ignored_like_the_reference(z):
    """Q1: a wrong CRC-16 must not change the result (the frame is then probed, not linked); a payload bit flip
    must surface as the same error the oracle reports."""
    from oracle import oracle
    s = make("c5_file_0")
    r0 = oracle.decode(s.flac, trace_frames=True)
    end = int(r0.frames["end_offset"][3])
    bad_crc = s.flac.copy()
    bad_crc[end - 1] ^= 0x5A
    got = z.decode(bad_crc)
    assert np.array_equal(got.samples, s.pcm)
    assert got.info["n_crc16_mismatch"] == 1
    got = z.decode(bad_crc, use_fast_path=False)
    assert np.array_equal(got.samples, s.pcm)
    assert got.info["n_crc16_mismatch"] == 1 and got.info["n_probed_frames"] >= 1
    # ZFLACB200_SKIP_FRAME_CRC16: the diagnostic is not computed, nothing else changes
    h = z.open_device(0, 0, host_data=bad_crc, frame_crc16=False)
    assert np.array_equal(h.to_host(), s.pcm) and h.info["n_crc16_mismatch"] == 0 and h.info["fast_path"] == 1
    assert set(h.frames()["crc16_ok"].tolist()) == {2}
    h.close()
    dst = np.zeros(s.pcm.size, dtype=s.pcm.dtype)
    assert z.decode_into(bad_crc, dst, frame_crc16=False)["n_crc16_mismatch"] == 0 and np.array_equal(dst, s.pcm)
    hdr = int(r0.frames["offset"][5]) + int(r0.frames["header_len"][5]) - 1  # CRC-8 byte of frame 5
    bad_crc8 = s.flac.copy()
    bad_crc8[hdr] ^= 0x33
    assert np.array_equal(z.decode(bad_crc8).samples, s.pcm)
    bad_payload = s.flac.copy()
    bad_payload[end - 40] ^= 0x10
    want = oracle.decode(bad_payload).error
    with pytest.raises(z.ZflacError) as ei:
        z.decode(bad_payload)
    assert type(ei.value).__name__ == want


def test_truncated_streams_match_oracle_errors(z):
    from oracle import oracle
    s = make("c5_file_1")
    for cut in (3, 30, 100, s.flac.size // 2, s.flac.size - 1):
        data = s.flac[:cut].copy()
        want = oracle.decode(data).error
        if want == "OK":
            assert np.array_equal(z.decode(data).samples, oracle.decode(data).samples)
        else:
            with pytest.raises(z.ZflacError) as ei:
                z.decode(data)
            assert type(ei.value).__name__ == want, cut

