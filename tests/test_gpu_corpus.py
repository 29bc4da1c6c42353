"""GPU parity tests proper: the CUDA path through the C ABI vs the oracle and the committed goldens.

Mirrors the reference's tests/basic.zig, std_subset.zig, std_uncommon.zig, std_faulty.zig."""
import hashlib

import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries, corpus_ids, load_kats

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


@pytest.mark.parametrize("kat", load_kats(), ids=lambda k: k["name"])
def test_rfc9639_known_answer(z, kat):
    r = z.decode(bytes.fromhex(kat["flac_hex"]))
    assert r.channels == kat["channels"]
    assert r.samples.dtype.itemsize == kat["container_bytes"]
    assert r.samples.tolist() == kat["samples"]


@pytest.mark.parametrize("entry", corpus_entries(), ids=corpus_ids())
def test_corpus_vs_golden_and_oracle(z, entry):
    from oracle import oracle
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    if entry["expect_error"]:
        with pytest.raises(z.ZflacError) as ei:
            z.decode(data)
        assert type(ei.value).__name__ == entry["expect_error"] == ref.error
        return
    got = z.decode(data)
    assert (got.channels, got.sample_rate, got.bits_per_sample, got.samples.dtype.itemsize) == (
        entry["channels"], entry["sample_rate"], entry["bits_per_sample"], entry["container_bytes"])
    assert got.samples.size == entry["n_samples"]
    assert hashlib.sha256(got.samples.tobytes()).hexdigest() == entry["pcm_sha256"]  # golden (reference's raw / MD5)
    assert np.array_equal(got.samples, ref.samples)                                    # oracle, sample for sample
    assert got.info["md5_computed"] == bytes.fromhex(entry["md5_streaminfo"]) == ref.md5_computed
    assert got.info["n_frames"] == entry["n_frames"]


@pytest.mark.parametrize("entry", [e for e in corpus_entries() if not e["expect_error"]][::4],
                         ids=[f"{e['group']}_{e['id']}" for e in corpus_entries() if not e["expect_error"]][::4])
def test_frame_index_matches_oracle_chain(z, entry):
    """K1 + K1b + chain: same frames, offsets, ends and sample offsets as the oracle's sequential walk;
    and the same result when every frame end is probed by decoding instead of linked by CRC-16."""
    from oracle import oracle
    data = corpus_bytes(entry)
    ref = oracle.decode(data, trace_frames=True)
    for fast, links in ((True, True), (False, True), (False, False)):
        s = z.open_device(0, 0, host_data=data, use_crc_links=links, use_fast_path=fast)
        fr = s.frames()
        assert np.array_equal(fr["offset"], ref.frames["offset"])
        assert np.array_equal(fr["end_offset"], ref.frames["end_offset"])
        assert np.array_equal(fr["sample_offset"], ref.frames["sample_offset"])
        assert np.array_equal(fr["block_size"], ref.frames["block_size"])
        assert np.array_equal(s.to_host(), ref.samples)
        assert s.info["n_crc16_mismatch"] == 0
        if not fast:
            assert s.info["fast_path"] == 0
        if not links:
            assert s.info["n_probed_frames"] == fr.size
        s.close()


def test_fast_path_covers_the_regular_corpus(z):
    """The device-side hinted chain (no host walk) must be the path that actually runs on ordinary streams:
    every error-free corpus file with a stated total and consecutive coded numbers."""
    taken = {}
    for entry in corpus_entries():
        if entry["expect_error"]:
            continue
        s = z.open_device(0, 0, host_data=corpus_bytes(entry))
        taken[f"{entry['group']}_{entry['id']}"] = s.info["fast_path"]
        s.close()
    assert taken["subset_01"] == 1 and taken["subset_28"] == 1
    assert sum(taken.values()) >= len(taken) * 3 // 4, taken
