"""GPU parity for the additive raw-frame / resync entry point zflacb200_open_frames (the reference's
Readme.md:54 "sync to next frame" TODO; flac-test-files uncommon/10-11, which zflac.decode itself rejects,
tests/std_uncommon.zig:53-59). The checker is the oracle run on the same frames behind a synthetic STREAMINFO that
states no total (tests/frames_util.py)."""
import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries
from frames_util import first_frame_params, wrap_frames

pytestmark = pytest.mark.gpu

OK_ENTRIES = [e for e in corpus_entries() if not e["expect_error"]]
OK_IDS = [f"{e['group']}_{e['id']}" for e in OK_ENTRIES]


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


def oracle_on_frames(frames: bytes):
    from oracle import oracle
    ch, depth, rate = first_frame_params(frames)
    if depth == 0:
        return None
    return oracle.decode(wrap_frames(frames, ch, depth, rate), trace_frames=True, keep_on_md5_mismatch=True)


def check_against(z, data: bytes, byte_offset: int, frame_start: int, links: bool = True):
    """decode_frames(data, byte_offset) must begin at frame_start and give what the oracle gives for data[frame_start:]."""
    ref = oracle_on_frames(data[frame_start:])
    if ref is None:  # depth code 0 = "see STREAMINFO": a raw frame stream cannot be decoded
        with pytest.raises(z.MissingStreaminfo):
            z.decode_frames(data, byte_offset, use_crc_links=links)
        return None
    if ref.error not in ("OK", "InvalidChecksum"):  # the wrapped stream states MD5 = 0: InvalidChecksum = every frame decoded
        with pytest.raises(z.ZflacError) as ei:
            z.decode_frames(data, byte_offset, use_crc_links=links)
        assert type(ei.value).__name__ == ref.error
        return None
    got = z.decode_frames(data, byte_offset, use_crc_links=links)
    assert got.info["audio_offset"] == frame_start
    assert got.info["md5_checked"] == 0 and got.info["fast_path"] == 0
    assert (got.channels, got.bits_per_sample, got.sample_rate) == (ref.channels, ref.bits_per_sample, ref.sample_rate)
    assert got.samples.dtype == ref.samples.dtype
    assert np.array_equal(got.samples, ref.samples)
    assert got.info["n_frames"] == len(ref.frames)
    assert got.info["n_crc16_mismatch"] == 0
    return got


@pytest.mark.parametrize("entry", OK_ENTRIES[::3], ids=OK_IDS[::3])
def test_stripped_corpus_file_decodes_like_the_whole_file(z, entry):
    """Signature and metadata cut away: same PCM as the reference's decode of the whole file (when the frames carry their
    own bit depth), and the same as the oracle on the re-wrapped frames in every case."""
    from oracle import oracle
    data = corpus_bytes(entry)
    whole = oracle.decode(data, trace_frames=True)
    raw = data[whole.audio_offset:]
    got = check_against(z, raw, 0, 0)
    if got is not None and got.bits_per_sample == entry["bits_per_sample"]:
        assert np.array_equal(got.samples, whole.samples)


@pytest.mark.parametrize("entry", OK_ENTRIES[1::6], ids=OK_IDS[1::6])
def test_resync_from_the_middle_of_a_frame(z, entry):
    """byte_offset inside frame k-1 of a complete file: decoding begins at frame k and yields the tail of the PCM."""
    from oracle import oracle
    data = corpus_bytes(entry)
    whole = oracle.decode(data, trace_frames=True)
    fr = whole.frames
    if len(fr) < 3:
        pytest.skip("too few frames")
    k = len(fr) // 2
    inside = int(fr["offset"][k - 1]) + 1
    for links in (True, False):
        got = check_against(z, data, inside, int(fr["offset"][k]), links)
        if got is not None and got.bits_per_sample == entry["bits_per_sample"]:
            assert np.array_equal(got.samples, whole.samples[int(fr["sample_offset"][k]):])
    # exactly at a frame start as well
    check_against(z, data, int(fr["offset"][k]), int(fr["offset"][k]))


def test_uncommon_10_file_starting_at_frame_header(z):
    entry = next(e for e in corpus_entries() if e["group"] == "uncommon" and e["id"] == "10")
    data = corpus_bytes(entry)
    with pytest.raises(z.InvalidSignature):  # the reference's own verdict (tests/std_uncommon.zig:53-55)
        z.decode(data)
    got = check_against(z, data, 0, 0)
    assert got.samples.size == 664669 and got.info["n_frames"] == 163


def test_uncommon_11_file_starting_with_unparsable_data(z):
    entry = next(e for e in corpus_entries() if e["group"] == "uncommon" and e["id"] == "11")
    data = corpus_bytes(entry)
    with pytest.raises(z.InvalidSignature):  # tests/std_uncommon.zig:57-59
        z.decode(data)
    got = z.decode_frames(data)
    start = got.info["audio_offset"]
    assert start == 895 == data.find(b"\xff\xf8")  # nothing that looks like a frame precedes it in this file
    got = check_against(z, data, 0, start)
    assert got.samples.size == 568394


def test_no_frame_at_all_and_bad_offset(z):
    with pytest.raises(z.EndOfStream):
        z.decode_frames(bytes(range(200)) * 40)
    with pytest.raises(z.ZflacError):
        z.decode_frames(b"\x00" * 16, 17)
