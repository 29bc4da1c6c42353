"""CPU: the re-wrapping helper the raw-frame GPU tests lean on gives the oracle the same frames the whole file holds."""
import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries
from frames_util import first_frame_params, wrap_frames


@pytest.mark.parametrize("fid", ["01", "14", "20"])
def test_rewrapped_frames_decode_like_the_whole_file(fid):
    from oracle import oracle
    entry = next(e for e in corpus_entries() if e["group"] == "subset" and e["id"] == fid)
    data = corpus_bytes(entry)
    whole = oracle.decode(data, trace_frames=True)
    raw = data[whole.audio_offset:]
    ch, depth, rate = first_frame_params(raw)
    assert (ch, depth) == (entry["channels"], entry["bits_per_sample"])
    again = oracle.decode(wrap_frames(raw, ch, depth, rate), trace_frames=True, keep_on_md5_mismatch=True)
    assert again.error == "InvalidChecksum"  # the synthetic STREAMINFO states MD5 = 0; every frame decoded
    assert np.array_equal(again.samples, whole.samples)
    assert np.array_equal(again.frames["block_size"], whole.frames["block_size"])


def test_uncommon_10_is_a_raw_frame_stream():
    from oracle import oracle
    entry = next(e for e in corpus_entries() if e["group"] == "uncommon" and e["id"] == "10")
    data = corpus_bytes(entry)
    assert oracle.decode(data).error == "InvalidSignature"
    r = oracle.decode(wrap_frames(data, *first_frame_params(data)), trace_frames=True, keep_on_md5_mismatch=True)
    assert r.error == "InvalidChecksum" and r.samples.size == 664669 and len(r.frames) == 163
