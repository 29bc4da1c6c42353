"""GPU parity of the MD5 gate (reference src/zflac.zig:267-280) in both of its homes: the host thread that hashes one
long stream's PCM as it lands in host memory (zflac_b200/csrc/hostmd5.h) and the device chain (k_md5). Both must yield
the oracle's digest for every packing rule (1, 2, 3, 4 bytes per sample, MSB-aligned depths), and a stream whose digest
is wrong must fail with InvalidChecksum either way."""
import os

import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries
from oracle import oracle

pytestmark = pytest.mark.gpu

# one file per packing rule: 8 bit, 12 bit (MSB shift in s16), 16 bit, 20 bit and 24 bit (3 bytes of an s32), 32 bit, 15 bit
WANTED = {("subset", "23"), ("subset", "22"), ("subset", "01"), ("subset", "37"), ("subset", "28"), ("subset", "43"),
          ("uncommon", "05"), ("uncommon", "07"), ("faulty", "03")}


def _entries():
    return [e for e in corpus_entries() if (e["group"], e["id"]) in WANTED]


@pytest.fixture(params=["host", "device"])
def md5_home(request):
    old = os.environ.get("ZFLACB200_HOST_MD5_MIN_BYTES")
    os.environ["ZFLACB200_HOST_MD5_MIN_BYTES"] = "0"  # every stream is "long"
    yield request.param == "device"
    if old is None:
        del os.environ["ZFLACB200_HOST_MD5_MIN_BYTES"]
    else:
        os.environ["ZFLACB200_HOST_MD5_MIN_BYTES"] = old


@pytest.mark.parametrize("entry", _entries(), ids=lambda e: f"{e['group']}_{e['id']}")
@pytest.mark.parametrize("fast", [True, False], ids=["fast", "robust"])
def test_md5_gate_matches_oracle(entry, fast, md5_home):
    import zflac_b200 as z
    data = corpus_bytes(entry)
    ref = oracle.decode(data)
    try:
        got = z.decode(data, use_fast_path=fast, md5_on_device=md5_home)
    except z.ZflacError as e:
        assert type(e).__name__ == ref.error
        return
    assert ref.error == "OK"
    assert got.info["md5_checked"] == 1
    assert got.info["md5_computed"] == ref.md5_computed == got.info["md5_streaminfo"]
    assert np.array_equal(got.samples, ref.samples)


@pytest.mark.parametrize("fast", [True, False], ids=["fast", "robust"])
def test_wrong_digest_is_invalid_checksum(fast, md5_home):
    import zflac_b200 as z
    e = next(x for x in corpus_entries() if (x["group"], x["id"]) == ("subset", "01"))
    data = bytearray(corpus_bytes(e))
    data[8 + 18 + 5] ^= 0x40  # one bit of the STREAMINFO MD5 (signature 4 + block header 4 + 18 bytes of fields)
    with pytest.raises(z.InvalidChecksum):
        z.decode(bytes(data), use_fast_path=fast, md5_on_device=md5_home)
    assert oracle.decode(bytes(data)).error == "InvalidChecksum"


def test_decode_into_hashes_on_the_host_while_copying(md5_home):
    """zflacb200_decode with a host destination: the digest comes from the pieces as they arrive (HashWorker)."""
    import zflac_b200 as z
    from tools import flacgen as fg
    s = fg.generate(fg.c2(20.0), want_pcm=True)
    dst = np.empty(s.pcm.size, dtype=s.pcm.dtype)
    info = z.decode_into(s.flac, dst, md5_on_device=md5_home)
    assert info["md5_checked"] == 1 and info["md5_computed"] == s.md5
    assert np.array_equal(dst, s.pcm)
    bad = s.flac.copy()
    bad[8 + 18 + 3] ^= 1
    with pytest.raises(z.InvalidChecksum):
        z.decode_into(bad, dst, md5_on_device=md5_home)
