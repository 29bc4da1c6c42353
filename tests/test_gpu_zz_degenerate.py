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
