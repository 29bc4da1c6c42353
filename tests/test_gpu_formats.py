"""GPU: zflacb200_read_as (additive) — the output conventions the reference leaves open
(Readme.md:53, src/zflac.zig:282-292) are exact functions of the reference's own PCM, so the checker is the oracle's
PCM pushed through the same definition in numpy. Bit-exact, floats included (int -> float32 rounds to nearest even on
both sides, the scaling is a power of two)."""
import numpy as np
import pytest

from conftest import corpus_bytes, corpus_entries

pytestmark = pytest.mark.gpu

OK = {f"{e['group']}_{e['id']}": e for e in corpus_entries() if not e["expect_error"]}


def pick(container_bytes, limit):
    ids = [k for k, e in OK.items() if e["container_bytes"] == container_bytes]
    return ids[:: max(1, len(ids) // limit)][:limit]


@pytest.fixture(scope="module")
def z():
    import zflac_b200
    return zflac_b200


def opened(z, fid):
    from oracle import oracle
    data = corpus_bytes(OK[fid])
    ref = oracle.decode(data)
    return z.open_device(0, 0, host_data=data), ref


@pytest.mark.parametrize("fid", pick(1, 3) + pick(2, 5) + pick(4, 5))
def test_f32_and_native(z, fid):
    s, ref = opened(z, fid)
    assert np.array_equal(s.to_host_as("native"), ref.samples)
    scale = np.float32(2.0 ** -(8 * ref.samples.dtype.itemsize - 1))
    want = ref.samples.astype(np.float32) * scale
    got = s.to_host_as("f32")
    assert got.dtype == np.float32 and got.size == ref.samples.size
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert float(np.abs(got).max()) <= 1.0
    s.close()


@pytest.mark.parametrize("fid", pick(4, 6))
def test_s24_packed(z, fid):
    s, ref = opened(z, fid)
    got = s.to_host_as("s24")
    v = ref.samples.astype(np.int32) >> 8
    want = np.stack([(v & 0xFF), (v >> 8) & 0xFF, (v >> 16) & 0xFF], axis=1).astype(np.uint8).reshape(-1)
    assert got.dtype == np.uint8 and np.array_equal(got, want)
    if OK[fid]["bits_per_sample"] <= 24:  # lossless: back to the reference's MSB-aligned s32
        g = got.reshape(-1, 3).astype(np.int32)
        back = ((g[:, 0] | (g[:, 1] << 8) | (g[:, 2] << 16)) << 8).astype(np.int32)
        assert np.array_equal(back, ref.samples)
    with pytest.raises(z.ZflacError):
        s.to_host_as("u8")
    s.close()


@pytest.mark.parametrize("fid", pick(1, 3))
def test_u8(z, fid):
    s, ref = opened(z, fid)
    got = s.to_host_as("u8")
    assert np.array_equal(got, (ref.samples.astype(np.int16) + 128).astype(np.uint8))
    with pytest.raises(z.ZflacError):
        s.to_host_as("s24")
    s.close()


def test_odd_lengths_hit_the_tail_loop(z):
    """Sample counts that are not a multiple of 4, on a synthetic stream (every container)."""
    from oracle import oracle
    from tools import flacgen as fg
    cases = (dict(bps=8, channels=1, lpc_order=2, qlp_precision=6, n_samples=4099, amplitude=0.3, noise=0.01),
             dict(bps=16, channels=3, n_samples=1001), dict(bps=24, channels=1, n_samples=4097, noise=0.0003),
             dict(bps=20, channels=2, n_samples=515, lpc_order=4))
    for k, kw in enumerate(cases):
        st = fg.generate(fg.params(block_size=512, seed=300 + k, **kw))
        data = st.flac
        ref = oracle.decode(data)
        assert ref.error == "OK"
        assert ref.samples.size % 4 != 0 and np.array_equal(ref.samples, st.pcm)
        s = z.open_device(0, 0, host_data=data)
        want = ref.samples.astype(np.float32) * np.float32(2.0 ** -(8 * ref.samples.dtype.itemsize - 1))
        assert np.array_equal(s.to_host_as("f32").view(np.uint32), want.view(np.uint32))
        if ref.samples.dtype.itemsize == 4:
            v = ref.samples >> 8
            want24 = np.stack([(v & 0xFF), (v >> 8) & 0xFF, (v >> 16) & 0xFF], axis=1).astype(np.uint8).reshape(-1)
            assert np.array_equal(s.to_host_as("s24"), want24)
        if ref.samples.dtype.itemsize == 1:
            assert np.array_equal(s.to_host_as("u8"), (ref.samples.astype(np.int16) + 128).astype(np.uint8))
        s.close()
