"""GPU parity at the sizes BASELINE.json quotes and on every instantiation of pass A.

The frame-serial kernel has six instantiations chosen by the launch's frame count, the stream's bit depth and (for the
first bit reader, kept as a diagnostic) the ZFLACB200_WORD_PIPELINE flag
(zflacb200_pass_a_launches): the tests below make each of them run and compare the result with the generator's PCM
sample for sample and with the oracle (whole stream where that takes seconds, a prefix / a sample of files otherwise).
Full-size configs: C3 (600 s, 8 channels, 24 bit, 12 500 frames, 1.84 GB of PCM) and C5 (10 000 files in one batch)."""
import numpy as np
import pytest

from oracle import oracle
from tools import flacgen as fg

pytestmark = pytest.mark.gpu

STAGED, QUAD, NARROW64, NARROW32, WIDE64, WIDE32 = 0, 1, 2, 3, 4, 5


def _launches():
    from zflac_b200 import _native
    lib = _native.load()
    return [int(lib.zflacb200_pass_a_launches(v)) for v in range(6)]


def _ran(before, variant):
    return _launches()[variant] > before[variant]


@pytest.mark.parametrize("bps,variant,word_pipeline", [(16, NARROW32, False), (24, WIDE32, False), (24, QUAD, True)],
                         ids=["narrow32_16bit", "wide32_24bit", "quad_24bit"])
def test_launch_larger_than_one_wave(bps, variant, word_pipeline):
    """90 000 frames of 16 samples in ONE launch: more than one wave of the deep-ring kernels holds."""
    import zflac_b200 as z
    p = fg.params(block_size=16, bps=bps, lpc_order=2, qlp_precision=7, max_partition_order=2, stereo_mode=4,
                  n_samples=16 * 90000 + 5, seed=4242 + bps, rice_method=1 if bps > 16 else 0)
    s = fg.generate(p, want_pcm=True)
    before = _launches()
    got = z.decode(s.flac, word_pipeline=word_pipeline)
    assert _ran(before, variant), (before, _launches())
    assert got.info["fast_path"] == 1 and got.info["n_frames"] == s.n_frames
    assert np.array_equal(got.samples, s.pcm)
    ref = oracle.decode(s.flac)
    assert ref.error == "OK" and np.array_equal(got.samples, ref.samples) and got.info["md5_computed"] == ref.md5_computed


def test_small_launches_take_the_deep_ring_kernels():
    import zflac_b200 as z
    before = _launches()
    z.decode(fg.generate(fg.c2(5.0), want_pcm=False).flac)
    assert _ran(before, NARROW64)
    s = fg.generate(fg.c3(1.0), want_pcm=True)
    before = _launches()
    assert np.array_equal(z.decode(s.flac).samples, s.pcm)
    assert _ran(before, WIDE64)
    before = _launches()
    assert np.array_equal(z.decode(s.flac, word_pipeline=True).samples, s.pcm)
    assert _ran(before, STAGED)


def test_word_pipeline_and_position_reader_agree_on_deep_streams():
    """Both pass-A implementations for 17..31-bit streams give the generator's PCM: 20 and 24 bits, escapes, wasted bits,
    every channel assignment, a Rice parameter range that reaches past 16."""
    import zflac_b200 as z
    cases = [fg.params(channels=2, bps=20, stereo_mode=4, n_samples=60000, rice_method=1, max_partition_order=8, signal=1, subframe_mix=1),
             fg.params(channels=3, bps=24, n_samples=50000, rice_method=1, subframe_mix=1, wasted_bits=1, escape_every=5, signal=1),
             fg.params(channels=2, bps=28, stereo_mode=3, n_samples=40000, rice_method=1, lpc_order=12, qlp_precision=14, noise=0.05),
             fg.c3(2.0)]
    for p in cases:
        s = fg.generate(p, want_pcm=True)
        for wp in (False, True):
            before = _launches()
            got = z.decode(s.flac, word_pipeline=wp)
            assert np.array_equal(got.samples, s.pcm), (int(p.bps), wp)
            assert _ran(before, STAGED if wp else WIDE64)


def test_c3_full_size():
    """BASELINE configs[2] at its quoted size: 600 s, 24 bit, 96 kHz, 8 channels, block size 4608, LPC order 32."""
    import torch
    import zflac_b200 as z
    s = fg.generate(fg.c3(600.0), want_pcm=True)
    d = torch.from_numpy(s.flac).cuda()
    h = z.open_device(d.data_ptr(), d.numel())            # MD5 gate on (3 bytes of every s32, shift undone)
    assert h.info["fast_path"] == 1 and h.info["n_frames"] == s.n_frames == 12500
    assert h.info["md5_computed"] == s.md5 == h.info["md5_streaminfo"]
    got = h.to_host()
    h.close()
    assert got.dtype == s.pcm.dtype and np.array_equal(got, s.pcm)
    head = fg.generate(fg.c3(20.0), want_pcm=False)       # same seed: the same frames, up to the short stream's last ones
    ref = oracle.decode(head.flac)
    n = ref.samples.size - 4 * 4608 * 8
    assert ref.error == "OK" and np.array_equal(got[:n], ref.samples[:n])


def test_c5_full_size_batch():
    """BASELINE configs[4] at its quoted size: 10 000 files in one batch call, MD5 gate on for every file."""
    import os
    import zflac_b200 as z
    files = fg.generate_batch([fg.c5_file(i) for i in range(10000)], want_pcm=True, n_threads=os.cpu_count() or 1)
    before = _launches()
    res = z.decode_batch([f.flac for f in files])
    assert _ran(before, NARROW32) or _ran(before, NARROW64)
    assert len(res) == len(files)
    for f, r in zip(files, res):
        assert r.info["md5_checked"] == 1 and r.info["md5_computed"] == f.md5
        assert r.samples.dtype == f.pcm.dtype and np.array_equal(r.samples, f.pcm)
    for i in range(0, 10000, 499):                        # the oracle on a sample of the files
        ref = oracle.decode(files[i].flac)
        assert ref.error == "OK" and np.array_equal(res[i].samples, ref.samples)
        assert (res[i].channels, res[i].sample_rate, res[i].bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
