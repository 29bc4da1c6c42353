"""CPU: the oracle decodes every synthetic stream to exactly the PCM the generator encoded (and the same MD5),
without leaving the reference's defined behaviour (overflow_flag == 0). Pins generator and oracle to each other."""
import numpy as np
import pytest

from oracle import oracle
from synth_cases import CASES, make


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_decodes_synthetic(name):
    s = make(name)
    r = oracle.decode(s.flac, trace_frames=True)
    assert r.error == "OK"
    assert r.overflow_flag == 0
    assert r.container_bytes == s.container_bytes
    assert (r.channels, r.bits_per_sample) == (s.params.channels, s.params.bps)
    assert np.array_equal(r.samples, s.pcm)
    assert r.md5_computed == r.md5_streaminfo == s.md5
    assert r.frames.size == s.n_frames


def test_generator_is_deterministic_and_thread_independent():
    from tools import flacgen as fg
    a = fg.generate(fg.c2(4.0, n_threads=1))
    b = fg.generate(fg.c2(4.0, n_threads=4))
    assert np.array_equal(a.flac, b.flac) and a.md5 == b.md5


def test_corrupted_frames_do_not_change_reference_semantics():
    """CRC-16 / CRC-8 damage is ignored by the reference (Q1): flipping a CRC-16 byte still decodes;
    flipping a payload bit changes samples and trips the MD5 gate."""
    s = make("c5_file_0")
    r0 = oracle.decode(s.flac, trace_frames=True)
    end = int(r0.frames["end_offset"][3])
    bad_crc = s.flac.copy()
    bad_crc[end - 1] ^= 0x5A
    assert oracle.decode(bad_crc).error == "OK"
    bad_payload = s.flac.copy()
    bad_payload[end - 40] ^= 0x10
    assert oracle.decode(bad_payload).error != "OK"


def test_cuda_generator_library_builds_and_exports_its_interface():
    """tools/flacgen_cuda.cu cross-compiles for sm_100a and exports the entry points tools/flacgen.py
    binds; without a device a call fails with a CUDA error instead of falling back to the CPU generator."""
    import ctypes as C
    from tools import flacgen as fg
    lib = C.CDLL(fg.build_cuda())
    for name in ("fgc_generate", "fgc_info", "fgc_copy", "fgc_trace", "fgc_free", "fgc_last_error"):
        assert hasattr(lib, name), name
    import torch
    if not torch.cuda.is_available():
        import pytest
        with pytest.raises(RuntimeError, match="code 3"):
            fg.generate_device(fg.c2(1.0))


def test_cpu_generator_trace_describes_every_subframe():
    from tools import flacgen as fg
    p = fg.c4_variable(3.0)
    s, t = fg.cpu_trace(p)
    assert t.shape == (s.n_frames * 2, 8)
    assert set(np.unique(t[:, 0])) <= set(range(0, 64)) and (t[:, 5] > 0).all()   # a type and a bit count each
    assert np.array_equal(fg.generate(p).flac, s.flac)                               # tracing changes nothing
