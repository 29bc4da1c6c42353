"""GPU parity of the pipelined zflacb200_decode (chunked upload / decode / read-back overlapped): same samples and
header fields as the oracle, MD5 gate included, and a clean fallback to the plain path when the hinted chain is refused.
The chunk size and the size threshold are lowered through the library's test knobs (read once per process, so this
module runs its checks in a subprocess)."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu

SCRIPT = r'''
import sys
import numpy as np
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(root)r + "/tests")
import zflac_b200 as z
from oracle import oracle
from tools import flacgen as fg
import synth_cases

def check(data, **kw):
    """zflacb200_decode (the one-call, pipelined entry point) against the oracle."""
    ref = oracle.decode(data)
    n = int(ref.samples.size) if ref.error == "OK" else 1 << 22
    dst = np.full(n + 64, 0x55, dtype=ref.samples.dtype if ref.error == "OK" else np.int32)
    try:
        info = z.decode_into(data, dst, **kw)
    except z.ZflacError as e:
        assert ref.error == type(e).__name__, (ref.error, type(e).__name__)
        return "error:" + ref.error
    assert ref.error == "OK", ref.error
    assert (info["channels"], info["sample_rate"], info["bits_per_sample"]) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
    assert info["n_samples"] == n and np.array_equal(dst[:n], ref.samples)
    assert np.all(dst[n:] == 0x55)  # nothing written past the stream's PCM
    return "fast" if info["fast_path"] else "robust"

def check_device(data, **kw):
    """zflacb200_open on a device-resident stream (chunked the same way, PCM and frame table left in HBM)."""
    import torch
    ref = oracle.decode(data, trace_frames=True)
    d = torch.from_numpy(np.ascontiguousarray(data)).cuda()
    try:
        h = z.open_device(d.data_ptr(), d.numel(), **kw)
    except z.ZflacError as e:
        assert ref.error == type(e).__name__, (ref.error, type(e).__name__)
        return "error:" + ref.error
    assert ref.error == "OK", ref.error
    assert (h.channels, h.sample_rate, h.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
    assert np.array_equal(h.to_host(), ref.samples)
    fr = h.frames()
    assert len(fr) == len(ref.frames) and np.array_equal(fr["offset"], ref.frames["offset"]) and np.array_equal(fr["end_offset"], ref.frames["end_offset"])
    fast = h.info["fast_path"]
    h.close()
    return "fast" if fast else "robust"

s = fg.generate(fg.c2(30.0), want_pcm=False)          # ~2.5 MB, 324 frames: a dozen chunks of 256 KiB
assert check(s.flac) == "fast"
assert check_device(s.flac) == "fast"
assert check_device(s.flac, verify_md5=False) == "fast"
assert check(s.flac, verify_md5=False) == "fast"
s3 = fg.generate(fg.c3(2.0), want_pcm=False)          # 8 channels, order 32, 24-bit (MSB-align shift 8)
assert check(s3.flac) == "fast"
bad = s.flac.copy(); bad[bad.size // 2] ^= 0x10       # payload damage: chain or MD5 refuses -> plain path decides
print("damaged:", check(bad), check_device(bad))
cut = s.flac[: s.flac.size * 2 // 3].copy()           # truncated
print("truncated:", check(cut), check_device(cut))
assert check_device(s3.flac) == "fast"
for name in ("c4_variable_blocksize", "c4_32bit_rice2_escapes", "unknown_total", "ch6_mixed", "bps12_midside"):
    print(name, check(synth_cases.make(name).flac), check_device(synth_cases.make(name).flac))
print("PIPELINE-OK")
'''


def test_pipelined_decode_matches_oracle():
    env = dict(os.environ, ZFLACB200_PIPELINE_MIN_BYTES="0", ZFLACB200_PIPELINE_CHUNK=str(256 << 10), ZFLACB200_DEVICE_CHUNKS="6")
    r = subprocess.run([sys.executable, "-c", SCRIPT % {"root": ROOT}], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "PIPELINE-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
