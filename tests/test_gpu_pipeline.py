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
assert check(s.flac, md5_on_device=True) == "fast"     # the device chain instead of the host hash worker
assert check_device(s.flac, md5_on_device=True) == "fast"
wrong = s.flac.copy(); wrong[8 + 18 + 7] ^= 0x02      # STREAMINFO MD5 off by one bit: every home of the gate must refuse
assert check(wrong) == "error:InvalidChecksum" and check(wrong, md5_on_device=True) == "error:InvalidChecksum"
assert check_device(wrong) == "error:InvalidChecksum" and check_device(wrong, md5_on_device=True) == "error:InvalidChecksum"
s3 = fg.generate(fg.c3(2.0), want_pcm=False)          # 8 channels, order 32, 24-bit (MSB-align shift 8)
assert check(s3.flac) == "fast"                        # 3 bytes of every s32 enter the hash, shift undone (Q4)
assert check(s3.flac, md5_on_device=True) == "fast"
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


ALLOC_SCRIPT = r'''
import sys
import numpy as np
sys.path.insert(0, %(root)r)
sys.path.insert(0, %(root)r + "/tests")
import zflac_b200 as z
from oracle import oracle
from tools import flacgen as fg
import synth_cases

streams = [fg.generate(fg.c2(4.0 + i), want_pcm=False).flac for i in range(4)]
streams += [synth_cases.make(n).flac for n in ("c3_8ch24_lpc32_po7", "c4_32bit_rice2_escapes", "unknown_total", "ch6_mixed")]
refs = [oracle.decode(s) for s in streams]
for rep in range(3):                      # sizes change from call to call: blocks are reused, split off the driver, trimmed
    for s, ref in zip(streams, refs):
        got = z.decode(s)
        assert np.array_equal(got.samples, ref.samples)
    res = z.decode_batch(streams)
    assert all(np.array_equal(r.samples, ref.samples) for r, ref in zip(res, refs))
    blob, offs, lens = z.pack_streams(streams)
    dst = np.zeros(z.batch_pcm_bound(blob, offs, lens) + (8 << 20), dtype=np.uint8)
    out = z.decode_batch_into(blob, offs, lens, dst)
    assert all(np.array_equal(out.samples(i), ref.samples) for i, ref in enumerate(refs))
    parts = [z.decode_range(streams[0], lo, hi) for lo, hi in ((0, len(streams[0]) // 2), (len(streams[0]) // 2, len(streams[0])))]
    assert z.stitch_ranges(parts, parts[0].info["audio_offset"])
    assert np.array_equal(np.concatenate([p.samples for p in parts]), refs[0].samples)
z._native.load().zflacb200_shutdown()     # returns every cached block; the next call starts from nothing
assert np.array_equal(z.decode(streams[1]).samples, refs[1].samples)
print("ALLOC-OK")
'''


@pytest.mark.parametrize("env", [{"ZFLACB200_CACHE_MAX_MIB": "4"}, {"ZFLACB200_ALLOC": "pool"}, {}], ids=["tiny_cache", "driver_pool", "default"])
def test_device_memory_paths(env):
    """The library's block cache under pressure (a 4 MiB idle limit trims after almost every free), the driver's
    stream-ordered pool (A/B switch) and the default: same results through every entry point, shutdown included."""
    r = subprocess.run([sys.executable, "-c", ALLOC_SCRIPT % {"root": ROOT}], env=dict(os.environ, **env), capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0 and "ALLOC-OK" in r.stdout, r.stdout[-2000:] + r.stderr[-4000:]
