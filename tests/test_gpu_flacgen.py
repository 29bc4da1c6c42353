"""the CUDA generator (tools/flacgen_cuda.cu) writes the SAME BYTES as the CPU generator
(tools/flacgen.cpp) — stream, PCM and MD5 — for every preset of tools/flacgen.py, single streams and batches, and what it
writes decodes (library and oracle) to the PCM it reports. On a mismatch the per-subframe trace of both generators says
which stage parted ways."""
import numpy as np
import pytest

from oracle import oracle
from tools import flacgen as fg

pytestmark = pytest.mark.gpu

TRACE_FIELDS = ("type", "wasted", "precision", "shift", "partition_order", "bits", "sample_checksum", "coef_checksum")


def _diagnose(p, cpu, dev_flac, dev_pcm):
    """Where the two generators part ways, as text for the assertion message."""
    lines = []
    if dev_pcm is not None and cpu.pcm is not None:
        a, b = cpu.pcm.view(np.uint8), dev_pcm
        if a.size != b.size:
            lines.append(f"PCM sizes differ: cpu {a.size} device {b.size}")
        else:
            bad = np.flatnonzero(a != b)
            if bad.size:
                cb = cpu.container_bytes
                s = int(bad[0]) // cb
                lines.append(f"PCM differs at sample {s} (per-channel index {s // int(p.channels)}), {bad.size} bytes differ")
    n = min(cpu.flac.size, dev_flac.size)
    bad = np.flatnonzero(cpu.flac[:n] != dev_flac[:n])
    lines.append(f"stream sizes: cpu {cpu.flac.size} device {dev_flac.size}; first differing byte: {int(bad[0]) if bad.size else None}"
                 f" ({bad.size} differ in the common prefix)")
    try:
        _, t_cpu = fg.cpu_trace(p)
        d = fg.generate_device(p, keep_trace=True)
        t_dev = d.trace()
        d.close()
        m = min(len(t_cpu), len(t_dev))
        rows = np.flatnonzero((t_cpu[:m] != t_dev[:m]).any(axis=1))
        lines.append(f"subframes: cpu {len(t_cpu)} device {len(t_dev)}; {rows.size} records differ")
        for r in rows[:6]:
            diff = {TRACE_FIELDS[k]: (int(t_cpu[r, k]), int(t_dev[r, k])) for k in range(8) if t_cpu[r, k] != t_dev[r, k]}
            lines.append(f"  subframe {r} (frame {r // int(p.channels)}, channel {r % int(p.channels)}): cpu type {int(t_cpu[r, 0])} -> (cpu, device) {diff}")
    except Exception as e:  # noqa: BLE001
        lines.append(f"(trace unavailable: {type(e).__name__} {e})")
    return "\n".join(lines)


def _check_single(p):
    cpu = fg.generate(p)
    d = fg.generate_device(p)
    try:
        flac, pcm = d.to_host()
        s = d.streams[0]
        assert d.flac_bytes == s["flac_bytes"] and s["flac_off"] == 0
        same = flac.size == cpu.flac.size and np.array_equal(flac, cpu.flac) and np.array_equal(pcm, cpu.pcm.view(np.uint8))
        assert same, _diagnose(p, cpu, flac, pcm)
        assert s["md5"] == cpu.md5 and s["n_frames"] == cpu.n_frames and s["container_bytes"] == cpu.container_bytes
    finally:
        d.close()
    return cpu


PRESETS = {
    "c2": lambda: fg.c2(20.0),
    "c2_odd_tail": lambda: fg.c2(3.37),
    "c3": lambda: fg.c3(3.0),
    "c4_variable": lambda: fg.c4_variable(10.0),
    "c4_32bit": lambda: fg.c4_32bit(6.0),
    "c4_wasted": lambda: fg.c4_wasted(10.0),
    "c4_bs65535": lambda: fg.c4_bs65535(8.0),
    "c5_file_0": lambda: fg.c5_file(0),
    "c5_file_17": lambda: fg.c5_file(17),
    "mono_8bit_mix": lambda: fg.params(channels=1, bps=8, n_samples=30000, subframe_mix=1),
    "three_ch_12bit": lambda: fg.params(channels=3, bps=12, n_samples=30000, subframe_mix=1, wasted_bits=1, escape_every=5),
    "stereo_20bit": lambda: fg.params(channels=2, bps=20, stereo_mode=4, n_samples=50000, rice_method=1, max_partition_order=8, signal=1),
    "unknown_total": lambda: fg.params(set_total=0, n_samples=20000, stereo_mode=1),
    "tiny": lambda: fg.params(n_samples=5, channels=2, stereo_mode=3),
}


@pytest.mark.parametrize("name", sorted(PRESETS))
def test_device_generator_writes_the_cpu_generators_bytes(name):
    _check_single(PRESETS[name]())


def test_batch_blob_matches_the_cpu_batch_and_pack_streams():
    import zflac_b200 as z
    ps = [fg.c5_file(i) for i in range(96)]
    cpu = fg.generate_batch(ps, want_pcm=True)
    d = fg.generate_device(ps, align=16)
    try:
        flac, pcm = d.to_host()
        blob, offs, lens = z.pack_streams([c.flac for c in cpu])
        assert [s["flac_off"] for s in d.streams] == [int(o) for o in offs]
        assert [s["flac_bytes"] for s in d.streams] == [int(n) for n in lens]
        assert np.array_equal(flac, blob[: flac.size]), "device blob differs from pack_streams() of the CPU streams"
        for c, s in zip(cpu, d.streams):
            assert s["md5"] == c.md5
            assert np.array_equal(pcm[s["pcm_off"]: s["pcm_off"] + s["pcm_bytes"]], c.pcm.view(np.uint8))
    finally:
        d.close()


def test_device_generated_stream_decodes_to_its_pcm():
    """The generator closes the loop without the CPU generator: library and oracle decode the device-made stream to the
    device-made PCM, MD5 gate on."""
    import zflac_b200 as z
    p = fg.c2(30.0)
    d = fg.generate_device(p)
    try:
        s = d.synths()[0]
    finally:
        d.close()
    got = z.decode(s.flac, verify_md5=True)
    assert np.array_equal(got.samples, s.pcm)
    ref = oracle.decode(s.flac)
    assert ref.error == "OK" and np.array_equal(ref.samples, s.pcm)


def test_full_size_c2_in_hbm_and_its_generator_time():
    """BASELINE configs[1] built in HBM: decodes (device in, device out) to the generator's PCM; the device part of the
    generator takes a fraction of a second (the MD5 of the PCM is a serial chain on one host core: reported apart)."""
    import torch
    import zflac_b200 as z
    p = fg.c2(3600.0)
    d = fg.generate_device(p)
    try:
        t = d.seconds
        print("generator seconds:", {k: round(v, 3) for k, v in t.items()})
        assert t["total"] - t["md5"] < 1.0
        flac = torch.empty(d.flac_bytes, dtype=torch.uint8, device="cuda")
        pcm = torch.empty(d.pcm_bytes, dtype=torch.uint8, device="cuda")
        d.copy_to(flac.data_ptr(), True, pcm.data_ptr(), True)
    finally:
        d.close()
    h = z.open_device(flac.data_ptr(), flac.numel(), verify_md5=True)
    try:
        assert h.info["md5_checked"] == 1 and h.pcm_bytes == pcm.numel()
        got = h.to_host()
        assert np.array_equal(got.view(np.uint8), pcm.cpu().numpy())
    finally:
        h.close()
