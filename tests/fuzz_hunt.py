"""Parity hunt with tests/bitcraft.py's random encoder, beyond what the test suite runs every time.

  python tests/fuzz_hunt.py make FIRST COUNT OUT.npz     # CPU: build the streams (slow, pure Python)
  python tests/fuzz_hunt.py run  OUT.npz                  # GPU: every stream through the batch call and each single-stream
                                                          # path, against the oracle
  python tests/fuzz_hunt.py damaged OUT.npz               # the same with one defect per stream (bit flip, cut, repeat):
                                                          # same error name, or the same samples where it still decodes
  python tests/fuzz_hunt.py damaged synth                 # twelve defects for each of tests/synth_cases.py's streams

Test infrastructure (not collected by pytest): the suite runs a few hundred seeds every time, this runs thousands once."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


DEPTHS = None  # make32: (32,)


def _one(seed):
    import bitcraft
    try:
        return seed, (bitcraft.random_stream(seed, DEPTHS)[0] if DEPTHS else bitcraft.random_stream(seed)[0])
    except Exception as e:  # a generator bug is a finding too
        return seed, repr(e)


def make(first, count, out):
    from multiprocessing import Pool
    with Pool(min(16, os.cpu_count() or 1)) as pool:
        res = pool.map(_one, range(first, first + count), chunksize=16)
    bad = [(s, d) for s, d in res if isinstance(d, str)]
    if bad:
        print("generator failures:", bad[:10])
    good = [(s, d) for s, d in res if not isinstance(d, str)]
    lens = np.array([d.size for _, d in good], dtype=np.int64)
    np.savez(out, seeds=np.array([s for s, _ in good]), lens=lens, blob=np.concatenate([d for _, d in good]))
    print(f"{len(good)} streams, {lens.sum() / 1e6:.1f} MB -> {out}")


def damage(data, seed, lo=42):
    """One defect per stream: a flipped bit, a zeroed or 0xFF byte, a cut, or a few bytes repeated — all after the metadata."""
    rng = np.random.default_rng(seed ^ 0x5EED)
    d = data.copy()
    if d.size <= lo + 2:
        return d
    at = int(rng.integers(lo, d.size))
    kind = int(rng.integers(0, 6))
    if kind <= 1:
        d[at] ^= np.uint8(1 << int(rng.integers(0, 8)))
    elif kind == 2:
        d[at] = 0
    elif kind == 3:
        d[at] = 0xFF
    elif kind == 4:
        d = d[:at].copy()
    else:
        n = int(rng.integers(1, 9))
        d = np.concatenate([d[:at], d[max(lo, at - n):at], d[at:]])
    return d


def run(path, damaged=False):
    from oracle import oracle as chk
    import zflac_b200 as z
    if path == "synth":  # the generator's presets (long frames, many of them: the sync scan and the frame chain), `damaged` x 12
        import synth_cases
        seeds, streams = [], []
        for k, name in enumerate(sorted(synth_cases.CASES)):
            flac = synth_cases.make(name).flac
            lo = int(chk.decode(flac).audio_offset)
            for j in range(12):
                seeds.append(1000 * k + j)
                streams.append(damage(flac, 1000 * k + j, lo))
        seeds = np.array(seeds)
        damaged = False
    else:
        f = np.load(path)
        seeds, lens, blob = f["seeds"], f["lens"], f["blob"]
        offs = np.concatenate([[0], np.cumsum(lens)])
        streams = [blob[offs[i]:offs[i + 1]].copy() for i in range(len(seeds))]
    if damaged:
        streams = [damage(d, int(s)) for d, s in zip(streams, seeds)]
    t0 = time.time()
    refs = [chk.decode(d, keep_on_md5_mismatch=True) for d in streams]
    keep = [i for i, r in enumerate(refs) if not r.overflow_flag and r.error != "UndefinedInReference"]
    import collections
    print(sorted(collections.Counter(refs[i].error for i in keep).items()))
    print(f"{len(keep)} of {len(streams)} streams are defined in the reference ({time.time() - t0:.1f} s)")
    bad = []

    def same(got, ref):
        return (not isinstance(got, Exception) and got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples)
                and (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample))

    t0 = time.time()
    for lo in range(0, len(keep), 512):
        idx = keep[lo:lo + 512]
        res = z.decode_batch([streams[i] for i in idx])
        for i, got in zip(idx, res):
            ref = refs[i]
            if ref.error != "OK":
                if not (isinstance(got, z.ZflacError) and type(got).__name__ == ref.error):
                    bad.append((int(seeds[i]), "batch", "expected " + ref.error))
            elif not same(got, ref):
                bad.append((int(seeds[i]), "batch", repr(got)[:60]))
    print(f"batch: {time.time() - t0:.1f} s, {len(bad)} findings")
    paths = [dict(), dict(use_fast_path=False), dict(word_pipeline=True), dict(single_pass=True), dict(md5_on_device=True)]
    t0 = time.time()
    for n, i in enumerate(keep):
        ref = refs[i]
        for kw in (paths[0], paths[1 + n % 4]):
            kw = dict(kw)
            try:
                if ref.error != "OK":
                    try:
                        z.decode(streams[i], **kw)
                        bad.append((int(seeds[i]), str(kw), "no " + ref.error))
                    except z.ZflacError as e:
                        if type(e).__name__ != ref.error:
                            bad.append((int(seeds[i]), str(kw), type(e).__name__ + " for " + ref.error))
                    if ref.error != "InvalidChecksum":
                        continue
                    kw["verify_md5"] = False  # without the gate: the samples the reference had when it refused them
                if not same(z.decode(streams[i], **kw), ref):
                    bad.append((int(seeds[i]), str(kw), "samples"))
            except z.ZflacError as e:
                bad.append((int(seeds[i]), str(kw), type(e).__name__))
    print(f"single: {time.time() - t0:.1f} s")
    # a block size that is no multiple of the partition count is refused (the reference reads stale residuals there)
    refused = [b for b in bad if "UndefinedInReference" in b[2] or (b[1] == "batch" and b[2].startswith("expected"))]
    print("FINDINGS", len(bad), bad[:40])
    print("of which refused as undefined (DESIGN §7, Q16) or their batch twins:", len(refused))
    return 1 if bad else 0


if __name__ == "__main__":
    if sys.argv[1] in ("make", "make32"):  # make32: 32 bits per sample only (int64 scratch, the single-pass decoder)
        if sys.argv[1] == "make32":
            DEPTHS = (32,)
        make(int(sys.argv[2]), int(sys.argv[3]), sys.argv[4])
    else:
        sys.exit(run(sys.argv[2], damaged=sys.argv[1] == "damaged"))
