#!/usr/bin/env python3
"""Per-file parity table on a GPU box: CUDA path vs oracle over the committed corpus."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import zflac_b200 as z  # noqa: E402
from oracle import oracle  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")
only = sys.argv[1:] if len(sys.argv) > 1 else None
bad = 0
for kat in json.load(open(os.path.join(G, "kat.json"))):
    try:
        r = z.decode(bytes.fromhex(kat["flac_hex"]))
        ok = r.samples.tolist() == kat["samples"]
        print(f"KAT {kat['name']}: {'OK' if ok else 'MISMATCH ' + str(r.samples.tolist()[:8])}", flush=True)
        bad += not ok
    except Exception as e:  # noqa: BLE001
        print(f"KAT {kat['name']}: EXC {type(e).__name__} {e}", flush=True)
        bad += 1
for e in json.load(open(os.path.join(G, "manifest.json")))["files"]:
    tag = f"{e['group']}_{e['id']}"
    if only and tag not in only:
        continue
    data = open(os.path.join(G, e["file"]), "rb").read()
    ref = oracle.decode(data)
    t = time.time()
    try:
        got = z.decode(data)
        dt = time.time() - t
        if ref.error != "OK":
            res = f"MISMATCH: gpu OK, oracle {ref.error}"
            bad += 1
        else:
            same = got.samples.dtype == ref.samples.dtype and np.array_equal(got.samples, ref.samples)
            fields = (got.channels, got.sample_rate, got.bits_per_sample) == (ref.channels, ref.sample_rate, ref.bits_per_sample)
            i = got.info
            res = (f"{'OK' if same and fields else 'MISMATCH'} n={got.samples.size} frames={i['n_frames']} cand={i['n_candidates']} "
                   f"probed={i['n_probed_frames']} crcbad={i['n_crc16_mismatch']} ms: scan={i['ms_scan']:.3f} link={i['ms_link']:.3f} "
                   f"dec={i['ms_decode']:.3f} epi={i['ms_epilogue']:.3f} md5={i['ms_md5']:.3f} tot={i['ms_total']:.3f} wall={dt*1e3:.1f}")
            if not (same and fields):
                bad += 1
                if got.samples.size == ref.samples.size:
                    d = np.nonzero(got.samples != ref.samples)[0]
                    res += f" first_diff@{d[0]} ndiff={d.size} got={got.samples[d[0]:d[0]+4]} ref={ref.samples[d[0]:d[0]+4]}"
                else:
                    res += f" size gpu={got.samples.size} ref={ref.samples.size}"
    except z.ZflacError as ex:
        name = type(ex).__name__
        ok = name == ref.error
        res = f"{'OK' if ok else 'MISMATCH'} error gpu={name} oracle={ref.error} {ex if not ok else ''}"
        bad += not ok
        if not ok and name == "InvalidChecksum":
            got = z.decode(data, verify_md5=False)
            if got.samples.size == ref.samples.size:
                d = np.nonzero(got.samples != ref.samples)[0]
                res += f" nomd5: ndiff={d.size}" + (f" first@{d[0]} got={got.samples[d[0]:d[0]+6]} ref={ref.samples[d[0]:d[0]+6]}" if d.size else "")
            else:
                res += f" nomd5: size gpu={got.samples.size} ref={ref.samples.size}"
            g2 = z.open_device(0, 0, host_data=data, verify_md5=False)
            res += f" frames={g2.info['n_frames']} probed={g2.info['n_probed_frames']}"
            try:
                z.decode(data)
            except z.ZflacError:
                pass
    print(f"{tag:12s} {res}", flush=True)
print(f"TOTAL BAD: {bad}")
sys.exit(1 if bad else 0)
