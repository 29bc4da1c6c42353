#!/usr/bin/env python3
"""Generate tests/golden/ from the reference checkout (run in the build container only).

/root/reference does not exist on the GPU box, so everything the -m gpu tests
need travels as fixtures:

  tests/golden/corpus/<group>_<nn>.flac   the IETF flac-test-files inputs the reference's tests
                                          decode (CC0 data, the reference's git submodule
                                          test-files/ietf-wg-cellar; NOT reference source code)
  tests/golden/manifest.json              per file: what the reference's own tests pin
  tests/golden/kat.json                   the 3 RFC 9639 known-answer vectors of tests/basic.zig

How each expectation is pinned (manifest["files"][i]["pinned_by"]):
  golden_raw      sha256 of tests/expected_samples/<name>.raw converted to the API layout
                  (8-bit raws are unsigned: -128, tests/std_subset.zig:23-30). Independent of the oracle.
  streaminfo_md5  the reference's own pass criterion (zflac.zig:279-280): the oracle's output
                  reproduced the STREAMINFO MD5; sha256 is of the oracle's API-layout bytes.
  error_name      the Zig error name the reference test expects (std_uncommon.zig / std_faulty.zig).
"""
import glob
import hashlib
import json
import os
import re
import shutil
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle  # noqa: E402

REF = "/root/reference"
OUT = os.path.join(ROOT, "tests", "golden")

EXPECTED_ERRORS = {  # tests/std_uncommon.zig:17-59, tests/std_faulty.zig:17-61
    ("uncommon", "01"): "InconsistentParameters", ("uncommon", "02"): "InconsistentParameters",
    ("uncommon", "03"): "InconsistentParameters", ("uncommon", "04"): "InconsistentParameters",
    ("uncommon", "10"): "InvalidSignature", ("uncommon", "11"): "InvalidSignature",
    ("faulty", "03"): "InvalidChecksum", ("faulty", "04"): "InconsistentParameters",
    ("faulty", "06"): "MissingStreaminfo", ("faulty", "08"): "InvalidFrameHeader",
    ("faulty", "09"): "InvalidFrameHeader", ("faulty", "11"): "InvalidMetadataHeader",
}


def parse_kats():
    """Pull the byte arrays and expected values out of tests/basic.zig."""
    src = open(os.path.join(REF, "tests", "basic.zig")).read()
    kats = []
    for m in re.finditer(r'test "(Example \d)" \{(.*?)\n\}\n', src, re.S):
        name, body = m.group(1), m.group(2)
        arr = re.search(r"const Example = \[_\]u8\{(.*?)\};", body, re.S).group(1)
        data = bytes(int(x, 16) for x in re.findall(r"0x([0-9a-fA-F]{2})", arr))
        kats.append({"name": name, "flac_hex": data.hex()})
    # expected values, transcribed from tests/basic.zig:17-21, :50-74, :92-94
    kats[0].update(channels=2, container_bytes=2, samples=[25588, 10416])
    ex2 = [10372, 6070, 18041, 10545, 14942, 8743, 17876, 10449, 15627, 9143, 17899, 10463, 16242, 9502,
           18077, 10569, 16824, 9840, 18263, 10680, 17295, 10113, -14418, -8428, -15201, -8895, -14508, -8476,
           -15195, -8896, -14818, -8653]
    tail = [0xc382, 0b1101110010010000, 0xc40b, 0b1101110100000010, 0xc14a, 0b1101101100111110]
    ex2 += [v - 65536 if v >= 32768 else v for v in tail]
    kats[1].update(channels=2, container_bytes=2, samples=ex2)
    kats[2].update(channels=1, container_bytes=1,
                   samples=[0, 79, 111, 78, 8, -61, -90, -68, -13, 42, 67, 53, 13, -27, -46, -38, -12, 14, 24, 19, 6,
                            -4, -5, 0])
    return kats


def main():
    os.makedirs(os.path.join(OUT, "corpus"), exist_ok=True)
    files = []
    for grp in ("subset", "uncommon", "faulty"):
        for path in sorted(glob.glob(f"{REF}/test-files/ietf-wg-cellar/{grp}/*.flac")):
            name = os.path.basename(path)[:-5]
            nn = name[:2]
            dst = f"corpus/{grp}_{nn}.flac"
            shutil.copyfile(path, os.path.join(OUT, dst))
            data = open(path, "rb").read()
            r = oracle.decode(data, trace_frames=True)
            entry = {"group": grp, "id": nn, "name": name, "file": dst, "bytes": len(data),
                     "flac_sha256": hashlib.sha256(data).hexdigest()}
            exp_err = EXPECTED_ERRORS.get((grp, nn))
            if exp_err:
                assert r.error == exp_err, (name, r.error, exp_err)
                entry.update(expect_error=exp_err, pinned_by="error_name")
            else:
                assert r.error == "OK", (name, r.error)
                assert r.overflow_flag == 0, name
                got = r.samples.view(np.uint8)
                raw = f"{REF}/tests/expected_samples/{name}.raw"
                pinned = "streaminfo_md5"
                if grp == "subset" and os.path.exists(raw):
                    exp = np.fromfile(raw, dtype=np.uint8)
                    if r.container_bytes == 1:
                        exp = (exp.astype(np.int16) - 128).astype(np.int8).view(np.uint8)
                    assert exp.size == got.size and np.array_equal(exp, got), f"golden raw mismatch: {name}"
                    sha = hashlib.sha256(exp.tobytes()).hexdigest()
                    pinned = "golden_raw"
                else:
                    sha = hashlib.sha256(got.tobytes()).hexdigest()
                entry.update(expect_error=None, pinned_by=pinned, channels=r.channels, sample_rate=r.sample_rate,
                             bits_per_sample=r.bits_per_sample, container_bytes=r.container_bytes,
                             n_samples=int(r.samples.size), pcm_sha256=sha, md5_streaminfo=r.md5_streaminfo.hex(),
                             n_frames=int(r.frames.size), audio_offset=int(r.audio_offset),
                             first_samples=[int(x) for x in r.samples[:16]])
            files.append(entry)
            print(f"{dst:28s} {entry['pinned_by']:15s} {exp_err or 'OK'}")
    manifest = {
        "source": "Senryoku/zflac @ 92313a2: test-files/ietf-wg-cellar (submodule aa7b0c6, CC0) + tests/expected_samples",
        "generator": "tools/make_golden.py",
        "missing_in_checkout": "subset 33-36, 44, 48-53, 55 and uncommon 06 (.MISSING_LARGE_BLOBS)",
        "files": files,
    }
    json.dump(manifest, open(os.path.join(OUT, "manifest.json"), "w"), indent=1)
    json.dump(parse_kats(), open(os.path.join(OUT, "kat.json"), "w"), indent=1)
    shutil.copyfile(f"{REF}/test-files/ietf-wg-cellar/LICENSE.txt", os.path.join(OUT, "corpus", "LICENSE.txt"))
    n_gold = sum(f["pinned_by"] == "golden_raw" for f in files)
    n_md5 = sum(f["pinned_by"] == "streaminfo_md5" for f in files)
    n_err = sum(f["pinned_by"] == "error_name" for f in files)
    print(f"{len(files)} files: {n_gold} golden_raw, {n_md5} streaminfo_md5, {n_err} error_name")


if __name__ == "__main__":
    main()
