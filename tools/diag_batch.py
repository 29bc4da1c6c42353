#!/usr/bin/env python3
"""Where does a batch step spend its HOST time?  python tools/diag_batch.py [files=10000] [reps=3]

Prints wall-clock per Python-level call next to the library's own host trace (ZFLACB200_TRACE=1).
"""
import os
import sys
import time

os.environ["ZFLACB200_TRACE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

import zflac_b200 as z  # noqa: E402
from tools import flacgen as fg  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10000
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
files = fg.generate_batch([fg.c5_file(i) for i in range(n)], want_pcm=False, n_threads=os.cpu_count() or 1)
blob, offs, lens = z.pack_streams([f.flac for f in files])
h_blob = torch.from_numpy(blob).pin_memory()
d_blob = h_blob.to("cuda")
b = z.open_batch(h_blob.numpy(), offs, lens, device_blob_ptr=d_blob.data_ptr())
h_pcm = torch.empty(b.pcm_bytes(), dtype=torch.uint8).pin_memory()
b.close()
for rep in range(reps):
    for mode in ("device", "e2e"):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        bb = z.open_batch(h_blob.numpy(), offs, lens, device_blob_ptr=d_blob.data_ptr() if mode == "device" else 0)
        t1 = time.perf_counter()
        if mode == "e2e":
            bb.read_all(h_pcm.numpy())
        t2 = time.perf_counter()
        bb.close()
        t3 = time.perf_counter()
        print(f"rep {rep} {mode}: open {1e3 * (t1 - t0):.1f} ms, read_all {1e3 * (t2 - t1):.1f} ms, close {1e3 * (t3 - t2):.1f} ms", flush=True)
