#!/usr/bin/env python3
"""Short driver for ncu: decodes one synthetic stream a few times with the stream resident in HBM.

    python tools/profile_run.py [workload=c2|c3|c5] [seconds=3600 | files] [reps=2] [md5=0]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import zflac_b200 as z  # noqa: E402
from tools import flacgen as fg  # noqa: E402

wl = sys.argv[1] if len(sys.argv) > 1 else "c2"
secs = float(sys.argv[2]) if len(sys.argv) > 2 else 3600.0
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
md5 = bool(int(sys.argv[4])) if len(sys.argv) > 4 else False
if wl == "c5":  # batch of short files (secs = number of files here), device-resident blob
    n = int(secs)
    files = fg.generate_batch([fg.c5_file(i) for i in range(n)], want_pcm=False, n_threads=os.cpu_count() or 1)
    blob, offs, lens = z.pack_streams([f.flac for f in files])
    h = torch.from_numpy(blob).pin_memory()
    d = h.to("cuda")
    for i in range(reps):
        t = time.time()
        b = z.open_batch(h.numpy(), offs, lens, device_blob_ptr=d.data_ptr(), verify_md5=md5)
        i_ = b.infos[0]
        b.close()
        print(f"rep {i}: files={n} ms scan={i_['ms_scan']:.3f} link={i_['ms_link']:.3f} dec={i_['ms_decode']:.3f} "
              f"epi={i_['ms_epilogue']:.3f} md5={i_['ms_md5']:.3f} wall={1e3 * (time.time() - t):.1f}", flush=True)
    torch.cuda.synchronize()
    sys.exit(0)
p = {"c2": fg.c2, "c3": fg.c3}[wl](secs)
s = fg.generate(p, want_pcm=False)
d = torch.from_numpy(s.flac).cuda()
for i in range(reps):
    t = time.time()
    h = z.open_device(d.data_ptr(), s.flac.size, verify_md5=md5)
    i_ = h.info
    h.close()
    print(f"rep {i}: n={i_['n_samples']} frames={i_['n_frames']} ms scan={i_['ms_scan']:.3f} link={i_['ms_link']:.3f} "
          f"dec={i_['ms_decode']:.3f} epi={i_['ms_epilogue']:.3f} md5={i_['ms_md5']:.3f} total={i_['ms_total']:.3f} wall={1e3 * (time.time() - t):.1f}",
          flush=True)
torch.cuda.synchronize()
