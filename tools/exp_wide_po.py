"""Experiment: pass A of a C3-shaped stream (24 bit, 8 channels, LPC 32) for several Rice partition orders, position reader
with int32 rows against the word pipeline (ZFLACB200_WORD_PIPELINE). Prints ms_decode of the device-resident decode."""
import sys
import numpy as np
import torch
sys.path.insert(0, ".")
import zflac_b200 as z
from zflac_b200 import _native
from tools import flacgen as fg

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 300.0
for po in (7, 5, 3, 0):
    p = fg.c3(secs)
    p.min_partition_order = p.max_partition_order = po
    d = fg.generate_device(p)
    flac = torch.empty(d.flac_bytes, dtype=torch.uint8, device="cuda")
    d.copy_to(flac.data_ptr(), True, 0, True)
    d.close()
    out = {}
    for name, extra in (("pos_wide", 0), ("word_pipeline", _native.WORD_PIPELINE)):
        ms = []
        for _ in range(4):
            lib = _native.load()
            import ctypes as C
            opt = z._options(_native.INPUT_ON_DEVICE | _native.SKIP_MD5 | _native.SKIP_FRAME_CRC16 | extra, 0)
            h, info = C.c_void_p(), _native.Info()
            st = lib.zflacb200_open(flac.data_ptr(), flac.numel(), C.byref(opt), C.byref(h), C.byref(info))
            assert st == 0, st
            ms.append((info.ms_decode, info.ms_epilogue))
            lib.zflacb200_close(h)
        out[name] = tuple(round(float(np.median([m[k] for m in ms[1:]])), 3) for k in (0, 1))
    print(f"partition order {po}: bytes {flac.numel()} (ms_decode, ms_epilogue) {out}", flush=True)
