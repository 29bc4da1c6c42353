#!/usr/bin/env python3
"""What the PCIe link of this box gives: H2D alone, D2H alone, both at once (pinned host memory, two streams)."""
import time
import torch

n_in, n_out = 284 << 20, 635 << 20
h_in = torch.empty(n_in, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n_out, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n_in, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(do_in, do_out, chunks=1):
    torch.cuda.synchronize()
    t = time.perf_counter()
    for c in range(chunks):
        if do_in:
            with torch.cuda.stream(s1):
                a, b = c * n_in // chunks, (c + 1) * n_in // chunks
                d_in[a:b].copy_(h_in[a:b], non_blocking=True)
        if do_out:
            with torch.cuda.stream(s2):
                a, b = c * n_out // chunks, (c + 1) * n_out // chunks
                h_out[a:b].copy_(d_out[a:b], non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t) * 1e3


for name, args in (("h2d 284MB", (True, False)), ("d2h 635MB", (False, True)), ("both", (True, True)), ("both, 24 chunks", (True, True, 24))):
    run(*args)
    ms = min(run(*args) for _ in range(5))
    print(f"{name}: {ms:.2f} ms", flush=True)
