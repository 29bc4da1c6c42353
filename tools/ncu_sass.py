#!/usr/bin/env python3
"""Per-SASS-instruction executed counts and stall samples of one kernel from an ncu report (needs --import-source on).

    python tools/ncu_sass.py report.ncu-rep kernel_regex [min_samples_pct=0.3]
Prints the instructions in address order with: warp executions, stall samples (% of the kernel's), opcode text.
"""
import csv
import io
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
min_pct = float(sys.argv[3]) if len(sys.argv) > 3 else 0.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = None
for i, r in enumerate(rows):
    if r and r[0] == "Address":
        hdr = i
        break
if hdr is None:
    print(out[:2000])
    sys.exit(1)
h = rows[hdr]
ia, isrc = h.index("Address"), h.index("Source")
iex = next(i for i, x in enumerate(h) if x.startswith("Warp Stall Sampling (All"))
iinst = next(i for i, x in enumerate(h) if x.startswith("Instructions Executed"))
data = []
for r in rows[hdr + 1:]:
    if len(r) <= max(iex, iinst):
        continue
    try:
        data.append((r[ia], r[isrc], int(r[iinst] or 0), int(r[iex] or 0)))
    except ValueError:
        continue
tot = sum(d[3] for d in data) or 1
print(f"samples {tot}")
for a, src, n, smp in data:
    pct = 100.0 * smp / tot
    if n == 0 and smp == 0:
        continue
    mark = "*" if pct >= min_pct and min_pct > 0 else " "
    print(f"{a[-5:]} {n:10d} {pct:6.2f}% {mark} {src[:90]}")
