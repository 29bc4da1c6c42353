#!/usr/bin/env python3
"""Per-source-line and per-opcode instruction breakdown of one kernel from an ncu report (needs -lineinfo + --import-source on).

    python tools/ncu_lines.py report.ncu-rep kernel_regex [top=40] [rows_per_warp_unit]
"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, kre = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
unit = float(sys.argv[4]) if len(sys.argv) > 4 else None
sel = ["--kernel-id", kre] if kre.startswith("::") else ["--kernel-name", f"regex:{kre}"]  # e.g. ::regex:k_predict:1
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", *sel, "--print-source", "cuda,sass"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "Line No"]
lines, ops = {}, collections.Counter()
for n, hi in enumerate(hdr):
    end = hdr[n + 1] - 2 if n + 1 < len(hdr) else len(rows)
    fn = rows[hi - 2][1].split("/")[-1]
    for r in rows[hi + 1:end]:
        if len(r) < 9:
            continue
        if r[0] and r[7] not in ("", "-"):
            try:
                lines[(fn, int(r[0]))] = (r[1][:100], int(r[7]), int(r[8]), int(r[6]))
            except ValueError:
                pass
        elif r[0] == "" and r[2].startswith("0x"):
            try:
                k = int(r[7])
            except ValueError:
                continue
            s = re.sub(r"^@!?U?P\d+\s+", "", r[3].strip())
            ops[s.split()[0]] += k
tot = sum(v[1] for v in lines.values()) or 1
ts = sum(v[3] for v in lines.values()) or 1
print(f"total warp instructions {tot}, samples {ts}")
for k, v in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
    per = f" {v[1] / unit:6.2f}/u" if unit else ""
    print(f"{k[0]}:{k[1]:4d} inst {v[1] / tot * 100:5.1f}%{per} lanes {v[2] / max(v[1], 1):5.1f} smp {v[3] / ts * 100:5.1f}%  {v[0]}")
ot = sum(ops.values()) or 1
print("--- opcodes")
for op, n in ops.most_common(30):
    per = f" {n / unit:6.2f}/u" if unit else ""
    print(f"{op:28s} {n / ot * 100:5.1f}%{per}")
