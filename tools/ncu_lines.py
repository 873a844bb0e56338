#!/usr/bin/env python
"""Per-source-line instruction / stall-sample breakdown of one kernel from an .ncu-rep (needs -lineinfo + --import-source on).
usage: ncu_lines.py report.ncu-rep kernel_regex [min_pct]"""
import csv, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass", "--kernel-name",
                      "regex:" + rx, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Line No")
ci, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
lines = [r for r in rows if r and r[0].isdigit() and len(r) > ci and r[ci].isdigit()]
tot = sum(int(r[ci]) for r in lines) or 1
num = lambda v: int(v) if v.isdigit() else 0
tots = sum(num(r[sm]) for r in lines) or 1
print(f"total warp instructions {tot}, samples {tots}")
for r in lines:
    if int(r[ci]) > tot * minp / 100 or num(r[sm]) > tots * minp / 100:
        print(f"{r[0]:>5} inst {100*int(r[ci])/tot:6.2f}%  stall-smp {100*num(r[sm])/tots:6.2f}% | {r[1].strip()[:110]}")
