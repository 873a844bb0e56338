#!/usr/bin/env python
"""Per-source-line warp-instruction / stall-sample breakdown of one kernel from an .ncu-rep (needs -lineinfo + --import-source on).
usage: ncu_lines.py report.ncu-rep kernel_regex [min_pct]   (the regex is matched against ncu's demangled function name,
e.g. 'fastCellsKernel<\\(int\\)64' ; the first matching launch is printed)"""
import csv, io, re, subprocess, sys
rep, rx = sys.argv[1], sys.argv[2]
minp = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
# the page is a sequence of blocks: "File Path", "Function Name", header ("Line No", ...), then rows
blocks, cur = [], None
for r in rows:
    if not r:
        continue
    if r[0] == "Function Name":
        cur = {"name": r[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif r[0] == "Line No" and cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and cur["hdr"] is not None and r[0] != "File Path":
        cur["rows"].append(r)
match = [b for b in blocks if re.search(rx, b["name"]) and b["hdr"]]
if not match:
    sys.exit("no kernel matches; have:\n" + "\n".join(sorted({b["name"][:120] for b in blocks})))
match = [b for b in match if b["name"] == match[0]["name"]]  # one function, one block per source file (inlined headers)
num = lambda v: int(v) if v.isdigit() else 0
allrows = []
for b in match:
    hdr = b["hdr"]
    ci, sm = hdr.index("Instructions Executed"), hdr.index("# Samples")
    allrows.append([(r[0], r[1], num(r[ci]), num(r[sm])) for r in b["rows"] if r[0].isdigit() and len(r) > ci])
tot = sum(x[2] for rows in allrows for x in rows) or 1
tots = sum(x[3] for rows in allrows for x in rows) or 1
print(f"{match[0]['name'][:100]}\ntotal warp instructions {tot}, samples {tots}")
for k, rows in enumerate(allrows):
    print(f"--- source file {k + 1}: {100 * sum(x[2] for x in rows) / tot:.1f} % of the instructions")
    for ln, src, ni, ns in rows:
        if ni > tot * minp / 100 or ns > tots * minp / 100:
            print(f"{ln:>5} inst {100*ni/tot:6.2f}%  stall-smp {100*ns/tots:6.2f}% | {src.strip()[:120]}")
