#!/usr/bin/env python
"""Top warp-stall reasons (stalled warps per issue-active cycle) of every distinct kernel in an .ncu-rep.  usage: ncu_stalls.py rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines())); H = rows[0]
st = [i for i, h in enumerate(H) if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
seen = set()
for r in rows[2:]:
    k = r[H.index("Kernel Name")].split("(")[0].split("::")[-1]
    if k in seen: continue
    seen.add(k)
    vals = []
    for i in st:
        try: vals.append((float(r[i]), H[i][len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
        except ValueError: pass
    vals.sort(reverse=True)
    print(k[:26].ljust(27), " ".join(f"{n}={v:.2f}" for v, n in vals[:6]))
