#!/usr/bin/env python
"""profiles/traffic_rNN.json (what bench.py's roofline.traffic and issue_roofline read) from an ncu_summary.py text of the
256-frame capture: DRAM bytes read + written and warp instructions (smsp__inst_executed.sum) per stage, first pipeline pass.
usage: ncu_traffic.py summary.txt > traffic.json"""
import json, sys
STAGE = {"pyrLevel0AlignedKernel": "pyramid", "pyrLevel0Kernel": "pyramid", "pyrResizeKernel": "pyramid", "fastCellsKernel": "fast_cells",
         "fastStripKernel": "fast_cells", "blurKernel": "blur", "blurKernelLd": "blur", "quadtreeKernel": "quadtree", "slotKernel": "describe",
         "describeKernel": "describe", "describeAsyncKernel": "describe", "matchKernel": "match", "matchConsecutiveKernel": "match", "unpackPairsKernel": "match"}
lines = open(sys.argv[1]).read().splitlines()
hdr = lines[0].split()
rows = [l.split() for l in lines[1:] if l.strip()]
agg, winst, seen = {}, {}, 0
for r in rows:
    name = r[0].split("<")[0]
    if name.startswith("pyrLevel0"):
        seen += 1
        if seen > 1:
            break
    if seen == 0 or name not in STAGE:
        continue
    i = 1
    while True:  # template arguments may add tokens before the first number
        try:
            float(r[i]); break
        except ValueError:
            i += 1
    vals = r[i:]
    col = {h: j for j, h in enumerate(hdr[1:])}  # kernel, then the numeric columns
    st = STAGE[name]
    a = agg.setdefault(st, {"read": 0.0, "write": 0.0})
    a["read"] += float(vals[col["rdMB"]]) * 1e6
    a["write"] += float(vals[col["wrMB"]]) * 1e6
    winst[st] = winst.get(st, 0.0) + float(vals[col["winst"]])
out = {"source": f"{sys.argv[1]} (ncu --set full --clock-control none, bench.py --steps 1 --warmup 1 --skip-secondary --subbatches 1 "
                 "--resident-handles 1, 256 frames 752x480)",
       "unit": "bytes per stage per 256-frame batch, dram__bytes_read.sum + dram__bytes_write.sum; winst = smsp__inst_executed.sum", "frames": 256}
for k, v in agg.items():
    v["total"] = v["read"] + v["write"]
    out[k] = v
out["winst"] = winst
print(json.dumps(out, indent=1))
