#!/usr/bin/env python
"""profiles/traffic_r01.json (what bench.py's roofline.traffic reads) from an ncu_summary.py text of the 256-frame capture:
DRAM bytes read + written per extraction stage, first pipeline pass.  usage: ncu_traffic.py summary.txt > traffic.json"""
import json, sys
STAGE = {"pyrLevel0AlignedKernel": "pyramid", "pyrLevel0Kernel": "pyramid", "pyrResizeKernel": "pyramid", "fastCellsKernel": "fast_cells",
         "blurKernel": "blur", "blurKernelLd": "blur", "quadtreeKernel": "quadtree", "slotKernel": "describe", "describeKernel": "describe"}
rows = [l.split() for l in open(sys.argv[1]).read().splitlines()[1:]]
agg, seen = {}, 0
for r in rows:
    name = r[0].split("<")[0]
    if name.startswith("pyrLevel0"):
        seen += 1
        if seen > 1:
            break
    if seen == 0 or name not in STAGE:
        continue
    i = 1
    while True:
        try:
            float(r[i]); break
        except ValueError:
            i += 1
    a = agg.setdefault(STAGE[name], {"read": 0.0, "write": 0.0})
    a["read"] += float(r[i + 1]) * 1e6
    a["write"] += float(r[i + 2]) * 1e6
out = {"source": f"{sys.argv[1]} (ncu --set full --clock-control none, bench.py --steps 1 --warmup 1 --skip-secondary --subbatches 1 "
                 "--resident-handles 1, 256 frames 752x480)",
       "unit": "bytes per stage per 256-frame batch, dram__bytes_read.sum + dram__bytes_write.sum", "frames": 256}
for k, v in agg.items():
    v["total"] = v["read"] + v["write"]
    out[k] = v
print(json.dumps(out, indent=1))
