#!/usr/bin/env python
"""One line per kernel launch of an .ncu-rep: time, DRAM bytes, issue %, occupancy, pipes.  usage: ncu_summary.py rep"""
import csv, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = rows[0]; idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "us"), ("dram__bytes_read.sum", "rdMB"), ("dram__bytes_write.sum", "wrMB"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu%"), ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "lsu%"), ("smsp__inst_executed.sum", "winst"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "bankconf"), ("launch__registers_per_thread", "regs"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram%"), ("lts__t_sector_hit_rate.pct", "l2hit%")]
units = rows[1]
names = [n for _, n in cols]
tscale = {"ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6}.get(units[idx["gpu__time_duration.sum"]], 1.0)  # -> us
print("kernel".ljust(28), " ".join(n.rjust(9) for n in names), " grid")
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0].split("::")[-1].replace(" ", "")[:28]
    vals = []
    for c, _ in cols:
        v = r[idx[c]] if c in idx else ""
        try:
            x = float(v) * (tscale if c == "gpu__time_duration.sum" else 1.0)
            vals.append(f"{x:9.2f}" if x < 1e6 else f"{x:9.3g}")
        except ValueError: vals.append(v.rjust(9))
    print(name.ljust(28), " ".join(vals), " ", r[idx["Grid Size"]])
