#!/usr/bin/env python
"""Pinned-host <-> device copy bandwidth on this box (the bound of bench.py's e2e leg).  usage: pcie_probe.py"""
import json, time, torch
n = 92_766_720
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(n // 4, dtype=torch.uint8).pin_memory(); d2 = torch.empty(n // 4, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=10):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
a = t(lambda: d.copy_(h, non_blocking=True)); b = t(lambda: h.copy_(d, non_blocking=True))
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
c = t(both)
print(json.dumps({"h2d_GBps": n / a / 1e9, "d2h_GBps": n / b / 1e9, "h2d_with_concurrent_d2h_GBps": n / c / 1e9, "bytes": n}))
