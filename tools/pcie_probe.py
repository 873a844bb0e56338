#!/usr/bin/env python
"""Pinned-host <-> device copy bandwidth with N ranks copying at once (the bound of bench.py's e2e leg and of its 1 -> 8 GPU curve).
    python tools/pcie_probe.py                                              # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py
Every rank copies the e2e leg's bytes per step (92.4 MB H2D, 20.1 MB D2H) on its own GPU: H2D alone, D2H alone, both at once on two
streams; all ranks start behind a barrier, the time is the max over ranks.  One JSON line with per-GPU and aggregate GB/s."""
import json, os, time
import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
H2D, D2H = 92_405_760, 20_064_128
h_in = torch.empty(H2D, dtype=torch.uint8).pin_memory(); d_in = torch.empty(H2D, dtype=torch.uint8, device=dev)
h_out = torch.empty(D2H, dtype=torch.uint8).pin_memory(); d_out = torch.empty(D2H, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)


def timed(fn, reps=10):
    fn(); torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        fn()
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    if world > 1:
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dt = float(t.item())
    return dt


def h2d():
    with torch.cuda.stream(s1):
        d_in.copy_(h_in, non_blocking=True)


def d2h():
    with torch.cuda.stream(s2):
        h_out.copy_(d_out, non_blocking=True)


def both():
    h2d(); d2h()


a, b, c = timed(h2d), timed(d2h), timed(both)
if rank == 0:
    print(json.dumps({"ranks": world, "h2d_bytes": H2D, "d2h_bytes": D2H,
                      "h2d_alone_GBps_per_gpu": H2D / a / 1e9, "d2h_alone_GBps_per_gpu": D2H / b / 1e9,
                      "both_seconds": c, "both_GBps_per_gpu": (H2D + D2H) / c / 1e9, "both_GBps_aggregate": world * (H2D + D2H) / c / 1e9,
                      "e2e_bound_frames_per_s": world * 256 / c}), flush=True)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
