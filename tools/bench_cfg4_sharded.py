#!/usr/bin/env python
"""Config 4 of BASELINE.json: 10,000 synthetic 1920x1080 frames, 4000 features, sharded over the ranks by contiguous frame
ranges, NO collective on the data path.  Every rank generates its share on the device first (20 GB / G, resident in HBM),
then extracts it in 250-frame batches alternating over two handles; timed with CUDA events after a barrier, max over ranks.
  python tools/bench_cfg4_sharded.py                                  (1 GPU)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29544 tools/bench_cfg4_sharded.py
Rank 0 prints one JSON line."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

TOTAL, W, H, NF, CHUNK = 10_000, 1920, 1080, 4000, 250


def main():
    import torch
    import torch.distributed as dist
    from orb_slam3_ros2_b200 import orbx
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lo, hi = TOTAL * rank // world, TOTAL * (rank + 1) // world
    frames = torch.empty((hi - lo, H, W), dtype=torch.uint8, device=dev)
    for k in range(lo, hi, 1000):
        n = min(1000, hi - k)
        orbx._check(orbx.lib().orbx_synth_frames_device(frames[k - lo].data_ptr(), k, n, W, H, 42, 0, local))
    exs = [orbx.ORBextractor(NF, 1.2, 8, 20, 7, device=local) for _ in range(2)]
    for e in exs:
        e.set_subbatches(2)
    chunks = [(k, min(CHUNK, hi - lo - k)) for k in range(0, hi - lo, CHUNK)]

    def run():
        total = 0
        for i, (k, n) in enumerate(chunks):
            e = exs[i % 2]
            e.extract_batch_device(frames[k].data_ptr(), n, W, H, (0, 1000))
            total += n
        return total

    run()  # warm-up: scratch allocation, tensor maps
    for e in exs:
        e.sync()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    s0 = torch.cuda.ExternalStream(exs[0].stream_ptr(), device=dev)
    s1 = torch.cuda.ExternalStream(exs[1].stream_ptr(), device=dev)
    e0, e1, j = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True), torch.cuda.Event()
    with torch.cuda.stream(s1):
        j.record()
    s0.wait_event(j)
    with torch.cuda.stream(s0):
        e0.record()
    j0 = torch.cuda.Event()
    with torch.cuda.stream(s0):
        j0.record()
    s1.wait_event(j0)  # both handles start after e0
    done = run()
    with torch.cuda.stream(s1):
        j.record()
    s0.wait_event(j)
    with torch.cuda.stream(s0):
        e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    n_kp = float(exs[(len(chunks) - 1) % 2].counts()[0].mean())
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    if rank == 0:
        print(json.dumps({"config": "cfg4", "workload": f"{TOTAL} frames {W}x{H}, {NF} features, contiguous frame ranges over {world} GPU(s), no collective",
                          "n_gpus": world, "frames_per_s": TOTAL / (ms * 1e-3), "ms_total": ms, "frames_per_rank": hi - lo,
                          "batch": CHUNK, "keypoints_per_frame": n_kp, "timing": "CUDA events on the handles' streams, max over ranks",
                          "data": "synthetic, generated on the device before the timed region (resident in HBM)"}), flush=True)
    if world > 1:
        dist.destroy_process_group()
    assert done == hi - lo


if __name__ == "__main__":
    main()
