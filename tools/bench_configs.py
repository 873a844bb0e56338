#!/usr/bin/env python
"""Throughput of BASELINE.json's other single-GPU configs (bench.py measures config 1's shape, the metric's own):
  cfg2  D435i 640x480, 1000 features, batches of 256 frames, extraction + frame-to-frame top-2
  cfg3  KITTI 1241x376 stereo pairs, 2000 features: left + right extraction (two handles, two streams) and, on a sample
        of pairs, the whole Frame::ComputeStereoMatches on the device-resident pyramids
  cfg4  1920x1080, 4000 features, batches of 250 frames generated on the device (one GPU's share of the 10k-frame job)
Inputs resident in HBM, CUDA events on the handle's stream, 3 warm-up + 10 timed batches rotating over 2 input batches.
One JSON line per config.  usage: python tools/bench_configs.py [cfg2 cfg3 cfg4]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def timed(torch, stream, fn, warm=3, reps=10):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        e0.record()
    for i in range(reps):
        fn(i)
    with torch.cuda.stream(stream):
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    import torch
    from orb_slam3_ros2_b200 import orbx, synth
    what = sys.argv[1:] or ["cfg2", "cfg3", "cfg4"]
    dev = torch.device("cuda", 0)

    def frames_on_device(k0, F, w, h, right=0):
        t = torch.empty((F, h, w), dtype=torch.uint8, device=dev)
        orbx._check(orbx.lib().orbx_synth_frames_device(t.data_ptr(), k0, F, w, h, 42, right, 0))
        return t

    if "cfg2" in what or "cfg4" in what:
        for name, w, h, nf, F, match in (("cfg2", 640, 480, 1000, 256, True), ("cfg4", 1920, 1080, 4000, 250, False)):
            if name not in what:
                continue
            ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
            ex.set_subbatches(2)
            imgs = [frames_on_device(b * F, F, w, h) for b in range(2)]
            cap = ex.capacity
            midx = [torch.empty((F - 1, cap), dtype=torch.int32, device=dev) for _ in range(4)]
            stream = torch.cuda.ExternalStream(ex.stream_ptr(), device=dev)

            def step(i):
                ex.extract_batch_device(imgs[i % 2].data_ptr(), F, w, h, (0, 1000))
                if match:
                    ex.match_consecutive_device(*[m.data_ptr() for m in midx])
            ms = timed(torch, stream, step)
            n = ex.counts()[0]
            print(json.dumps({"config": name, "workload": f"{w}x{h}, {nf} features, batch {F}" + (", + frame-to-frame top-2" if match else ""),
                              "frames_per_s": F / (ms * 1e-3), "ms_per_batch": ms, "keypoints_per_frame": float(n.mean())}), flush=True)
            del ex, imgs, midx

    if "cfg3" in what:
        w, h, nf, F = 1241, 376, 2000, 128
        exl, exr = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
        left = [frames_on_device(b * F, F, w, h, 0) for b in range(2)]
        right = [frames_on_device(b * F, F, w, h, 1) for b in range(2)]
        sl = torch.cuda.ExternalStream(exl.stream_ptr(), device=dev)

        def step(i):
            exl.extract_batch_device(left[i % 2].data_ptr(), F, w, h, (0, 0))
            exr.extract_batch_device(right[i % 2].data_ptr(), F, w, h, (0, 0))
            exr.sync()  # the pair is complete when both handles are (host-side join; the left handle's events time it)
        ms = timed(torch, sl, step)
        # Frame::ComputeStereoMatches per pair (KITTI00-02.yaml: fx = 718.856, bf = 386.1448)
        fx, bf = 718.856, 386.1448
        b = bf / fx
        pairs = 16
        data = [(exl.download(f), exr.download(f)) for f in range(pairs)]
        t0 = time.perf_counter()
        matched = 0
        for f, ((_, kl, dl), (_, kr, dr)) in enumerate(data):
            ur, depth = orbx.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, b, frame=f)
            matched += int((depth > 0).sum())
        dt = (time.perf_counter() - t0) / pairs
        # the whole config: both extractions + ComputeStereoMatches of all pairs in one call (results to host arrays)
        out = (np.zeros((F, exl.capacity), np.float32), np.zeros((F, exl.capacity), np.float32))

        def full(i):
            exl.extract_batch_device(left[i % 2].data_ptr(), F, w, h, (0, 0))
            exr.extract_batch_device(right[i % 2].data_ptr(), F, w, h, (0, 0))
            orbx.compute_stereo_matches_batch(exl, exr, bf, b, out=out)
        for i in range(3):
            full(i)
        t0 = time.perf_counter()
        for i in range(10):
            full(i)
        full_ms = (time.perf_counter() - t0) / 10 * 1e3
        n = exl.counts()[0]
        matched_b = float(np.mean([(out[1][f, :n[f]] > 0).sum() for f in range(F)]))
        print(json.dumps({"config": "cfg3", "workload": f"{w}x{h} stereo pairs, {nf} features, batch {F} pairs (left + right extraction)",
                          "pairs_per_s": F / (ms * 1e-3), "ms_per_batch": ms,
                          "with_stereo_matches_batch": {"pairs_per_s": F / (full_ms * 1e-3), "ms_per_batch": full_ms,
                                                        "matched_per_pair": matched_b,
                                                        "timing": "host wall clock, extraction of both sides + orbx_compute_stereo_matches_batch, uRight/depth downloaded"},
                          "stereo_matches": {"ms_per_pair_host_call": dt * 1e3, "matched_per_pair": matched / pairs,
                                             "note": "orbx_compute_stereo_matches per pair through host keypoint arrays (single-frame API)"}}),
              flush=True)


if __name__ == "__main__":
    main()
