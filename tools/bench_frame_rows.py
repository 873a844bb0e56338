#!/usr/bin/env python
"""Timing of the Frame-side rows on a device-resident batch (256 frames 752x480, 1000 features, EuRoC pinhole camera):
  undistort   Frame::UndistortKeyPoints of every frame (orbx_undistort_batch, result left on the device)
  init        ORBmatcher::SearchForInitialization of every consecutive pair (orbx_search_for_initialization_batch,
              window 100, nnratio 0.9, orientation check), vnMatches12 / counts / vbPrevMatched downloaded
next to the oracle (CPU port, one thread) on a sample of the same batch.  Host wall clock around the C-ABI calls after
3 warm-up calls, median of 10.  One JSON line.  usage: python tools/bench_frame_rows.py [frames]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def median_ms(fn, warm=3, reps=10):
    for _ in range(warm):
        fn()
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        fn()
        t.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(t))


def main():
    import torch
    from orb_slam3_ros2_b200 import orbx, synth
    from oracle import pyoracle as po
    F = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    _, w, h, fx, fy, cx, cy, dist = synth.UNDISTORT_CASES[0]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    dist = np.array(dist, np.float32)
    dev = torch.device("cuda", 0)
    imgs = torch.empty((F, h, w), dtype=torch.uint8, device=dev)
    orbx._check(orbx.lib().orbx_synth_frames_device(imgs.data_ptr(), 0, F, w, h, 42, 0, 0))
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    ex.extract_batch_device(imgs.data_ptr(), F, w, h, (0, 1000))
    ex.sync()
    bounds = orbx.image_bounds(w, h, K, dist)

    def undistort():
        orbx.undistort_batch(ex, K, dist, download=False)
        ex.sync()

    out = {}

    def init():
        out["r"] = orbx.search_for_initialization_batch(ex, bounds, 100, 0.9, True, use_undistorted=True)

    t_un = median_ms(undistort)
    t_init = median_ms(init)
    nm, m12, prev = out["r"]
    # oracle on a sample of pairs
    sample = min(16, F - 1)
    n, _ = ex.counts()
    un = orbx.undistort_batch(ex, K, dist)
    kd = [ex.download(f) for f in range(sample + 1)]
    t0 = time.perf_counter()
    equal = True
    for p in range(sample):
        a, b = un[p, :n[p]], un[p + 1, :n[p + 1]]
        want = po.search_for_initialization(a, kd[p][2], b, kd[p + 1][2], bounds, np.stack([a["x"], a["y"]], 1), 100, 0.9, True)
        equal = bool(equal and want[0] == nm[p] and np.array_equal(want[1], m12[p, :n[p]]))
    cpu_ms = (time.perf_counter() - t0) * 1e3 / sample
    print(json.dumps({"workload": f"{F} frames {w}x{h}, 1000 features, EuRoC pinhole camera", "undistort_batch_ms": t_un,
                      "undistort_keypoints_per_s": float(n.sum()) / (t_un * 1e-3),
                      "search_for_initialization_batch_ms": t_init, "pairs_per_s": (F - 1) / (t_init * 1e-3),
                      "matches_per_pair": float(nm.mean()), "oracle_ms_per_pair_1thread": cpu_ms, "sample_pairs": sample,
                      "equal_to_oracle": equal, "timing": "host wall clock around the C-ABI call, median of 10"}))


if __name__ == "__main__":
    main()
