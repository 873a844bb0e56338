#!/usr/bin/env python
"""The reduced workload tools/sanitize.sh runs under compute-sanitizer: every kernel family of liborbx.so once, at small sizes,
through the C ABI with host buffers only (no torch: nothing but this library's kernels and copies is checked), each result
compared with the oracle so that a tool run is also a parity run.  Prints one line per case."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from orb_slam3_ros2_b200 import orbx, synth  # noqa: E402
from oracle import pyoracle  # noqa: E402


def same(a, b):
    return len(a) == len(b) and np.array_equal(a, b)


GUARD = os.environ.get("ORBX_GUARD") == "1"


def guards(*handles):
    if GUARD:
        for h in handles:
            h.check_guards()  # raises if a kernel wrote behind a device buffer


def main():
    # 1. single frames: strip FAST + minThFAST redo (low contrast) + overflow hand-over to the per-cell kernel (white noise)
    rng = np.random.default_rng(3)
    cases = [("textured 320x240", synth.frame(0, 320, 240), 500), ("white noise 200x160", rng.integers(0, 256, (160, 200), dtype=np.uint8), 300),
             ("low contrast 300x220", ((synth.frame(3, 300, 220).astype(np.int32) - 128) // 6 + 128).astype(np.uint8), 400),
             ("textured 752x480", synth.frame(1, 752, 480), 1000)]
    for name, im, nf in cases:
        ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
        mono, kp, desc = ex(im, None, (0, 1000))
        pm, pk, pd = pyoracle.PortExtractor(nfeatures=nf).extract(im, (0, 1000))
        assert mono == pm and same(kp, pk) and same(desc, pd), name
        guards(ex)
        print(f"extract {name}: {len(kp)} keypoints == oracle", flush=True)
        ex.close()
    # 2. a small batch through host buffers, sub-batches on, then frame-to-frame matching
    ims = synth.frames(0, 8, 320, 240)
    ex = orbx.ORBextractor(500, 1.2, 8, 20, 7)
    mono, n, kps, desc = ex.extract_batch(ims, (0, 1000))
    for f in range(8):
        pm, pk, pd = pyoracle.PortExtractor(nfeatures=500).extract(ims[f], (0, 1000))
        assert n[f] == len(pk) and same(kps[f, :n[f]], pk) and same(desc[f, :n[f]], pd), f
    m = ex.match_consecutive()
    w = pyoracle.match_top2(desc[0, :n[0]], desc[1, :n[1]])
    assert np.array_equal(np.asarray(m[0][0][:n[0]], np.int64), np.asarray(w[0], np.int64)), "consecutive match idx1"
    guards(ex)
    print(f"batch of 8 + consecutive top-2: == oracle ({int(n.sum())} keypoints)", flush=True)
    # the same batch ten more times, sub-batches on and off: the order-free lists and atomics of the strip FAST, the quadtree's
    # rounds and the persistent describe loop must give the same bytes every time (a race shows up as a difference)
    for rep in range(10):
        ex.set_subbatches(2 if rep % 2 else 1)
        mono2, n2, kps2, desc2 = ex.extract_batch(ims, (0, 1000))
        assert np.array_equal(mono, mono2) and np.array_equal(n, n2) and np.array_equal(kps, kps2) and np.array_equal(desc, desc2), rep
    guards(ex)
    print("batch repeated 10 x (sub-batches 1 / 2): bit-identical", flush=True)
    # 3. brute-force top-2, ties and nd < 2
    q = rng.integers(0, 256, (70, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (1500, 32), dtype=np.uint8)
    db[7] = db[900] = q[3]
    for d in (db, db[:1]):
        got, want = orbx.match_top2(q, d), pyoracle.match_top2(q, d)
        assert all(np.array_equal(np.asarray(g, np.int64), np.asarray(x, np.int64)) for g, x in zip(got, want))
    print("match_top2 70 x 1500 (+ nd = 1): == oracle", flush=True)
    # 4. accept rules, every rule
    d1, d2 = [a.ravel().astype(np.int32) for a in np.meshgrid(np.arange(0, 130), np.arange(0, 257, 3), indexing="ij")]
    for rule in range(8):
        assert np.array_equal(orbx.accept_top2(rule, d1, d2, 0.7), pyoracle.accept_top2(rule, d1, d2, 0.7)), rule
    print("accept rules 0..7: == oracle", flush=True)
    # 5. stereo: left / right extraction + the whole ComputeStereoMatches on the device-resident pyramids
    exl, exr = orbx.ORBextractor(600, 1.2, 8, 20, 7), orbx.ORBextractor(600, 1.2, 8, 20, 7)
    il, ir = synth.frame(2, 480, 200), synth.frame(2, 480, 200, right=True)
    _, kl, dl = exl(il, None, (0, 0))
    _, kr, dr = exr(ir, None, (0, 0))
    ur, depth = orbx.compute_stereo_matches(exl, exr, kl, dl, kr, dr, 386.1448, 0.53716)
    ol, orr = pyoracle.PortExtractor(nfeatures=600), pyoracle.PortExtractor(nfeatures=600)
    ol.extract(il, (0, 0)); orr.extract(ir, (0, 0))
    t = ol.tables()
    wur, wdepth = pyoracle.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], [ol.pyramid(l) for l in range(8)],
                                                  [orr.pyramid(l) for l in range(8)], 386.1448, 0.53716)[:2]
    assert np.array_equal(ur, wur) and np.array_equal(depth, wdepth)
    guards(exl, exr)
    print(f"ComputeStereoMatches: {int((depth > 0).sum())} matches == oracle", flush=True)
    print("sanitize_case ok", flush=True)


if __name__ == "__main__":
    main()
