"""oracle/minicv (restated OpenCV primitives) against cv2 4.13.0 golden vectors (tools/gen_golden.py).

The reference has no tests or fixtures for this path, so these goldens — produced by
the real OpenCV build the survey declared as the oracle version — are what pins the oracle."""
import numpy as np
import pytest

from conftest import load_golden


def level_sizes(w, h, nlevels=8, sf=1.2):
    s = np.float32(1.0)
    out = []
    for _ in range(nlevels):
        inv = np.float32(1.0) / s
        out.append((int(np.rint(np.float32(w) * inv)), int(np.rint(np.float32(h) * inv))))
        s = np.float32(np.float64(s) * np.float64(np.float32(sf)))
    return out


def test_resize_chain_matches_cv2(oracle):
    g = load_golden("prim_resize.npz")
    names = sorted({k.rsplit("_", 1)[0] for k in g.files})
    assert names
    for name in names:
        prev = g[f"{name}_src"]
        sizes = level_sizes(prev.shape[1], prev.shape[0])
        for l in range(1, 5):
            want = g[f"{name}_L{l}"]
            assert want.shape == (sizes[l][1], sizes[l][0])
            got = oracle.resize_linear(prev, sizes[l][0], sizes[l][1])
            assert np.array_equal(got, want), (name, l)
            prev = got


def test_border_reflect101_matches_cv2(oracle):
    g = load_golden("prim_border.npz")
    for k in [k for k in g.files if k.endswith("_src")]:
        assert np.array_equal(oracle.border101(g[k], 19), g[k.replace("_src", "_b19")]), k


def test_fast_matches_cv2_set_order_response(oracle):
    g = load_golden("prim_fast.npz")
    checked = 0
    for k in [k for k in g.files if k.endswith("_src")]:
        for th in (20, 7, 0, 60):
            want = g[k.replace("_src", f"_t{th}")]
            got = oracle.fast(g[k], th, True)
            assert np.array_equal(got, want.reshape(-1, 3)), (k, th, len(got), len(want))
            checked += len(want)
    assert checked > 1000


def test_gaussian_blur_matches_cv2(oracle):
    g = load_golden("prim_blur.npz")
    for k in [k for k in g.files if k.endswith("_src")]:
        assert np.array_equal(oracle.gauss7(g[k]), g[k.replace("_src", "_g7")]), k


def test_fast_atan2_bit_exact(oracle):
    g = load_golden("prim_atan2.npz")
    got = oracle.fast_atan2(g["y"], g["x"])
    assert np.array_equal(got.view(np.uint32), g["angle"].view(np.uint32))


def test_undistort_points_bit_exact(oracle):
    """minicv's cv::undistortPoints (the primitive inside Frame::UndistortKeyPoints, Frame.cc:766) == cv2 4.13.0, float bits."""
    from orb_slam3_ros2_b200 import synth
    g = load_golden("prim_undistort.npz")
    for name, *_ in synth.UNDISTORT_CASES:
        _, _, K, dist, pts = synth.undistort_case(name)
        assert np.array_equal(pts, g[f"{name}__pts"]), name
        got = oracle.undistort_points(pts, K, dist)
        assert np.array_equal(got.view(np.uint32), g[f"{name}__out"].view(np.uint32)), name


def test_knn_top2_order_matches_bfmatcher(oracle):
    g = load_golden("match.npz")
    i1, d1, i2, d2 = oracle.match_top2(g["q"], g["t"])
    assert np.array_equal(i1, g["idx"][:, 0]) and np.array_equal(i2, g["idx"][:, 1])
    assert np.array_equal(d1, g["dist"][:, 0].astype(np.int32)) and np.array_equal(d2, g["dist"][:, 1].astype(np.int32))
    assert d1[0] == 0 and d2[0] == 0 and (i1[0], i2[0]) == (10, 200)  # planted duplicates: lowest index first
    # single train row: only one neighbour exists (knnMatch returns 1 element; Frame.cc:1151 skips it)
    i1, d1, i2, d2 = oracle.match_top2(g["q"], g["t"][:1])
    assert (i1 == 0).all() and (i2 == -1).all() and (d2 == 256).all()


def test_descriptor_distance_swar(oracle):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (200, 32), dtype=np.uint8)
    for x, y in zip(a, b):
        assert oracle.hamming(x, y) == int(np.unpackbits(x ^ y).sum())
    assert oracle.hamming(a[0], a[0]) == 0 and oracle.hamming(a[0], ~a[0]) == 256


def test_stage_resize_then_gray_matches_cv2(oracle):
    """Input staging (System.cc:420-425 resize, Tracking.cc:1569-1582 cvtColor) against the cv2 4.13 goldens."""
    g = load_golden("stage.npz")
    names = sorted(k[:-5] for k in g.files if k.endswith("_meta"))
    assert len(names) >= 9
    for name in names:
        ch, rgb, nw, nh = [int(v) for v in g[name + "_meta"]]
        got = oracle.stage_gray(g[name + "_src"], bool(rgb), (nw, nh) if nw else None)
        assert np.array_equal(got, g[name + "_gray"]), name
