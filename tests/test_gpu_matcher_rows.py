"""GPU parity (through the C ABI) of the matcher-side rows: accept rules, guided grid search, ComputeStereoMatches."""
import numpy as np
import pytest

from orb_slam3_ros2_b200 import synth
from test_oracle_matcher_rows import make_grid_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx as m
    m.lib()
    return m


def test_accept_rules_match_oracle_exhaustively(orbx, oracle):
    # every (d1, d2) pair of the ranges that matter, every rule, three ratios
    d1, d2 = np.meshgrid(np.arange(0, 130, dtype=np.int32), np.arange(0, 257, dtype=np.int32), indexing="ij")
    d1, d2 = d1.ravel(), d2.ravel()
    rng = np.random.default_rng(0)
    l1, l2 = rng.integers(0, 3, d1.size), rng.integers(0, 3, d1.size)
    idx2 = np.where(d2 == 256, -1, 5).astype(np.int64)
    for rule in range(8):  # 7 = TH_LOW * ratioHamming (ORBmatcher.cc:523, :636): the ratio argument carries ratioHamming
        for nn in (0.6, 0.7, 0.9, 0.75, 1.0, 1.5):
            got = orbx.accept_top2(rule, d1, d2, nn, idx2=idx2, level1=l1, level2=l2)
            want = oracle.accept_top2(rule, d1, d2, nn, idx2=idx2, level1=l1, level2=l2)
            assert np.array_equal(got, want), (rule, nn)
    assert np.array_equal(orbx.accept_top2(1, d1, d2, 0.7), oracle.accept_top2(1, d1, d2, 0.7))
    assert len(orbx.accept_top2(1, d1[:0], d2[:0])) == 0


@pytest.mark.parametrize("seed,nt,nq,filters", [(1, 600, 200, False), (2, 2000, 1500, True), (5, 1, 10, False), (6, 5000, 3000, True)])
def test_grid_search_matches_oracle(orbx, oracle, seed, nt, nq, filters):
    kps, desc, bounds, qs, dq, skip, ur = make_grid_case(seed, nt=max(nt, 80), nq=nq)
    kps, desc, skip, ur = kps[:nt], desc[:nt], skip[:nt], ur[:nt]
    if not filters:
        skip = ur = None
    got = orbx.search_grid_top2(kps, desc, bounds, qs, dq, skip, ur)
    want = oracle.search_grid_top2(kps, desc, bounds, qs, dq, skip, ur)
    for g, w, name in zip(got, want, ["idx1", "d1", "idx2", "d2", "lvl1", "lvl2"]):
        assert np.array_equal(g, w), name


def test_grid_search_on_extracted_frames(orbx, oracle):
    """SearchForInitialization-shaped use: level-0 keypoints of frame 0 searched in frame 1 with windowSize 100."""
    ex = orbx.ORBextractor(2000, 1.2, 8, 20, 7)
    _, k0, d0 = ex(synth.frame(0, 752, 480), None, (0, 1000))
    _, k1, d1 = ex(synth.frame(1, 752, 480), None, (0, 1000))
    sel = k0["octave"] == 0
    qs = np.zeros(int(sel.sum()), dtype=orbx.GRID_QUERY_DTYPE)
    qs["u"], qs["v"], qs["r"] = k0["x"][sel], k0["y"][sel], 100.0
    qs["min_level"], qs["max_level"] = 0, 0
    bounds = (0.0, 752.0, 0.0, 480.0)
    got = orbx.search_grid_top2(k1, d1, bounds, qs, d0[sel])
    want = oracle.search_grid_top2(k1, d1, bounds, qs, d0[sel])
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    acc = orbx.accept_top2(orbx.ACCEPT_INIT, got[1], got[3], 0.9)
    assert acc.sum() > 100
    # frame 1 is frame 0 shifted by (3, 1) px: accepted matches must agree with that motion
    dx = k0["x"][sel][acc] - k1["x"][got[0][acc]]
    assert abs(np.median(dx) - 3.0) < 0.5


@pytest.mark.parametrize("w,h,nf", [(1241, 376, 2000), (640, 360, 1000)])
def test_compute_stereo_matches_bit_exact(orbx, oracle, w, h, nf):
    exl, exr = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    iml, imr = synth.frame(0, w, h), synth.frame(0, w, h, right=True)
    _, kl, dl = exl(iml, None, (0, 0))
    _, kr, dr = exr(imr, None, (0, 0))
    bf, b = 386.1448, 0.53716  # config/Stereo/KITTI00-02.yaml
    ur, dp = orbx.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, b)
    ol, orr = oracle.PortExtractor(nfeatures=nf), oracle.PortExtractor(nfeatures=nf)
    ol.extract(iml, (0, 0)); orr.extract(imr, (0, 0))
    t = ol.tables()
    wu, wd = oracle.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], [ol.pyramid(l) for l in range(8)],
                                           [orr.pyramid(l) for l in range(8)], bf, b)
    assert np.array_equal(ur.view(np.uint32), wu.view(np.uint32))  # bit-exact floats
    assert np.array_equal(dp.view(np.uint32), wd.view(np.uint32))
    assert (ur >= 0).sum() > 300
    # empty right side
    ur0, dp0 = orbx.compute_stereo_matches(exl, exr, kl, dl, kr[:0], dr[:0], bf, b)
    assert (ur0 == -1).all() and (dp0 == -1).all()


def test_compute_stereo_matches_batch_equals_per_pair_and_oracle(orbx, oracle):
    """Config 3: KITTI-shape pairs in one call, everything read on the device == the per-pair call == the oracle."""
    w, h, nf, F = 1241, 376, 2000, 12
    exl, exr = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    exl.set_subbatches(2)
    left, right = synth.frames(0, F, w, h), synth.frames(0, F, w, h, right=True)
    exl.extract_batch(left, (0, 0), download=False)
    exr.extract_batch(right, (0, 0), download=False)
    bf, b = 386.1448, 0.53716
    ur, dp = orbx.compute_stereo_matches_batch(exl, exr, bf, b)
    assert ur.shape == (F, exl.capacity)
    t = None
    for f in range(F):
        _, kl, dl = exl.download(f)
        _, kr, dr = exr.download(f)
        u1, d1 = orbx.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, b, frame=f)
        n = len(kl)
        assert np.array_equal(ur[f, :n].view(np.uint32), u1.view(np.uint32)), f
        assert np.array_equal(dp[f, :n].view(np.uint32), d1.view(np.uint32)), f
        if f in (0, F - 1):
            ol, orr = oracle.PortExtractor(nfeatures=nf), oracle.PortExtractor(nfeatures=nf)
            ol.extract(left[f], (0, 0)); orr.extract(right[f], (0, 0))
            t = ol.tables()
            wu, wd = oracle.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], [ol.pyramid(l) for l in range(8)],
                                                   [orr.pyramid(l) for l in range(8)], bf, b)
            assert np.array_equal(ur[f, :n].view(np.uint32), wu.view(np.uint32))
            assert np.array_equal(dp[f, :n].view(np.uint32), wd.view(np.uint32))
            # the synthetic right image is the left one shifted by 16 px: matched disparities sit there
            m = ur[f, :n] >= 0
            assert m.sum() > 300 and abs(np.median(kl["x"][m] - ur[f, :n][m]) - 16.0) < 1.0


def test_extract_color_staging_matches_cv2_goldens_and_oracle(orbx, oracle):
    """orbx_extract_color: resize + cvtColor fused ahead of operator() (System.cc:420-425, Tracking.cc:1569-1582)."""
    from conftest import load_golden
    g = load_golden("stage.npz")
    ex = orbx.ORBextractor(300, 1.2, 8, 20, 7)
    names = sorted(k[:-5] for k in g.files if k.endswith("_meta"))
    for name in names:
        ch, rgb, nw, nh = [int(v) for v in g[name + "_meta"]]
        src = g[name + "_src"]
        if name == "cube":
            n = src.shape[1] // 97 * 97
            src = np.ascontiguousarray(src.reshape(-1, 3)[:n].reshape(-1, 97, 3))  # a 2-D image of the lattice
            want = oracle.stage_gray(src, True, None)
            assert np.array_equal(want.ravel(), g["cube_gray"].ravel()[:want.size])
        else:
            want = g[name + "_gray"]
        ex.extract_color(src, bool(rgb), (nw, nh) if nw else None)
        assert np.array_equal(ex.staged_gray(), want), name
    # end to end on a real frame: colour frame whose grey conversion is known, resized 2x down
    base = synth.frame(0, 1504, 960)
    rgb = np.stack([base, base, base], axis=2)  # R = G = B = v  ->  grey = v exactly (weights sum to 2^15)
    mono, kp, desc = ex.extract_color(rgb, True, (752, 480))
    gray = oracle.stage_gray(rgb, True, (752, 480))
    assert np.array_equal(ex.staged_gray(), gray)
    wm, wk, wd = oracle.PortExtractor(nfeatures=300).extract(gray, (0, 1000))
    assert mono == wm and np.array_equal(kp, wk) and np.array_equal(desc, wd)


@pytest.mark.parametrize("name", [c[0] for c in synth.UNDISTORT_CASES])
def test_undistort_keypoints_and_image_bounds_bit_exact(orbx, oracle, name):
    """Frame::UndistortKeyPoints / ComputeImageBounds (Frame.cc:747-809) on the device == oracle == cv2 golden, float bits."""
    import os
    from test_oracle_refslices import undistort_keypoint_case
    w, h, K, dist, kps = undistort_keypoint_case(name)
    got = orbx.undistort_keypoints(kps, K, dist)
    want = oracle.undistort_keypoints(kps, K, dist)
    assert got.tobytes() == want.tobytes()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "prim_undistort.npz"))
    if dist[0] != 0:
        assert np.array_equal(np.stack([got["x"], got["y"]], 1).view(np.uint32), g[f"{name}__out"].view(np.uint32))
    else:
        assert got.tobytes() == kps.tobytes()
    assert orbx.image_bounds(w, h, K, dist).tobytes() == oracle.image_bounds(w, h, K, dist).tobytes()
    assert len(orbx.undistort_keypoints(kps[:0], K, dist)) == 0
    with pytest.raises(orbx.OrbxError):  # zero focal length, 3 coefficients: rejected, as cv::undistortPoints asserts
        orbx.undistort_keypoints(kps[:4], np.zeros((3, 3), np.float32), dist)
    with pytest.raises(orbx.OrbxError):
        orbx.undistort_keypoints(kps[:4], K, np.array([0.1, 0.2, 0.3], np.float32))
    # a different projection matrix P than K (cv::undistortPoints' last argument)
    P = K.copy(); P[0, 0] *= 0.9; P[1, 1] *= 1.1; P[0, 2] += 3.5
    assert orbx.undistort_keypoints(kps[:500], K, dist, P).tobytes() == oracle.undistort_keypoints(kps[:500], K, dist, P).tobytes()


def test_undistort_batch_on_device_resident_keypoints(orbx, oracle):
    """orbx_undistort_batch: mvKeysUn of a whole extracted batch without leaving the device == per-frame oracle, then the
    guided search runs on the undistorted keypoints inside the undistorted image bounds (Frame.cc:323-367 order)."""
    _, w, h, fx, fy, cx, cy, dist = synth.UNDISTORT_CASES[0]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    dist = np.array(dist, np.float32)
    frames = np.stack([synth.frame(k, w, h) for k in range(6)])
    e = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    mono, n, kps, desc = e.extract_batch(frames, (0, 1000))
    un = orbx.undistort_batch(e, K, dist)
    for f in range(len(frames)):
        assert un[f, :n[f]].tobytes() == oracle.undistort_keypoints(kps[f, :n[f]], K, dist).tobytes(), f
    assert orbx.undistort_batch(e, K, dist, download=False) != 0
    zero = np.zeros(4, np.float32)
    assert orbx.undistort_batch(e, K, zero)[0, :n[0]].tobytes() == kps[0, :n[0]].tobytes()  # Frame.cc:749-753
    bounds = orbx.image_bounds(w, h, K, dist)
    qs = np.zeros(int(n[0]), dtype=orbx.GRID_QUERY_DTYPE)
    qs["u"], qs["v"], qs["r"] = un[0, :n[0]]["x"], un[0, :n[0]]["y"], 100.0
    qs["min_level"] = qs["max_level"] = kps[0, :n[0]]["octave"]
    got = orbx.search_grid_top2(un[1, :n[1]], desc[1, :n[1]], bounds, qs, desc[0, :n[0]])
    want = oracle.search_grid_top2(un[1, :n[1]], desc[1, :n[1]], bounds, qs, desc[0, :n[0]])
    for g_, w_ in zip(got, want):
        assert np.array_equal(g_, w_)
    e.close()


@pytest.mark.parametrize("nf,shift,window,nnratio,ori,corrupt", [(1000, 1, 100, 0.9, True, 0), (1000, 2, 100, 0.9, False, 0),
                                                                  (2000, 1, 30, 0.7, True, 3), (5000, 3, 100, 0.9, True, 5),
                                                                  (1000, 5, 10, 0.9, True, 0), (1000, 1, 1000, 0.9, True, 7)])
def test_search_for_initialization_bit_exact(orbx, oracle, nf, shift, window, nnratio, ori, corrupt):
    """ORBmatcher::SearchForInitialization, whole function (ORBmatcher.cc:648-763), on the device == oracle (itself pinned to the
    reference's own body): vnMatches12, the return value and the updated vbPrevMatched, over two rounds."""
    from test_oracle_refslices import init_pair_case
    k1, d1, k2, d2, bounds, prev = init_pair_case(oracle, nf, shift, corrupt=corrupt)
    for _ in range(2):
        got = orbx.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, nnratio, ori)
        want = oracle.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, nnratio, ori)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
        prev = got[2]
    if corrupt:
        assert oracle.port_lib().orbo_search_for_initialization_withdrawn() >= 0
    # degenerate inputs: empty second frame, empty first frame
    n, m12, _ = orbx.search_for_initialization(k1, d1, k2[:0], d2[:0], bounds, prev, window, nnratio, ori)
    assert n == 0 and (m12 == -1).all()
    assert orbx.search_for_initialization(k1[:0], d1[:0], k2, d2, bounds, prev[:0], window, nnratio, ori)[0] == 0


def test_search_for_initialization_batch_equals_per_pair(orbx, oracle):
    """Consecutive pairs of a device-resident batch (on the undistorted keypoints) == the per-pair oracle."""
    _, w, h, fx, fy, cx, cy, dist = synth.UNDISTORT_CASES[0]
    K = np.array([[fx, 0, cx], [0, fy, cy], [0, 0, 1]], np.float32)
    dist = np.array(dist, np.float32)
    frames = np.stack([synth.frame(k, w, h) for k in range(5)])
    e = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    mono, n, kps, desc = e.extract_batch(frames, (0, 1000))
    for und in (False, True):
        pts = orbx.undistort_batch(e, K, dist) if und else kps
        bounds = orbx.image_bounds(w, h, K, dist) if und else (0.0, float(w), 0.0, float(h))
        nm, m12, prev = orbx.search_for_initialization_batch(e, bounds, 100, 0.9, True, use_undistorted=und)
        assert len(nm) == len(frames) - 1
        for p in range(len(frames) - 1):
            a, b = pts[p, :n[p]], pts[p + 1, :n[p + 1]]
            p0 = np.stack([a["x"], a["y"]], 1)
            want = oracle.search_for_initialization(a, desc[p, :n[p]], b, desc[p + 1, :n[p + 1]], bounds, p0, 100, 0.9, True)
            assert nm[p] == want[0] and np.array_equal(m12[p, :n[p]], want[1]), (und, p)
            assert prev[p, :n[p]].tobytes() == want[2].tobytes()
            assert nm[p] > 50
    # a new extraction invalidates the undistorted keypoints of the previous batch
    e.extract_batch(frames[:2], (0, 1000))
    with pytest.raises(orbx.OrbxError):
        orbx.search_for_initialization_batch(e, bounds, 100, 0.9, True, use_undistorted=True)
    e.close()


@pytest.mark.parametrize("kw", [dict(k=0, seed=0), dict(k=3, seed=1, nf=2000, w=1241, h=376), dict(k=1, seed=2, mono=False, stereo=True, dz=0.3),
                                dict(k=2, seed=3, mono=False, stereo=True, dz=-0.3), dict(k=4, seed=4, rot=0.35), dict(k=5, seed=5, taken_frac=0.5),
                                dict(k=6, seed=6, w=640, h=360, nf=1250, mono=False, stereo=True)])
def test_search_by_projection_frame_frame_matches_oracle(orbx, oracle, kw):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono), whole function on the device (k_project.cu), against the
    restatement that tests/test_oracle_refslices.py pins to the reference's own body: nmatches and every current feature's map point."""
    from test_oracle_matcher_rows import make_projection_case
    c = make_projection_case(oracle, **kw)
    for ori in (True, False):
        want = oracle.search_by_projection_ff(**dict(c, check_orientation=ori))
        got = orbx.search_by_projection([c], c["cam"], c["bounds"], c["scale_factors"], c["th"], c["mono"], c["mbf"], c["mb"], ori)[0]
        assert got[0] == want[0], (kw, ori, got[0], want[0])
        assert np.array_equal(got[1], want[1]), (kw, ori, np.flatnonzero(got[1] != want[1])[:10])


def test_search_by_projection_batch_and_crowded_windows(orbx, oracle):
    """Several pairs in one call (one warp per pair in the order-dependent walk) and windows with more candidates than the per-feature
    list holds (th = 60: the walk re-enumerates the window in place)."""
    from test_oracle_matcher_rows import make_projection_case
    cases = [make_projection_case(oracle, k=k, seed=10 + k, nf=nf) for k, nf in ((0, 1000), (1, 500), (2, 1500), (3, 1000))]
    c0 = cases[0]
    for th in (15, 60):
        got = orbx.search_by_projection(cases, c0["cam"], c0["bounds"], c0["scale_factors"], th, True)
        for c, g in zip(cases, got):
            want = oracle.search_by_projection_ff(**dict(c, th=th))
            assert g[0] == want[0] and np.array_equal(g[1], want[1]), th
    # argument errors: an octave outside the scale table, counts above the capacity
    bad = dict(c0, kps_last=c0["kps_last"].copy())
    bad["kps_last"]["octave"][0] = 9
    with pytest.raises(orbx.OrbxError):
        orbx.search_by_projection([bad], c0["cam"], c0["bounds"], c0["scale_factors"], 15, True)


@pytest.mark.parametrize("kw,th,far", [(dict(k=0, seed=0), 3.0, False), (dict(k=1, seed=1, stereo=True), 3.0, False), (dict(k=2, seed=2, nf=2000, w=1241, h=376, stereo=True), 5.0, True),
                                      (dict(k=3, seed=3, taken_frac=0.7), 1.0, False), (dict(k=4, seed=4, w=640, h=360, nf=1250), 15.0, True)])
def test_search_by_projection_map_points_matches_oracle(orbx, oracle, kw, th, far):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) (Tracking::SearchLocalPoints), whole function on the
    device, against the restatement pinned to the reference's own body; th = 15 makes windows longer than the per-point list."""
    from test_oracle_matcher_rows import make_local_points_case
    c = make_local_points_case(oracle, **kw)
    for nn in (0.8, 0.6):
        want = oracle.search_by_projection_map(**c, th=th, far_points=far, th_far_points=30.0, nnratio=nn)
        got = orbx.search_by_projection_map([c], c["bounds"], c["scale_factors"], th, far, 30.0, nn)[0]
        assert got[0] == want[0], (kw, th, nn, got[0], want[0])
        assert np.array_equal(got[1], want[1]), (kw, th, nn, np.flatnonzero(got[1] != want[1])[:10])


def test_search_by_projection_map_points_batch(orbx, oracle):
    from test_oracle_matcher_rows import make_local_points_case
    cases = [make_local_points_case(oracle, k=k, seed=20 + k, nf=nf) for k, nf in ((0, 1000), (1, 600), (2, 1400))]
    got = orbx.search_by_projection_map(cases, cases[0]["bounds"], cases[0]["scale_factors"], 3.0)
    for c, g in zip(cases, got):
        want = oracle.search_by_projection_map(**c, th=3.0)
        assert g[0] == want[0] and np.array_equal(g[1], want[1])
