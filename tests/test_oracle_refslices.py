"""Pins oracle/match_oracle.cpp (the restatements the GPU parity tests compare against) to the reference's OWN function
bodies: oracle/_ref/librefslices.so holds Frame::ComputeStereoMatches, AssignFeaturesToGrid, GetFeaturesInArea, PosInGrid,
UndistortKeyPoints, ComputeImageBounds and ORBmatcher::DescriptorDistance cut out of /root/reference/ORB_SLAM3/src/{Frame,ORBmatcher}.cc by oracle/Makefile and
compiled unchanged (built where /root/reference exists; the prebuilt library travels to the GPU box)."""
import numpy as np
import pytest

from orb_slam3_ros2_b200 import synth
from test_oracle_matcher_rows import make_grid_case, make_local_points_case, make_projection_case


@pytest.fixture(scope="module")
def slices(oracle):
    if not oracle.have_refslices():
        pytest.skip("oracle/_ref/librefslices.so not built (needs /root/reference at build time)")
    return oracle


def test_thresholds_and_descriptor_distance(slices):
    from orb_slam3_ros2_b200 import orbx  # constants only; no compute call
    assert slices.ref_thresholds() == (orbx.TH_LOW, orbx.TH_HIGH) == (50, 100)
    rng = np.random.default_rng(1)
    a = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    b[:20] = a[:20]
    b[20:40] = ~a[20:40]
    for i in range(300):
        assert slices.ref_descriptor_distance(a[i], b[i]) == slices.hamming(a[i], b[i])
    assert slices.ref_descriptor_distance(a[0], b[0]) == 0 and slices.ref_descriptor_distance(a[20], b[20]) == 256


@pytest.mark.parametrize("w,h,nf,k", [(1241, 376, 2000, 0), (640, 360, 1000, 5), (752, 480, 1000, 2)])
def test_compute_stereo_matches_restatement_equals_reference_body(slices, w, h, nf, k):
    ol, orr = slices.PortExtractor(nfeatures=nf), slices.PortExtractor(nfeatures=nf)
    _, kl, dl = ol.extract(synth.frame(k, w, h), (0, 0))
    _, kr, dr = orr.extract(synth.frame(k, w, h, right=True), (0, 0))
    t = ol.tables()
    pl, pr = [ol.pyramid(l) for l in range(8)], [orr.pyramid(l) for l in range(8)]
    bf, b = 386.1448, 0.53716  # config/Stereo/KITTI00-02.yaml
    got = slices.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], pl, pr, bf, b)
    want = slices.ref_compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], pl, pr, bf, b)
    assert np.array_equal(got[0].view(np.uint32), want[0].view(np.uint32))  # mvuRight, bit for bit
    assert np.array_equal(got[1].view(np.uint32), want[1].view(np.uint32))  # mvDepth
    assert (want[0] >= 0).sum() > 200
    # a different baseline changes maxD and the depth scale
    got2 = slices.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], pl, pr, 40.0, 0.8)
    want2 = slices.ref_compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], pl, pr, 40.0, 0.8)
    assert np.array_equal(got2[0].view(np.uint32), want2[0].view(np.uint32))
    assert np.array_equal(got2[1].view(np.uint32), want2[1].view(np.uint32))


@pytest.mark.parametrize("seed,nt,nq", [(0, 600, 200), (1, 3000, 500), (2, 50, 64), (3, 5000, 300)])
def test_grid_search_restatement_equals_reference_bodies(slices, seed, nt, nq):
    kps, desc, bounds, qs, dq, skip, ur = make_grid_case(seed, nt, nq)
    got = slices.search_grid_top2(kps, desc, bounds, qs, dq)
    want = slices.ref_grid_top2(kps, desc, bounds, qs, dq)
    for g, w_ in zip(got[:4], want[:4]):
        assert np.array_equal(g, w_)
    assert want[4].max() > 5  # the windows really hold candidates


def test_grid_search_on_extracted_frames_equals_reference_bodies(slices):
    from test_oracle_matcher_rows import GQ
    e = slices.PortExtractor(nfeatures=1000)
    _, k0, d0 = e.extract(synth.frame(0, 752, 480), (0, 1000))
    _, k1, d1 = e.extract(synth.frame(1, 752, 480), (0, 1000))
    qs = np.zeros(len(k0), dtype=GQ)
    qs["u"], qs["v"], qs["r"] = k0["x"], k0["y"], 100.0   # SearchForInitialization's window (ORBmatcher.cc:665)
    qs["min_level"] = qs["max_level"] = k0["octave"]
    bounds = (0.0, 752.0, 0.0, 480.0)
    got = slices.search_grid_top2(k1, d1, bounds, qs, d0)
    want = slices.ref_grid_top2(k1, d1, bounds, qs, d0)
    for g, w_ in zip(got[:4], want[:4]):
        assert np.array_equal(g, w_)


def bow_pair_case(oracle, seed=0, n=1000, levelsup=2):
    """A keyframe / frame pair for SearchByBoW: two overlapping descriptor sets through the synthetic vocabulary."""
    k, L, _, parent, node_desc, weight, _ = synth.bow_case("k10L4_levelsup2")
    rng = np.random.default_rng(seed)
    d_kf = synth.descriptors_near_words(parent, node_desc, n, seed=20 + seed)
    d_f = d_kf[rng.permutation(n)[: n - 100]].copy()
    flips = rng.integers(0, 256, (len(d_f), 6))
    for i in range(len(d_f)):
        for bpos in flips[i]:
            d_f[i, bpos >> 3] ^= np.uint8(1 << (bpos & 7))
    d_f[:30] = d_f[30:60]                      # duplicates: distance ties inside a node, already-matched skips
    a_kf = rng.uniform(0, 360, n).astype(np.float32)
    a_f = ((a_kf[rng.integers(0, n, len(d_f))] + rng.choice([0.0, 0.0, 0.0, 45.0, 200.0], len(d_f))) % 360).astype(np.float32)
    valid = rng.choice([0, 1, 1, 1, 2], n).astype(np.uint8)
    fv_kf = oracle.bow_transform(parent, node_desc, weight, L, d_kf, levelsup)
    fv_f = oracle.bow_transform(parent, node_desc, weight, L, d_f, levelsup)
    return d_kf, valid, a_kf, fv_kf, d_f, a_f, fv_f


@pytest.mark.parametrize("seed,levelsup,nnratio,ori", [(0, 2, 0.7, True), (1, 2, 0.9, True), (2, 3, 0.7, False), (3, 1, 0.6, True),
                                                      (4, 4, 0.75, True)])
def test_search_by_bow_restatement_equals_reference_body(slices, seed, levelsup, nnratio, ori):
    case = bow_pair_case(slices, seed, 1000, levelsup)
    n_ref, m_ref = slices.ref_search_by_bow(*case, nnratio=nnratio, check_orientation=ori)
    n_got, m_got = slices.search_by_bow(*case, nnratio=nnratio, check_orientation=ori)
    assert n_got == n_ref and np.array_equal(m_got, m_ref)
    assert n_ref == int((m_ref >= 0).sum()) and n_ref > 50


def undistort_keypoint_case(name):
    """Keypoints with every field set (only pt may change) at the points of the undistortion case `name`."""
    from orb_slam3_ros2_b200.orbx import KP_DTYPE
    w, h, K, dist, pts = synth.undistort_case(name)
    rng = np.random.default_rng(11)
    kps = np.zeros(len(pts), dtype=KP_DTYPE)
    kps["x"], kps["y"] = pts[:, 0], pts[:, 1]
    kps["size"] = 31.0 * 1.2 ** rng.integers(0, 8, len(pts))
    kps["angle"] = rng.random(len(pts)) * 360
    kps["response"] = rng.integers(7, 120, len(pts))
    kps["octave"] = rng.integers(0, 8, len(pts))
    kps["class_id"] = -1
    return w, h, K, dist, kps


@pytest.mark.parametrize("name", [c[0] for c in synth.UNDISTORT_CASES])
def test_undistort_keypoints_and_image_bounds_equal_reference_bodies(slices, name):
    """Frame::UndistortKeyPoints / ComputeImageBounds: restatement == the reference's own bodies (Frame.cc:747-809), and the
    undistorted coordinates == the cv2 golden (the un-vendored primitive inside)."""
    import os
    w, h, K, dist, kps = undistort_keypoint_case(name)
    got = slices.undistort_keypoints(kps, K, dist)
    want = slices.ref_undistort_keypoints(kps, K, dist)
    assert got.tobytes() == want.tobytes()
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "prim_undistort.npz"))
    if dist[0] != 0:
        assert np.array_equal(np.stack([got["x"], got["y"]], 1).view(np.uint32), g[f"{name}__out"].view(np.uint32))
    else:
        assert got.tobytes() == kps.tobytes()  # Frame.cc:749-753
    b_got, b_want = slices.image_bounds(w, h, K, dist), slices.ref_image_bounds(w, h, K, dist)
    assert b_got.tobytes() == b_want.tobytes()
    assert slices.undistort_keypoints(kps[:0], K, dist).size == 0


def init_pair_case(oracle, nfeatures=1000, shift=1, w=752, h=480, corrupt=0):
    """Two extracted frames of the synthetic stream (frame `shift` = frame 0 moved by (3, 1) * shift px) as
    Tracking::MonocularInitialization hands them to SearchForInitialization (Tracking.cc:2479-2481, :2512): vbPrevMatched = the
    first frame's own keypoint positions."""
    e = oracle.PortExtractor(nfeatures=nfeatures)
    _, k1, d1 = e.extract(synth.frame(0, w, h), (0, 1000))
    _, k2, d2 = e.extract(synth.frame(shift, w, h), (0, 1000))
    if corrupt:  # duplicate descriptors in frame 2 so that several i1 compete for one i2 (the withdrawal branch, :705-709)
        rng = np.random.default_rng(corrupt)
        lvl0 = np.flatnonzero(k2["octave"] == 0)
        for _ in range(60):
            a, b = rng.choice(lvl0, 2, replace=False)
            d2[b] = d2[a]
        # and near-copies in frame 1: an earlier keypoint a gets a neighbour b's descriptor with 3 bits flipped, so a first takes
        # b's partner and b, arriving later with the smaller distance, takes it back (:705-709)
        lvl0 = np.flatnonzero(k1["octave"] == 0)
        for b in lvl0[10::7]:
            near = lvl0[(lvl0 < b) & (np.abs(k1["x"][lvl0] - k1["x"][b]) < 8) & (np.abs(k1["y"][lvl0] - k1["y"][b]) < 8)]
            if len(near):
                a = near[0]
                d1[a] = d1[b]
                d1[a, rng.integers(0, 32, 3)] ^= np.uint8(1) << rng.integers(0, 8, 3).astype(np.uint8)
    prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    return k1, d1, k2, d2, (0.0, float(w), 0.0, float(h)), prev


@pytest.mark.parametrize("nf,shift,window,nnratio,ori,corrupt", [(1000, 1, 100, 0.9, True, 0), (1000, 2, 100, 0.9, False, 0),
                                                                  (2000, 1, 30, 0.7, True, 3), (5000, 3, 100, 0.9, True, 5),
                                                                  (1000, 5, 10, 0.9, True, 0)])
def test_search_for_initialization_restatement_equals_reference_body(slices, nf, shift, window, nnratio, ori, corrupt):
    k1, d1, k2, d2, bounds, prev = init_pair_case(slices, nf, shift, corrupt=corrupt)
    got = slices.search_for_initialization(k1, d1, k2, d2, bounds, prev, window, nnratio, ori)
    want = slices.ref_search_for_initialization(k1, d1, k2, d2, bounds, prev, window, nnratio, ori)
    assert got[0] == want[0] and np.array_equal(got[1], want[1]) and got[2].tobytes() == want[2].tobytes()
    assert got[0] == int((got[1] >= 0).sum())
    if corrupt:
        assert slices.port_lib().orbo_search_for_initialization_withdrawn() > 0  # the withdrawal branch ran
    if window >= 30:
        assert got[0] > 50
    # second round, as the initializer retries with the updated vbPrevMatched
    got2 = slices.search_for_initialization(k1, d1, k2, d2, bounds, got[2], window, nnratio, ori)
    want2 = slices.ref_search_for_initialization(k1, d1, k2, d2, bounds, want[2], window, nnratio, ori)
    assert got2[0] == want2[0] and np.array_equal(got2[1], want2[1]) and got2[2].tobytes() == want2[2].tobytes()


def bow_kf_pair_case(oracle, seed=0, n=1000, levelsup=2):
    """Two keyframes for SearchByBoW(KF1, KF2): the pair of bow_pair_case plus MapPoint flags for the second keyframe."""
    d1, valid1, a1, fv1, d2, a2, fv2 = bow_pair_case(oracle, seed, n, levelsup)
    valid2 = np.random.default_rng(100 + seed).choice([0, 1, 1, 1, 1, 2], len(d2)).astype(np.uint8)
    return d1, valid1, a1, fv1, d2, valid2, a2, fv2


@pytest.mark.parametrize("seed,levelsup,nnratio,ori", [(0, 2, 0.75, True), (1, 2, 0.9, True), (2, 3, 0.75, False), (3, 1, 0.6, True),
                                                      (4, 4, 0.75, True)])
def test_search_by_bow_keyframes_restatement_equals_reference_body(slices, seed, levelsup, nnratio, ori):
    """ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, ...) (ORBmatcher.cc:765-905): restatement == the reference's own body."""
    case = bow_kf_pair_case(slices, seed, 1000, levelsup)
    n_ref, m_ref = slices.ref_search_by_bow_kf(*case, nnratio=nnratio, check_orientation=ori)
    n_got, m_got = slices.search_by_bow_kf(*case, nnratio=nnratio, check_orientation=ori)
    assert n_got == n_ref and np.array_equal(m_got, m_ref)
    assert n_ref == int((m_ref >= 0).sum()) and n_ref > 30
    sel = m_ref[m_ref >= 0]
    assert len(np.unique(sel)) == len(sel)                      # vbMatched2: a point of keyframe 2 is given away once
    assert (case[1][m_ref >= 0] == 1).all() and (case[5][sel] == 1).all()  # only good MapPoints on both sides


# row a11: orbx_accept_rule -> the reference sites whose `if` it restates (ORBmatcher.cc line; 1151 = Frame.cc:1151)
ACCEPT_SITES = {0: [123], 1: [327], 2: [702], 3: [848], 4: [1569, 1649, 1770], 5: [1312, 1437], 6: [1151], 7: [523, 636]}


@pytest.mark.parametrize("rule", sorted(ACCEPT_SITES))
def test_accept_rules_equal_the_reference_conditions(slices, rule):
    """The restated accept rules (oracle/match_oracle.cpp, the checker of the GPU acceptKernel) against the reference's own `if`
    lines, compiled from the sliced sources, over EVERY (d1, d2) in [0, 256]^2 and the ratios the reference uses (constructor
    calls: 0.6, 0.7, 0.75, 0.8, 0.9; ratioHamming 1.0 and the loop-closing values)."""
    d1, d2 = [a.ravel().astype(np.int32) for a in np.meshgrid(np.arange(257), np.arange(257), indexing="ij")]
    ratios = [0.6, 0.7, 0.75, 0.8, 0.9, 1.0] if rule != 7 else [1.0, 0.9, 1.5, 0.85, 2.0]
    for site in ACCEPT_SITES[rule]:
        for ratio in ratios:
            for lv in ([(0, 0), (2, 3), (3, 3)] if rule == 0 else [(0, 0)]):
                for results in ([2, 1, 0] if rule == 6 else [2]):
                    want = slices.ref_accept_grid(site, ratio, lv[0], lv[1], results).ravel()
                    idx2 = np.full(d1.shape, 5 if results >= 2 else -1, np.int64)
                    got = slices.accept_top2(rule, d1, d2, ratio, idx2=idx2, level1=np.full(d1.shape, lv[0], np.int32),
                                             level2=np.full(d1.shape, lv[1], np.int32))
                    assert np.array_equal(np.asarray(got, bool), want), (rule, site, ratio, lv, results)


@pytest.mark.parametrize("kw", [dict(k=0, seed=0), dict(k=3, seed=1, nf=2000, w=1241, h=376), dict(k=1, seed=2, mono=False, stereo=True, dz=0.3),
                                dict(k=2, seed=3, mono=False, stereo=True, dz=-0.3), dict(k=4, seed=4, rot=0.35), dict(k=5, seed=5, taken_frac=0.5),
                                dict(k=6, seed=6, w=640, h=360, nf=1250, mono=False, stereo=True)])
def test_search_by_projection_frame_frame_restatement_equals_reference_body(slices, kw):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono) (ORBmatcher.cc:1676-1887, TrackWithMotionModel's matcher):
    the restatement against the reference's own body on shim Frame / MapPoint / Sophus classes - nmatches and, per current feature,
    the map point it ends up holding; monocular and stereo (uRight gate, forward / backward level windows), with and without the
    rotation histogram, with a roll that empties some histogram bins."""
    c = make_projection_case(slices, **kw)
    for ori in (True, False):
        want = slices.ref_search_by_projection_ff(**dict(c, check_orientation=ori))
        got = slices.search_by_projection_ff(**dict(c, check_orientation=ori))
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), (kw, ori, got[0], want[0])
        assert want[0] > 50


@pytest.mark.parametrize("kw,th,far", [(dict(k=0, seed=0), 3.0, False), (dict(k=1, seed=1, stereo=True), 3.0, False), (dict(k=2, seed=2, nf=2000, w=1241, h=376, stereo=True), 5.0, True),
                                      (dict(k=3, seed=3, taken_frac=0.7), 1.0, False), (dict(k=4, seed=4, w=640, h=360, nf=1250), 15.0, True)])
def test_search_by_projection_map_points_restatement_equals_reference_body(slices, kw, th, far):
    """ORBmatcher::SearchByProjection(F, vpMapPoints, th, bFarPoints, thFarPoints) (ORBmatcher.cc:43-210, Tracking::SearchLocalPoints):
    restatement == the reference's own body, for the ratios the reference uses (0.8 in SearchLocalPoints), th = 1 (no factor), 3, 5, 15."""
    c = make_local_points_case(slices, **kw)
    for nn in (0.8, 0.6, 0.9):
        want = slices.ref_search_by_projection_map(**c, th=th, far_points=far, th_far_points=30.0, nnratio=nn)
        got = slices.search_by_projection_map(**c, th=th, far_points=far, th_far_points=30.0, nnratio=nn)
        assert got[0] == want[0] and np.array_equal(got[1], want[1]), (kw, th, nn, got[0], want[0])
        assert want[0] > 50
