"""CPU tests (no GPU) of the matcher-side oracle rows (oracle/match_oracle.cpp): accept rules, guided grid search and
Frame::ComputeStereoMatches, each against hand-checked known answers or an independent numpy restatement."""
import numpy as np
import pytest

from orb_slam3_ros2_b200 import synth

KP = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
               ("class_id", "<i4")])
GQ = np.dtype([("u", "<f4"), ("v", "<f4"), ("r", "<f4"), ("min_level", "<i4"), ("max_level", "<i4"), ("ur", "<f4")])


def test_accept_rules_known_answers(oracle):
    # (rule, d1, d2, nnratio, extra) -> expected; boundaries of TH_LOW = 50, TH_HIGH = 100 and of the ratio test
    A = oracle.accept_top2
    # BoW KF->F (ORBmatcher.cc:327-329): d1 <= 50 and d1 < 0.7 * d2 (float)
    assert list(A(1, [50, 51, 35, 34, 0], [100, 100, 50, 50, 256], 0.7)) == [True, False, False, True, True]
    # BoW KF<->KF (:848-850): strict d1 < 50
    assert list(A(3, [50, 49], [100, 100], 0.7)) == [False, True]
    # init (:697-699): same numbers as BoW KF->F
    assert list(A(2, [50, 51, 35, 34], [100, 100, 50, 50], 0.9)) == [True, False, True, True]
    # projection to map (:123-128): ratio applies only when both candidates are on the same level
    assert list(A(0, [100, 101, 90, 90, 90], [100, 200, 100, 100, 256], 0.8, level1=[1, 1, 2, 2, 3], level2=[1, 1, 2, 3, -1])) == \
        [False, False, False, True, True]
    assert list(A(4, [100, 101], [0, 0])) == [True, False]
    assert list(A(5, [50, 51], [0, 0])) == [True, False]
    # fisheye knn (Frame.cc:1149): needs two results; 7 < 10 * 0.7 is False in double (7.0 < 7.000000000000001 is True!)
    got = A(6, [7, 69, 70, 10], [10, 100, 100, 256], idx2=[3, 3, 3, -1])
    assert list(got) == [bool(np.float64(7.0) < np.float64(np.float32(10)) * 0.7), True, bool(70.0 < 100 * 0.7), False]


def _grid_search_numpy(kps, desc, bounds, queries, dq, skip=None, ur=None):
    """Independent restatement (numpy/python) of Frame.cc:657-736 + ORBmatcher.cc:75-117."""
    f32 = np.float32
    mnx, mxx, mny, mxy = [f32(b) for b in bounds]
    iw, ih = f32(64) / f32(mxx - mnx), f32(48) / f32(mxy - mny)
    grid = {}
    for i, k in enumerate(kps):
        vx, vy = f32(f32(k["x"] - mnx) * iw), f32(f32(k["y"] - mny) * ih)
        px = int(np.sign(vx) * np.floor(abs(vx) + f32(0.5)))  # roundf: half away from zero
        py = int(np.sign(vy) * np.floor(abs(vy) + f32(0.5)))
        if 0 <= px < 64 and 0 <= py < 48:
            grid.setdefault((px, py), []).append(i)
    pop = np.unpackbits(np.arange(256, dtype=np.uint8)[:, None], axis=1).sum(1)
    out = []
    for qi, q in enumerate(queries):
        x, y, r = f32(q["u"]), f32(q["v"]), f32(q["r"])
        best = [256, -1, -1, 256, -1, -1]  # d1 i1 l1 d2 i2 l2
        x0 = max(0, int(np.floor(f32(f32(f32(x - mnx) - r) * iw))))
        x1 = min(63, int(np.ceil(f32(f32(f32(x - mnx) + r) * iw))))
        y0 = max(0, int(np.floor(f32(f32(f32(y - mny) - r) * ih))))
        y1 = min(47, int(np.ceil(f32(f32(f32(y - mny) + r) * ih))))
        if not (x0 >= 64 or x1 < 0 or y0 >= 48 or y1 < 0):
            check = q["min_level"] > 0 or q["max_level"] >= 0
            for ix in range(x0, x1 + 1):
                for iy in range(y0, y1 + 1):
                    for i in grid.get((ix, iy), []):
                        k = kps[i]
                        if check and (k["octave"] < q["min_level"] or (q["max_level"] >= 0 and k["octave"] > q["max_level"])):
                            continue
                        if not (abs(f32(k["x"] - x)) < r and abs(f32(k["y"] - y)) < r):
                            continue
                        if skip is not None and skip[i]:
                            continue
                        if ur is not None and ur[i] > 0 and abs(f32(q["ur"] - ur[i])) > r:
                            continue
                        d = int(pop[desc[i] ^ dq[qi]].sum())
                        if d < best[0]:
                            best = [d, i, int(k["octave"]), best[0], best[1], best[2]]
                        elif d < best[3]:
                            best[3:] = [d, i, int(k["octave"])]
        out.append(best)
    o = np.array(out, dtype=np.int32)
    return o[:, 1], o[:, 0], o[:, 4], o[:, 3], o[:, 2], o[:, 5]


def make_grid_case(seed, nt=600, nq=200, w=752, h=480):
    rng = np.random.default_rng(seed)
    kps = np.zeros(nt, dtype=KP)
    kps["x"] = rng.uniform(-5, w + 5, nt).astype(np.float32)
    kps["y"] = rng.uniform(-5, h + 5, nt).astype(np.float32)
    kps["x"][:40] = np.round(kps["x"][:40])  # integer coordinates sit on rounding boundaries of PosInGrid
    kps["y"][:40] = np.round(kps["y"][:40])
    kps["x"][40:80] = kps["x"][0]            # crowded cells: order inside a cell matters for ties
    kps["y"][40:80] = kps["y"][0]
    kps["octave"] = rng.integers(0, 8, nt)
    desc = rng.integers(0, 256, (nt, 32), dtype=np.uint8)
    desc[40:80, 4:] = desc[40, 4:]           # near-duplicates -> distance ties
    qs = np.zeros(nq, dtype=GQ)
    src = rng.integers(0, nt, nq)
    qs["u"] = kps["x"][src] + rng.uniform(-3, 3, nq).astype(np.float32)
    qs["v"] = kps["y"][src] + rng.uniform(-3, 3, nq).astype(np.float32)
    qs["r"] = rng.choice([7.0, 15.0, 30.0, 100.0], nq).astype(np.float32)
    lv = kps["octave"][src]
    mode = rng.integers(0, 3, nq)
    qs["min_level"] = np.where(mode == 0, 0, np.where(mode == 1, lv - 1, lv))
    qs["max_level"] = np.where(mode == 0, -1, lv)
    qs["ur"] = qs["u"] - 10
    qs["u"][:5] = [-200, 2000, 100, 100, 0]   # windows partly / fully outside the grid
    qs["v"][:5] = [100, 100, -300, 900, 0]
    dq = desc[src].copy()
    flip = rng.integers(0, 256, (nq, 32), dtype=np.uint8) & rng.integers(0, 256, (nq, 32), dtype=np.uint8) & 0x11
    dq ^= flip.astype(np.uint8)
    skip = (rng.random(nt) < 0.1).astype(np.uint8)
    ur = np.where(rng.random(nt) < 0.5, kps["x"] - rng.uniform(0, 40, nt), -1).astype(np.float32)
    return kps, desc, (0.0, float(w), 0.0, float(h)), qs, dq, skip, ur


@pytest.mark.parametrize("filters", [False, True])
def test_grid_search_oracle_equals_numpy_restatement(oracle, filters):
    kps, desc, bounds, qs, dq, skip, ur = make_grid_case(3, nt=300, nq=80)
    if not filters:
        skip = ur = None
    got = oracle.search_grid_top2(kps, desc, bounds, qs, dq, skip, ur)
    want = _grid_search_numpy(kps, desc, bounds, qs, dq, skip, ur)
    for g, w in zip(got, want):
        assert np.array_equal(g, w)
    assert (got[0] >= 0).sum() > 40


def stereo_case(oracle, w=640, h=360, nf=1000):
    exl, exr = oracle.PortExtractor(nfeatures=nf), oracle.PortExtractor(nfeatures=nf)
    _, kl, dl = exl.extract(synth.frame(0, w, h), (0, 0))
    _, kr, dr = exr.extract(synth.frame(0, w, h, right=True), (0, 0))
    t = exl.tables()
    pl = [exl.pyramid(l) for l in range(8)]
    pr = [exr.pyramid(l) for l in range(8)]
    return kl, dl, kr, dr, t["scale"], t["inv_scale"], pl, pr


def test_compute_stereo_matches_known_disparity(oracle):
    """The synthetic right image is the left one shifted by exactly 16 px, so every refined match must report
    uRight = uL - 16 up to the parabola's sub-pixel term, and depth = bf / disparity."""
    kl, dl, kr, dr, sf, isf, pl, pr = stereo_case(oracle)
    bf, b = 386.1448, 0.53716  # config/Stereo/KITTI00-02.yaml: fx * b, b
    ur, dp = oracle.compute_stereo_matches(kl, dl, kr, dr, sf, isf, pl, pr, bf, b)
    ok = ur >= 0
    assert ok.sum() > 300
    disp = kl["x"][ok] - ur[ok]
    assert np.abs(disp - 16.0).max() < 2.5 and abs(np.median(disp) - 16.0) < 0.2
    assert np.array_equal(dp[ok], (np.float32(bf) / disp.astype(np.float32)).astype(np.float32))
    assert (dp[~ok] == -1).all() and (ur[~ok] == -1).all()
    # no right keypoints -> nothing matched, no crash
    ur0, dp0 = oracle.compute_stereo_matches(kl, dl, kr[:0], dr[:0], sf, isf, pl, pr, bf, b)
    assert (ur0 == -1).all() and (dp0 == -1).all()


def make_projection_case(oracle, k=0, w=752, h=480, nf=1000, seed=0, mono=True, dz=0.0, rot=0.0, taken_frac=0.05, stereo=False):
    """Two consecutive synthetic frames (frame k+1 is frame k shifted by (-3, -1) px), the last frame's keypoints back-projected with
    pseudo-random depths into map points, a current pose that moves them by about the image shift (+ an optional z motion / roll)."""
    from orb_slam3_ros2_b200 import synth
    rng = np.random.default_rng(seed)
    ex = oracle.PortExtractor(nfeatures=nf)
    _, kl, dl = ex.extract(synth.frame(k, w, h), (0, 1000))
    _, kc, dc = ex.extract(synth.frame(k + 1, w, h), (0, 1000))
    sf = ex.tables()["scale"]
    fx, fy, cx, cy = 458.654, 457.296, 367.215, 248.375  # config/Monocular/EuRoC.yaml
    n = len(kl)
    z = rng.uniform(4.0, 9.0, n).astype(np.float32)
    world = np.stack([(kl["x"] - cx) / fx * z, (kl["y"] - cy) / fy * z, z], axis=1).astype(np.float32)  # LastFrame pose = identity
    flags = (rng.random(n) < 0.85).astype(np.uint8) | ((rng.random(n) < 0.7).astype(np.uint8) << 1)
    mp_desc = dl.copy()
    flip = rng.random(n) < 0.3  # some map point descriptors differ from the last observation by a few bits
    for i in np.flatnonzero(flip):
        for b in rng.integers(0, 256, 6):
            mp_desc[i, b // 8] ^= np.uint8(1 << (b % 8))
    half = np.float32(rot / 2)
    pose_cw = np.array([0, 0, np.sin(half), np.cos(half), -3.0 * 6.0 / fx, -1.0 * 6.0 / fy, dz], dtype=np.float32)
    pose_lw = np.array([0, 0, 0, 1, 0, 0, 0], dtype=np.float32)
    taken = (rng.random(len(kc)) < taken_frac).astype(np.uint8)
    ur = None
    if stereo:
        ur = np.where(rng.random(len(kc)) < 0.6, kc["x"] - rng.uniform(2, 40, len(kc)), -1).astype(np.float32)
    return dict(kps_last=kl, last_flags=flags, world=world, mp_desc=mp_desc, kps_cur=kc, desc_cur=dc, pose_cw=pose_cw, pose_lw=pose_lw,
                cam=(fx, fy, cx, cy), bounds=(0.0, float(w), 0.0, float(h)), scale_factors=sf, th=15 if mono else 7, mono=mono,
                mbf=40.0 if stereo else 0.0, mb=0.08 if stereo else 0.0, cur_u_right=ur, cur_taken=taken)


def test_search_by_projection_frame_frame_known_properties(oracle):
    c = make_projection_case(oracle, k=2, seed=1)
    n, m = oracle.search_by_projection_ff(**c)
    assert n > 300  # most projected map points find their feature again
    hit = np.flatnonzero(m >= 0)
    assert len(hit) <= n  # later points may overwrite earlier ones (nmatches counts both)
    kl, kc = c["kps_last"], c["kps_cur"]
    dx = kc["x"][hit] - kl["x"][m[hit]]
    assert np.median(np.abs(dx + 3.0 * c["scale_factors"][kl["octave"][m[hit]]] ** 0)) < 15  # matches sit near the shifted position
    assert not np.any(c["cur_taken"][hit])          # features that already held a map point are never given another
    assert np.all(c["last_flags"][m[hit]] & 1)      # only valid map points are projected
    n0, m0 = oracle.search_by_projection_ff(**dict(c, check_orientation=False))
    assert n0 >= n                                   # the rotation histogram only removes matches


def make_local_points_case(oracle, k=0, w=752, h=480, nf=1000, seed=0, stereo=False, taken_frac=0.3, n_extra=400):
    """Tracking::SearchLocalPoints' input: map points projected into the frame (here: the frame's own keypoints jittered by a few pixels,
    plus points that match nothing), predicted levels around the observed octave, viewing cosines on both sides of 0.998, some points
    out of view / bad / far, part of the features already matched (TrackWithMotionModel ran before)."""
    from orb_slam3_ros2_b200 import synth
    rng = np.random.default_rng(seed)
    ex = oracle.PortExtractor(nfeatures=nf)
    _, kc, dc = ex.extract(synth.frame(k, w, h), (0, 1000))
    sf = ex.tables()["scale"]
    n = len(kc)
    pick = rng.permutation(n)[: int(0.8 * n)]
    m = len(pick) + n_extra
    track = np.zeros((m, 5), np.float32)
    level = np.zeros(m, np.int32)
    desc = np.zeros((m, 32), np.uint8)
    track[:len(pick), 0] = kc["x"][pick] + rng.normal(0, 2.0, len(pick))
    track[:len(pick), 1] = kc["y"][pick] + rng.normal(0, 2.0, len(pick))
    level[:len(pick)] = np.clip(kc["octave"][pick] + rng.integers(0, 2, len(pick)), 0, 7)
    desc[:len(pick)] = dc[pick]
    for i in range(len(pick)):  # a few bits of noise
        for b in rng.integers(0, 256, int(rng.integers(0, 25))):
            desc[i, b // 8] ^= np.uint8(1 << (b % 8))
    track[len(pick):, 0] = rng.uniform(0, w, n_extra); track[len(pick):, 1] = rng.uniform(0, h, n_extra)
    level[len(pick):] = rng.integers(0, 8, n_extra)
    desc[len(pick):] = rng.integers(0, 256, (n_extra, 32), dtype=np.uint8)
    track[:, 2] = track[:, 0] - rng.uniform(2, 40, m)            # mTrackProjXR
    track[:, 3] = np.where(rng.random(m) < 0.5, 0.9995, 0.95)    # mTrackViewCos
    track[:, 4] = rng.uniform(1, 80, m)                          # mTrackDepth
    flags = (rng.random(m) < 0.9).astype(np.uint8) | ((rng.random(m) < 0.9).astype(np.uint8) << 1) | ((rng.random(m) < 0.03).astype(np.uint8) << 2)
    order = rng.permutation(m)
    taken = (rng.random(n) < taken_frac).astype(np.uint8)
    ur = np.where(rng.random(n) < 0.6, kc["x"] - rng.uniform(2, 40, n), -1).astype(np.float32) if stereo else None
    return dict(track=track[order], level=level[order], flags=flags[order], mp_desc=desc[order], kps_cur=kc, desc_cur=dc,
                bounds=(0.0, float(w), 0.0, float(h)), scale_factors=sf, cur_u_right=ur, cur_taken=taken)


def test_search_by_projection_map_points_known_properties(oracle):
    c = make_local_points_case(oracle, k=1, seed=2)
    n, m = oracle.search_by_projection_map(**c, th=3.0)
    hit = np.flatnonzero(m >= 0)
    assert n > 150 and len(hit) <= n
    assert not np.any(c["cur_taken"][hit])
    assert np.all(c["flags"][m[hit]] & 1) and not np.any(c["flags"][m[hit]] & 4)
    nfar, _ = oracle.search_by_projection_map(**c, th=3.0, far_points=True, th_far_points=20.0)
    assert nfar < n  # far points are skipped
