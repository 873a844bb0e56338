"""Randomised GPU parity sweeps: sizes and counts drawn at random around the launch-configuration boundaries of the kernels (staging
block of 256 rows, query tiles of 128 / 1024, database chunks, strip / cell geometry, sub-batch partitions, stereo bands, colour /
resize staging), each result compared with the oracle.  The fixed-case tests pin known shapes; these look for shapes nobody thought
of (the sweep in test_gpu_extract.py found a resize launch that needed exactly 48 KB of dynamic shared memory)."""
import numpy as np
import pytest

from orb_slam3_ros2_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx
    orbx.lib()
    return orbx


@pytest.mark.parametrize("seed", range(24))
def test_top2_random_shapes(orbx, oracle, seed):
    rng = np.random.default_rng(2000 + seed)
    nq = int(rng.choice([1, 2, 31, 127, 128, 129, 255, 256, 257, 1000, 1023, 1024, 1025, 2000, 2049]))
    nd = int(rng.choice([1, 2, 3, 255, 256, 257, 511, 512, 513, 1000, 4095, 4096, 4097, 20000, 65537]))
    ent = int(rng.choice([256, 16, 4]))  # low entropy = many distance ties
    q = rng.integers(0, ent, (nq, 32)).astype(np.uint8)
    db = rng.integers(0, ent, (nd, 32)).astype(np.uint8)
    if nd > 4 and nq > 2:
        db[nd // 2] = q[0]
        db[nd - 1] = q[0]  # exact duplicates: index order decides
    got, want = orbx.match_top2(q, db), oracle.match_top2(q, db)
    for g, w in zip(got, want):
        assert np.array_equal(np.asarray(g, np.int64), np.asarray(w, np.int64)), (nq, nd, ent)


@pytest.mark.parametrize("seed", range(6))
def test_batch_random_shapes_with_subbatches_and_consecutive_matching(orbx, oracle, seed):
    rng = np.random.default_rng(3000 + seed)
    h = int(rng.integers(120, 400))
    w = int(rng.integers(max(h, 160), 700))
    F = int(rng.choice([1, 2, 7, 16, 17, 33]))
    nf = int(rng.choice([100, 500, 1200]))
    lap = [(0, 1000), (0, 0), (w // 3, w // 2)][seed % 3]
    imgs = synth.frames(seed, F, w, h)
    ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    ex.set_subbatches(int(rng.choice([1, 2, 3])))
    mono, n, kps, desc = ex.extract_batch(imgs, lap)
    po = oracle.PortExtractor(nfeatures=nf)
    ref = [po.extract(imgs[f], lap) for f in range(F)]
    for f in range(F):
        pm, pk, pd = ref[f]
        assert mono[f] == pm and n[f] == len(pk), (f, w, h, F, nf)
        assert np.array_equal(kps[f, :n[f]], pk) and np.array_equal(desc[f, :n[f]], pd), (f, w, h, F, nf)
    if F >= 2:
        i1, d1, i2, d2 = ex.match_consecutive()
        for f in range(min(F - 1, 3)):
            want = oracle.match_top2(ref[f][2], ref[f + 1][2])
            got = (i1[f, :n[f]], d1[f, :n[f]], i2[f, :n[f]], d2[f, :n[f]])
            for g, x in zip(got, want):
                assert np.array_equal(np.asarray(g, np.int64), np.asarray(x, np.int64)), (f, w, h, F)
    ex.close()


@pytest.mark.parametrize("seed", range(6))
def test_stereo_random_shapes(orbx, oracle, seed):
    rng = np.random.default_rng(4000 + seed)
    h = int(rng.integers(150, 420))
    w = int(rng.integers(max(h, 300), 1300))
    nf = int(rng.choice([300, 1000, 2000]))
    bf = float(rng.uniform(20, 400))
    b = float(rng.uniform(0.05, 0.6))
    exl, exr = orbx.ORBextractor(nf, 1.2, 8, 20, 7), orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    il, ir = synth.frame(seed, w, h), synth.frame(seed, w, h, right=True)
    _, kl, dl = exl(il, None, (0, 0))
    _, kr, dr = exr(ir, None, (0, 0))
    ur, depth = orbx.compute_stereo_matches(exl, exr, kl, dl, kr, dr, bf, b)
    ol, orr = oracle.PortExtractor(nfeatures=nf), oracle.PortExtractor(nfeatures=nf)
    ol.extract(il, (0, 0))
    orr.extract(ir, (0, 0))
    t = ol.tables()
    wur, wdepth = oracle.compute_stereo_matches(kl, dl, kr, dr, t["scale"], t["inv_scale"], [ol.pyramid(l) for l in range(8)],
                                                [orr.pyramid(l) for l in range(8)], bf, b)
    assert np.array_equal(ur, wur) and np.array_equal(depth, wdepth), (w, h, nf, bf, b)
    exl.close()
    exr.close()


@pytest.mark.parametrize("seed", range(8))
def test_color_staging_random_shapes(orbx, oracle, seed):
    rng = np.random.default_rng(5000 + seed)
    h = int(rng.integers(100, 500))
    w = int(rng.integers(max(h, 120), 900))
    ch = int(rng.choice([1, 3, 4]))
    rgb = bool(seed % 2)
    new = None if seed % 3 == 0 else (int(rng.integers(max(120, w // 3), w + 60)), int(rng.integers(max(100, h // 3), h + 40)))
    if new is not None and new[0] < new[1]:
        new = (new[1], new[0])
    base = synth.frame(seed, w, h)
    img = base if ch == 1 else np.stack([np.roll(base, 3 * c, axis=1) ^ np.uint8(17 * c) for c in range(ch)], axis=2)
    ex = orbx.ORBextractor(600, 1.2, 8, 20, 7)
    mono, kp, desc = ex.extract_color(img, rgb=rgb, new_size=new, vLappingArea=(0, 1000))
    gray = oracle.stage_gray(img, rgb=rgb, new_size=new)
    if not (ch == 1 and new is None):
        assert np.array_equal(ex.staged_gray(), gray), (w, h, ch, new)
    pm, pk, pd = oracle.PortExtractor(nfeatures=600).extract(gray, (0, 1000))
    assert mono == pm and np.array_equal(kp, pk) and np.array_equal(desc, pd), (w, h, ch, new)
    ex.close()
