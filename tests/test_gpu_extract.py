"""GPU parity of the extractor (through the C ABI) against the oracle: stage taps (pyramid, blur, FAST candidates),
end-to-end operator() results against the committed reference fixtures, the port, and — where oracle/_ref
travelled — the reference binary itself.  Integer/byte/index results must be bit-exact; angles are produced by
the same non-FMA float sequence as the oracle, so they are compared bit-exactly too (north_star allows 1e-3 deg)."""
import hashlib

import numpy as np
import pytest

from conftest import load_golden
from orb_slam3_ros2_b200 import synth

pytestmark = pytest.mark.gpu

E2E = ["euroc752", "d435i640", "kittiL", "kittiR", "hd1080", "milkv360", "init5x", "small320"]


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx
    orbx.lib()
    return orbx


def _same(a, b, what):
    assert len(a) == len(b), f"{what}: count {len(a)} vs {len(b)}"
    assert np.array_equal(a, b), f"{what}: first diff at {np.flatnonzero(np.any(a != b, axis=-1) if a.ndim > 1 else a != b)[:5]}"


def test_pyramid_blur_candidates_match_oracle(orbx, oracle):
    im = synth.frame(5, 752, 480)
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    po = oracle.PortExtractor(nfeatures=1000)
    ex(im, None, (0, 1000))
    po.extract(im, (0, 1000))
    for l in range(8):
        want = po.pyramid(l)
        got = ex.pyramid_level(0, l)
        assert got.shape == want.shape
        assert np.array_equal(got, want), f"pyramid level {l}: {np.count_nonzero(got != want)} bytes differ"
        assert np.array_equal(ex.mvImagePyramid[l], want[19:-19, 19:-19])
        blur = oracle.gauss7(np.ascontiguousarray(want[19:-19, 19:-19]))
        assert np.array_equal(ex.blurred_level(0, l), blur), f"blur level {l}"
        _same(ex.candidates(0, l), po.candidates(l), f"FAST candidates level {l}")


@pytest.mark.parametrize("name", E2E)
def test_operator_matches_reference_fixture(orbx, name):
    g = load_golden(f"e2e_{name}.npz")
    w, h, k, nf = int(g["w"]), int(g["h"]), int(g["k"]), int(g["nfeatures"])
    lap = tuple(int(v) for v in g["lap"])
    im = synth.frame(k, w, h, 42, bool(g["right"]))
    assert hashlib.sha1(im.tobytes()).hexdigest() == str(g["img_sha1"])
    for mode in ("strict", "native"):
        ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7, fp_mode=mode)
        mono, kp, desc = ex(im, None, lap)
        same = "native_same" in g.files and bool(g["native_same"])
        src = "strict" if (mode == "strict" or same) else "native"
        assert mono == int(g[f"{mode}_mono"])
        _same(kp, g[f"{src}_kp"], f"{name}/{mode} keypoints")
        _same(desc, g[f"{src}_desc"], f"{name}/{mode} descriptors")
        if mode == "strict":
            for l in range(8):
                p = ex.pyramid_level(0, l)
                assert p.shape == tuple(g["pyr_shape"][l])
                assert hashlib.sha1(p.tobytes()).hexdigest() == str(g["pyr_sha1"][l])
        ex.close()


def _vs(orbx, oracle, im, nf, lap, mode="strict", ini=20, mn=7, scale=1.2, levels=8):
    ex = orbx.ORBextractor(nf, scale, levels, ini, mn, fp_mode=mode)
    mono, kp, desc = ex(im, None, lap)
    checker = oracle.RefExtractor if oracle.have_ref(mode) else oracle.PortExtractor
    rm, rk, rd = checker(nfeatures=nf, scale_factor=scale, nlevels=levels, ini_th=ini, min_th=mn, mode=mode).extract(im, lap)
    assert mono == rm
    _same(kp, rk, "keypoints")
    _same(desc, rd, "descriptors")
    ex.close()
    return len(kp)


def test_adversarial_inputs_match_reference(orbx, oracle):
    rng = np.random.default_rng(9)
    total = 0
    # white noise: thousands of candidates per level and many (count, UL.x) ties in the quadtree's sort
    total += _vs(orbx, oracle, rng.integers(0, 256, (240, 320), dtype=np.uint8), 500, (0, 1000))
    total += _vs(orbx, oracle, rng.integers(0, 256, (376, 1241), dtype=np.uint8), 2000, (0, 0))
    total += _vs(orbx, oracle, rng.integers(0, 256, (480, 640), dtype=np.uint8), 5000, (0, 1000), "native")
    # low contrast: minThFAST fallback in most cells, levels with fewer candidates than N
    low = (synth.frame(3, 400, 300).astype(np.int32) - 128) // 6 + 128
    total += _vs(orbx, oracle, low.astype(np.uint8), 1000, (100, 200))
    # flat image: no keypoints at all
    assert _vs(orbx, oracle, np.full((240, 320), 77, dtype=np.uint8), 500, (0, 1000)) == 0
    # tall image, odd sizes.  (A level with round(width/height) = 0 makes the REFERENCE index an empty vector,
    # ORBextractor.cc:559-586 — e.g. 301x517 segfaults it — so 360x517 is the tallest shape it survives.)
    total += _vs(orbx, oracle, synth.frame(7, 360, 517), 700, (0, 1000))
    total += _vs(orbx, oracle, synth.frame(8, 1000, 130), 600, (300, 600))
    # few features: N_l small, the quadtree stops in its first rounds
    total += _vs(orbx, oracle, synth.frame(9, 640, 480), 40, (0, 1000))
    assert total > 8000


def test_other_constructor_settings_match_reference(orbx, oracle):
    """Scale factors / level counts / thresholds other than the shipped YAMLs (the class accepts any): wider resize
    strips, 12-level pyramids whose top levels are smaller than one FAST cell, high and inverted thresholds."""
    total = 0
    total += _vs(orbx, oracle, synth.frame(11, 752, 480), 1000, (0, 1000), scale=1.5, levels=5)
    total += _vs(orbx, oracle, synth.frame(12, 640, 480), 800, (0, 1000), scale=1.1, levels=12)
    total += _vs(orbx, oracle, synth.frame(13, 1241, 376), 1500, (0, 0), scale=1.9, levels=4, mode="native")
    total += _vs(orbx, oracle, synth.frame(14, 333, 251), 300, (0, 1000), scale=1.3, levels=6)
    total += _vs(orbx, oracle, synth.frame(15, 640, 480), 1000, (0, 1000), ini=60, mn=30)
    total += _vs(orbx, oracle, synth.frame(16, 640, 480), 1000, (0, 1000), ini=7, mn=20)
    total += _vs(orbx, oracle, synth.frame(17, 640, 480), 1000, (0, 1000), ini=140, mn=130)
    assert total > 3000


def test_empty_image_returns_minus_one(orbx):
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    mono, kp, desc = ex(None)
    assert mono == -1 and len(kp) == 0 and len(desc) == 0


def test_tables_match_reference_constants(orbx, oracle):
    for nf in (1000, 1250, 2000, 4000, 5000):
        ex = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
        t = oracle.PortExtractor(nfeatures=nf).tables()
        assert np.array_equal(ex.features_per_level(), t["nfeatures"])
        assert np.array_equal(ex.GetScaleFactors(), t["scale"])
        assert np.array_equal(ex.GetInverseScaleFactors(), t["inv_scale"])
        assert np.array_equal(ex.GetScaleSigmaSquares(), t["sigma2"])
        assert np.array_equal(ex.GetInverseScaleSigmaSquares(), t["inv_sigma2"])
        assert ex.GetLevels() == 8


def test_batch_equals_single_frame_and_device_synth(orbx, oracle):
    import torch
    F, w, h = 6, 640, 480
    imgs = synth.frames(10, F, w, h)
    dev = torch.empty((F, h, w), dtype=torch.uint8, device="cuda")
    orbx._check(orbx.lib().orbx_synth_frames_device(dev.data_ptr(), 10, F, w, h, 42, 0, -1))
    assert np.array_equal(dev.cpu().numpy(), imgs), "device synthetic generator differs from synth.py"
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    mono, n, kps, desc = ex.extract_batch(imgs, (0, 1000))
    po = oracle.PortExtractor(nfeatures=1000)
    for f in range(F):
        pm, pk, pd = po.extract(imgs[f], (0, 1000))
        assert mono[f] == pm and n[f] == len(pk)
        _same(kps[f, :n[f]], pk, f"batch frame {f} keypoints")
        _same(desc[f, :n[f]], pd, f"batch frame {f} descriptors")
    # device-resident path gives the same
    ex.extract_batch_device(dev.data_ptr(), F, w, h, (0, 1000))
    ex.sync()
    for f in (0, F - 1):
        m2, k2, d2 = ex.download(f)
        assert m2 == mono[f]
        _same(k2, kps[f, :n[f]], "device batch keypoints")
        _same(d2, desc[f, :n[f]], "device batch descriptors")
    # frame-to-frame matching over the batch (config 2)
    i1, d1, i2, d2m = ex.match_consecutive()
    for f in range(F - 1):
        want = oracle.match_top2(desc[f, :n[f]], desc[f + 1, :n[f + 1]])
        got = (i1[f, :n[f]], d1[f, :n[f]], i2[f, :n[f]], d2m[f, :n[f]])
        for g_, w_ in zip(got, want):
            assert np.array_equal(g_, w_)
    assert orbx.accept_bow(d1[0, :n[0]], d2m[0, :n[0]]).sum() > 100  # the synthetic stream really overlaps


@pytest.mark.parametrize("nsub", [2, 3])
def test_subbatch_streams_give_identical_results(orbx, nsub):
    """orbx_set_subbatches: parts of the batch on concurrent stream pairs; frames are independent, so nothing changes."""
    F, w, h = 26, 640, 360
    imgs = synth.frames(3, F, w, h)
    ex = orbx.ORBextractor(1250, 1.2, 8, 20, 7)
    want = ex.extract_batch(imgs, (300, 600))
    wm = ex.match_consecutive()
    ex.set_subbatches(nsub)
    for _ in range(2):  # twice: the work counters and scratch of the parts must reset correctly
        got = ex.extract_batch(imgs, (300, 600))
        gm = ex.match_consecutive()
        n = want[1]
        assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], n)
        for f in range(F):
            assert np.array_equal(got[2][f, :n[f]], want[2][f, :n[f]]) and np.array_equal(got[3][f, :n[f]], want[3][f, :n[f]])
        for a, b in zip(gm, wm):
            for f in range(F - 1):
                assert np.array_equal(a[f, :n[f]], b[f, :n[f]])
    assert ex.timing()[0] > 0


@pytest.mark.parametrize("w,h,nf", [(752, 480, 40), (752, 480, 100), (1241, 376, 100), (1241, 376, 150)])
def test_small_nfeatures_match_reference(orbx, oracle, w, h, nf):
    """A level returns up to 4 * nIni keypoints when its target N is smaller than that (the first breadth pass of
    DistributeOctTree splits every root, ORBextractor.cc:610-681): capacity max(N + 4, 4 * nIni + 1) per level."""
    n = _vs(orbx, oracle, synth.frame(7, w, h), nf, (0, 1000))
    assert n >= nf


def test_extract_frame_one_sync_path_equals_extract(orbx):
    """orbx_extract_frame (the call the C++ drop-in class makes: results and the padded pyramid through one pinned staging buffer, one
    host synchronisation) returns exactly what orbx_extract + orbx_get_pyramid return; an empty image gives -1."""
    ex = orbx.ORBextractor(1000, 1.2, 8, 20, 7)
    for k, (w, h), lap in ((0, (752, 480), (0, 1000)), (3, (640, 360), (0, 0)), (5, (700, 500), (300, 600))):
        im = synth.frame(k, w, h)
        view = np.zeros((h, w + 24), np.uint8)  # a strided view: stride != width
        view[:, :w] = im
        mono, kp, desc = ex(im, None, lap)
        planes = [ex.pyramid_level(0, l) for l in range(8)]
        for src in (im, view[:, :w]):
            m2, k2, d2, p2 = orbx.extract_frame(ex, src, lap, pyramid=True)
            assert m2 == mono
            _same(k2, kp, "keypoints")
            _same(d2, desc, "descriptors")
            assert all(np.array_equal(a, b) for a, b in zip(p2, planes))
        m3, k3, d3, p3 = orbx.extract_frame(ex, im, lap, pyramid=False)
        assert m3 == mono and p3 is None and np.array_equal(k3, kp) and np.array_equal(d3, desc)
    assert orbx.extract_frame(ex, None)[0] == -1
    ex.close()


@pytest.mark.parametrize("seed", range(64))
def test_random_geometry_sweep_matches_reference(orbx, oracle, seed):
    """Random frame sizes (64 configurations; cell widths from 35 to 69 px, clipped last cells, levels smaller than one cell), feature counts, thresholds,
    scale factors, level counts and contents (texture, texture + noise, low contrast): every strip / cell geometry the FAST kernels can
    meet, against the reference.  Shapes keep round(width / height) >= 1 on every level (below that the reference indexes an empty vector)."""
    rng = np.random.default_rng(1000 + seed)
    h = int(rng.integers(70, 600))
    w = int(rng.integers(max(h, 90), 1000))
    levels = int(rng.integers(1, 11))
    scale = float(rng.choice([1.1, 1.2, 1.25, 1.4, 1.7]))
    while levels > 1 and h / scale ** (levels - 1) < 48:  # a level of <= 32 rows makes the REFERENCE resize a vector to a negative size
        levels -= 1
    nf = int(rng.choice([60, 300, 1000, 2500]))
    ini = int(rng.integers(5, 60))
    mn = int(rng.integers(2, ini + 1))
    im = synth.frame(seed, w, h).astype(np.int32)
    kind = seed % 3
    if kind == 1:
        im = im + rng.integers(-25, 26, im.shape)
    elif kind == 2:
        im = (im - 128) // 4 + 128
    im = np.clip(im, 0, 255).astype(np.uint8)
    lap = [(0, 1000), (0, 0), (w // 3, 2 * w // 3)][seed % 3]
    _vs(orbx, oracle, im, nf, lap, mode="native" if seed % 4 == 3 else "strict", ini=ini, mn=mn, scale=scale, levels=levels)
