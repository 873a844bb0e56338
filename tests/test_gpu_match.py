"""GPU parity of Hamming top-2 matching (through the C ABI) against the oracle's strict-< DescriptorDistance loop
and the committed cv2.BFMatcher.knnMatch goldens: indices and distances bit-exact, ties by lowest index."""
import numpy as np
import pytest

from conftest import load_golden
from orb_slam3_ros2_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx
    orbx.lib()
    return orbx


def _eq(got, want):
    for g, w in zip(got, want):
        assert np.array_equal(np.asarray(g, dtype=np.int64), np.asarray(w, dtype=np.int64))


def test_top2_against_knnmatch_golden(orbx):
    g = load_golden("match.npz")
    q, db = g["q"], g["t"]
    i1, d1, i2, d2 = orbx.match_top2(q, db)
    assert np.array_equal(i1, g["idx"][:, 0]) and np.array_equal(i2, g["idx"][:, 1])
    assert np.array_equal(d1, g["dist"][:, 0].astype(np.int32)) and np.array_equal(d2, g["dist"][:, 1].astype(np.int32))


@pytest.mark.parametrize("nq,nd", [(1, 1), (1, 2), (7, 3), (100, 255), (100, 256), (100, 257), (1000, 1006), (1025, 5000),
                                   (2000, 70001)])
def test_top2_matches_oracle(orbx, oracle, nq, nd):
    rng = np.random.default_rng(nq * 31 + nd)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (nd, 32), dtype=np.uint8)
    # plant duplicates / ties / extremes
    if nd >= 3:
        db[nd - 1] = db[0]
        q[0] = db[0]                       # distance 0, twice: indices (0, nd-1)
        db[1] = ~q[min(1, nq - 1)]         # distance 256
    _eq(orbx.match_top2(q, db), oracle.match_top2(q, db))


def test_top2_low_entropy_ties(orbx, oracle):
    rng = np.random.default_rng(3)
    # descriptors drawn from 16 distinct rows: every query has many equidistant candidates
    base = rng.integers(0, 256, (16, 32), dtype=np.uint8)
    db = base[rng.integers(0, 16, 20000)]
    q = base[rng.integers(0, 16, 300)]
    _eq(orbx.match_top2(q, db), oracle.match_top2(q, db))


def test_empty_and_single_row_database(orbx):
    q = np.zeros((5, 32), dtype=np.uint8)
    i1, d1, i2, d2 = orbx.match_top2(q, np.zeros((0, 32), dtype=np.uint8))
    assert (i1 == -1).all() and (i2 == -1).all() and (d1 == 256).all() and (d2 == 256).all()
    i1, d1, i2, d2 = orbx.match_top2(q, np.full((1, 32), 255, dtype=np.uint8))
    assert (i1 == 0).all() and (d1 == 256).all() and (i2 == -1).all()  # knnMatch would return one match only


def test_large_database_known_answers(orbx):
    """Property check at a size the CPU oracle would not finish quickly: query i is DB row r_i with 20 bit
    flips, so (idx1, d1) = (r_i, 20) and d2 is far away (random 256-bit rows: ~128 +- 8)."""
    import torch
    nd, nq = 3_000_000, 2000
    db = synth.descriptor_db(nd)
    q, rows = synth.queries_from_db(nq, nd)
    i1, d1, i2, d2 = orbx.match_top2(q, db)
    assert np.array_equal(i1, rows) and (d1 == 20).all()
    assert (d2 > 60).all() and (i2 != i1).all()
    # planted exact duplicates across chunk boundaries: tie -> lowest index first
    db2 = db.copy()
    dup = [1, nd // 3 + 5, nd // 2, nd - 1]
    for r in dup:
        db2[r] = q[7]
    j1, e1, j2, e2 = orbx.match_top2(q[7:8], db2)
    assert (j1[0], e1[0], j2[0], e2[0]) == (dup[0], 0, dup[1], 0)
    del torch


def test_sharded_merge_emulated_on_one_gpu(orbx, oracle):
    """The exchange step of config 5 without NCCL: G shards matched one after another on this GPU with global
    index bases, their candidate keys merged by the same merge kernel the all-gather feeds."""
    import torch
    rng = np.random.default_rng(11)
    nd, nq, G = 50_000, 257, 4
    db = rng.integers(0, 256, (nd, 32), dtype=np.uint8)
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    b = [orbx.shard_range(nd, r, G) for r in range(G)]
    db[b[1][0]] = q[3]; db[b[2][1] - 1] = q[3]; db[b[3][0] + 9] = q[3]   # ties across shard boundaries
    db[b[0][1] - 1] = q[4]; db[b[1][0] + 1] = q[4]
    want = oracle.match_top2(q, db)
    dq = torch.from_numpy(q).cuda()
    ddb = torch.from_numpy(db).cuda()
    k1 = torch.empty((G, nq), dtype=torch.int64, device="cuda")
    k2 = torch.empty((G, nq), dtype=torch.int64, device="cuda")
    out = [torch.empty(nq, dtype=torch.int64, device="cuda"), torch.empty(nq, dtype=torch.int32, device="cuda"),
           torch.empty(nq, dtype=torch.int64, device="cuda"), torch.empty(nq, dtype=torch.int32, device="cuda")]
    for r, (lo, hi) in enumerate(b):
        orbx._check(orbx.lib().orbx_match_top2_device(dq.data_ptr(), nq, ddb.data_ptr() + lo * 32, hi - lo, lo,
                                                      out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                                      out[3].data_ptr(), None))
        torch.cuda.synchronize()
        i1, d1, i2, d2 = [t.cpu().numpy().astype(np.int64) for t in out]
        assert ((i1 >= lo) & (i1 < hi)).all()  # global indices
        k1[r] = torch.from_numpy((d1 << 40) | i1).cuda()
        k2[r] = torch.from_numpy((d2 << 40) | i2).cuda()
    orbx._check(orbx.lib().orbx_merge_top2_device(k1.data_ptr(), k2.data_ptr(), G, nq, out[0].data_ptr(),
                                                  out[1].data_ptr(), out[2].data_ptr(), out[3].data_ptr(), None))
    torch.cuda.synchronize()
    _eq([t.cpu().numpy() for t in out], want)
    # the host restatement used by the gloo tests agrees with the device merge
    hk1, hk2 = orbx.merge_top2_host(k1.cpu().numpy().view(np.uint64), k2.cpu().numpy().view(np.uint64))
    assert np.array_equal((hk1 & np.uint64((1 << 40) - 1)).astype(np.int64), want[0])
    assert np.array_equal((hk2 >> np.uint64(40)).astype(np.int64), want[3])


def test_stereo_banded_matches_oracle(orbx, oracle):
    w, h, nf = 1241, 376, 2000
    exl = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    exr = orbx.ORBextractor(nf, 1.2, 8, 20, 7)
    _, kl, dl = exl(synth.frame(0, w, h), None, (0, 0))
    _, kr, dr = exr(synth.frame(0, w, h, right=True), None, (0, 0))
    sf = exl.GetScaleFactors()
    max_d = 718.856  # mbf/mb of config/Stereo/KITTI00-02.yaml (fx)
    bi, bd = orbx.match_stereo_banded(kl, dl, kr, dr, sf, h, max_d)
    wi, wd = oracle.match_stereo_banded(kl, dl, kr, dr, sf, h, max_d)
    assert np.array_equal(bi, wi) and np.array_equal(bd, wd)
    good = bd < 75  # thOrbDist (Frame.cc:815)
    assert good.sum() > 500
    # the synthetic right image is the left shifted by 16 px: accepted matches sit at disparity ~16
    disp = kl["x"][good] - kr["x"][bi[good]]
    assert np.median(disp) == pytest.approx(16.0, abs=1.5)
