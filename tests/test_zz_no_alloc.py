"""orbx_match_top2_device_noalloc (the entry point that allocates nothing; no counterpart in the reference)
against orbx_match_top2_device on the same device buffers: bit-exact, on calls in a row that share the
static key array (the later ones smaller, so the array still holds the earlier call's keys), plus the size limit.
Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_KEYS = 1 << 20  # ORBX_NOALLOC_MAX_KEYS (include/orbx.h)


@pytest.fixture(scope="module")
def orbx():
    from orb_slam3_ros2_b200 import orbx
    orbx.lib()
    return orbx


def _outs(torch, nq):
    return [torch.full((nq,), -7, dtype=torch.int64, device="cuda"), torch.full((nq,), -7, dtype=torch.int32, device="cuda"),
            torch.full((nq,), -7, dtype=torch.int64, device="cuda"), torch.full((nq,), -7, dtype=torch.int32, device="cuda")]


def _call(orbx, name, dq, nq, ddb, nd, base, outs):
    fn = getattr(orbx.lib(), name)
    orbx._check(fn(dq.data_ptr(), nq, ddb.data_ptr() if nd else None, nd, base, *[o.data_ptr() for o in outs], None))


def _case(rng, nq, nd):
    q = rng.integers(0, 256, (nq, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (max(nd, 1), 32), dtype=np.uint8)
    if nd >= 4000:  # exact copies and ties, some across a chunk boundary (chunks are multiples of 256 rows)
        db[17] = q[3]; db[nd - 1] = q[3]; db[nd // 2] = q[3]
        db[255] = q[5]; db[256] = q[5]
        db[1000] = q[min(9, nq - 1)] ^ np.uint8(1)
    return q, db


def test_noalloc_equals_match_top2_device_on_calls_in_a_row(orbx):
    import torch
    rng = np.random.default_rng(2026)
    # (nq, nd, index base): big first, then smaller ones that leave most of the array holding the first call's keys
    shapes = [(1500, 200_000, 0), (257, 50_000, 123_456_789), (1000, 1000, 0), (33, 1, 5), (64, 0, 0), (1, 300, 0), (1500, 200_000, 7)]
    prepared = []
    for nq, nd, base in shapes:
        q, db = _case(rng, nq, nd)
        dq, ddb = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
        want = _outs(torch, nq)
        _call(orbx, "orbx_match_top2_device", dq, nq, ddb, nd, base, want)
        prepared.append((nq, nd, base, dq, ddb, want, _outs(torch, nq)))
    torch.cuda.synchronize()
    free_before = None
    for k, (nq, nd, base, dq, ddb, want, got) in enumerate(prepared):
        if k == 1:  # the first call may load the module; from the second on the device's free memory must not move
            free_before = torch.cuda.mem_get_info()[0]
        _call(orbx, "orbx_match_top2_device_noalloc", dq, nq, ddb, nd, base, got)
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free_before, "the no-alloc entry point changed the device's free memory"
    for nq, nd, base, dq, ddb, want, got in prepared:
        for g, w in zip(got, want):
            assert torch.equal(g, w), f"nq={nq} nd={nd} base={base}"
    # eight queries of the second shape against a numpy brute force (the comparison above is not two wrongs agreeing)
    nq, nd, base, dq, ddb, want, got = prepared[1]
    q, db = dq.cpu().numpy(), ddb.cpu().numpy()
    d = np.unpackbits(q[:8, None, :] ^ db[None, :, :], axis=2).sum(axis=2)
    order = np.argsort(d, axis=1, kind="stable")
    assert np.array_equal(got[0].cpu().numpy()[:8], order[:, 0] + base) and np.array_equal(got[2].cpu().numpy()[:8], order[:, 1] + base)
    assert np.array_equal(got[1].cpu().numpy()[:8], d[np.arange(8), order[:, 0]])


def test_noalloc_same_call_twice_is_stable(orbx):
    import torch
    rng = np.random.default_rng(7)
    nq, nd = 700, 30_000
    q, db = _case(rng, nq, nd)
    dq, ddb = torch.from_numpy(q).cuda(), torch.from_numpy(db).cuda()
    want = _outs(torch, nq)
    _call(orbx, "orbx_match_top2_device", dq, nq, ddb, nd, 0, want)
    a, b = _outs(torch, nq), _outs(torch, nq)
    _call(orbx, "orbx_match_top2_device_noalloc", dq, nq, ddb, nd, 0, a)
    _call(orbx, "orbx_match_top2_device_noalloc", dq, nq, ddb, nd, 0, b)
    torch.cuda.synchronize()
    for x, y, w in zip(a, b, want):
        assert torch.equal(x, w) and torch.equal(y, w)


def test_noalloc_rejects_sizes_beyond_the_static_array(orbx):
    import torch
    nq = MAX_KEYS + 1  # one chunk x nq keys already exceeds the array
    dq = torch.zeros((nq, 32), dtype=torch.uint8, device="cuda")
    ddb = torch.zeros((2, 32), dtype=torch.uint8, device="cuda")
    outs = _outs(torch, nq)
    with pytest.raises(orbx.OrbxError) as e:
        _call(orbx, "orbx_match_top2_device_noalloc", dq, nq, ddb, 2, 0, outs)
    assert e.value.code == orbx.ORBX_ERR_CAPACITY
    assert "ORBX_NOALLOC_MAX_KEYS" in str(e.value) and str(nq) in str(e.value)
    torch.cuda.synchronize()
    assert all(int(o[0]) == -7 for o in outs)  # nothing was launched
    # the same sizes are fine for the allocating entry point
    _call(orbx, "orbx_match_top2_device", dq, nq, ddb, 2, 0, outs)
    torch.cuda.synchronize()
    assert int(outs[0][0]) == 0 and int(outs[2][0]) == 1 and int(outs[1][0]) == 0
    # exactly at the limit: accepted
    nq = MAX_KEYS
    outs = _outs(torch, nq)
    _call(orbx, "orbx_match_top2_device_noalloc", dq, nq, ddb, 2, 0, outs)
    torch.cuda.synchronize()
    assert int(outs[0][-1]) == 0 and int(outs[2][-1]) == 1
