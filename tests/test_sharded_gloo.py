"""world_size-2 (gloo, CPU) test of the sharded matcher's exchange protocol (config 5): every rank owns
rows [r*N/G, (r+1)*N/G) of the database, computes a local top-2 with GLOBAL row numbers packed as
key = distance << 40 | row, all-gathers the 2 x nq keys per rank and merges by plain ordering of the keys — which must
reproduce cv::BFMatcher::knnMatch's (distance asc, trainIdx asc) order across shard boundaries (Frame.cc:1144).

On the GPU the local top-2 and the merge are CUDA kernels and the exchange is ncclAllGather (orbx_match_top2_sharded);
here the local top-2 comes from the oracle (this is a test) so that the HOST logic — shard ranges, key packing,
gather layout, merge rule, empty shards — is covered without a GPU."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT  # noqa: F401  (puts the repo root on sys.path)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_problem(nd, nq, seed):
    rng = np.random.default_rng(seed)
    db = rng.integers(0, 256, (nd, 32), dtype=np.uint8)
    q = db[rng.integers(0, nd, nq)].copy()
    flips = rng.integers(0, 256, (nq, 6))
    for i in range(nq):
        for b in flips[i]:
            q[i, b // 8] ^= np.uint8(1 << (b % 8))
    # exact duplicates planted on both sides of the shard boundary: ties on (distance) must resolve to the lower row
    half = nd // 2
    if half >= 4:
        db[half - 1] = db[half + 2] = db[3]
        q[0] = db[3]
    return q, db


def _worker(rank, world, port, nd, nq, seed, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import pyoracle
        from orb_slam3_ros2_b200 import orbx
        q, db = _make_problem(nd, nq, seed)
        lo, hi = orbx.shard_range(nd, rank, world)
        no_key = np.uint64(0xFFFFFFFFFFFFFFFF)
        k1 = np.full(nq, no_key, dtype=np.uint64)
        k2 = np.full(nq, no_key, dtype=np.uint64)
        if hi > lo:
            i1, d1, i2, d2 = [np.asarray(a, dtype=np.int64) for a in pyoracle.match_top2(q, db[lo:hi])]
            k1 = np.where(i1 >= 0, (d1.astype(np.uint64) << np.uint64(40)) | (i1 + lo).astype(np.uint64), no_key)
            k2 = np.where(i2 >= 0, (d2.astype(np.uint64) << np.uint64(40)) | (i2 + lo).astype(np.uint64), no_key)
        # payload layout of orbx_match_top2_sharded: [k1[nq] | k2[nq]] per rank
        mine = torch.from_numpy(np.concatenate([k1, k2]).view(np.int64))
        gathered = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(gathered, mine)
        g = np.stack([t.numpy().view(np.uint64) for t in gathered])  # [rank][2*nq]
        m1, m2 = orbx.merge_top2_host(g[:, :nq], g[:, nq:])
        np.save(os.path.join(out_dir, f"m_{rank}.npy"), np.stack([m1, m2]))
    finally:
        dist.destroy_process_group()


def _run(world, nd, nq, seed, tmp_path):
    port = _free_port()
    mp.spawn(_worker, args=(world, port, nd, nq, seed, str(tmp_path)), nprocs=world, join=True)
    from oracle import pyoracle
    q, db = _make_problem(nd, nq, seed)
    i1, d1, i2, d2 = [np.asarray(a, dtype=np.int64) for a in pyoracle.match_top2(q, db)]
    mask = np.uint64((1 << 40) - 1)
    for r in range(world):
        m = np.load(os.path.join(str(tmp_path), f"m_{r}.npy"))
        for k, (wi, wd) in enumerate(((i1, d1), (i2, d2))):
            valid = wi >= 0
            assert np.array_equal((m[k] & mask).astype(np.int64)[valid], wi[valid]), f"rank {r} idx{k + 1}"
            assert np.array_equal((m[k] >> np.uint64(40)).astype(np.int64)[valid], wd[valid]), f"rank {r} d{k + 1}"
            assert np.all(m[k][~valid] == np.uint64(0xFFFFFFFFFFFFFFFF))
    return i1


def test_two_ranks_merge_equals_global_top2(tmp_path, oracle):
    i1 = _run(2, 4001, 97, 5, tmp_path)
    assert i1[0] == 3  # the planted duplicates at rows 3, nd/2-1, nd/2+2: lowest row wins across the boundary


def test_two_ranks_tiny_database(tmp_path, oracle):
    # nd = 1: rank 0 owns nothing ([0,0)), rank 1 owns the only row; second-best is missing everywhere
    _run(2, 1, 5, 9, tmp_path)


def test_shard_ranges_partition():
    from orb_slam3_ros2_b200 import orbx
    for n in (0, 1, 7, 10_000_000):
        for g in (1, 2, 4, 8):
            r = [orbx.shard_range(n, k, g) for k in range(g)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(g - 1))
            assert max(b - a for a, b in r) - min(b - a for a, b in r) <= 1
