"""world_size-2 NCCL run of the sharded top-2 matcher (config 5's exchange step) — needs two GPUs on the box."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = textwrap.dedent("""
    import os, sys
    import numpy as np, torch, torch.distributed as dist
    sys.path.insert(0, %r)
    from orb_slam3_ros2_b200 import orbx, synth
    from oracle import pyoracle
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
    uid = [orbx.ShardedMatcher.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    sm = orbx.ShardedMatcher(rank, world, uid[0], lr)
    n_db, nq = 200_003, 257
    db = synth.descriptor_db(n_db)
    q, _ = synth.queries_from_db(nq, n_db)
    # planted exact duplicates of query rows on both sides of the shard boundary: ties must resolve by global index
    lo, hi = orbx.shard_range(n_db, rank, world)
    b = orbx.shard_range(n_db, 0, world)[1]
    db[b - 1] = q[0]; db[b] = q[0]; db[5] = q[1]; db[n_db - 1] = q[1]
    dq = torch.from_numpy(q).cuda(); ddb = torch.from_numpy(db[lo:hi].copy()).cuda()
    out = [torch.empty(nq, dtype=torch.int64, device="cuda"), torch.empty(nq, dtype=torch.int32, device="cuda"),
           torch.empty(nq, dtype=torch.int64, device="cuda"), torch.empty(nq, dtype=torch.int32, device="cuda")]
    sm.match(dq.data_ptr(), nq, ddb.data_ptr(), hi - lo, lo, *[o.data_ptr() for o in out])
    want = pyoracle.match_top2(q, db, threads=4)
    for g, w in zip(out, want):
        assert np.array_equal(g.cpu().numpy().astype(np.int64), np.asarray(w).astype(np.int64)), rank
    assert out[0][0].item() == b - 1 and out[2][0].item() == b and out[0][1].item() == 5 and out[2][1].item() == n_db - 1
    dist.barrier()
    sm.close()
    dist.destroy_process_group()
    print("rank", rank, "ok")
""")


def test_two_rank_nccl_allgather_merge_equals_global_top2(tmp_path):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    script = tmp_path / "worker.py"
    script.write_text(WORKER % ROOT)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29611", str(script)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == 2
