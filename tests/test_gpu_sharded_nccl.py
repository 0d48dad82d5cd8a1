"""Sharded all-pairs 2-NN over NCCL on >= 2 GPUs (one process per GPU): orbm_comm_init_rank + orbm_knn2_sharded
(local search with global indices, ncclAllGather, merge kernel) == the oracle over the whole database.  Skipped on a
single-GPU box; tests/test_sharded_gloo.py covers the host logic on CPU."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, ok):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import match_oracle as MO
    from orb_slam_b200 import Matcher, ShardedKnn2, shard_range
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world, device_id=dev)
    rng = np.random.default_rng(5)
    nkf, per, nq = 37, 500, 1500
    db = rng.integers(0, 256, (nkf * per, 32), dtype=np.uint8)
    q = db[rng.integers(0, len(db), nq)].copy()
    q ^= np.packbits(rng.random((nq, 256)) < 0.08, axis=1)
    a, b = shard_range(nkf, rank, world)
    m = Matcher(rank)
    sh = ShardedKnn2.from_torch_distributed(m, dev)
    d_db = torch.from_numpy(db[a * per:b * per].copy()).to(dev); d_q = torch.from_numpy(q).to(dev)
    gathered = torch.empty((world, nq, 2), dtype=torch.int32, device=dev)
    top2 = torch.empty((nq, 2), dtype=torch.int32, device=dev); match = torch.empty(nq, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        sh.search(d_q.data_ptr(), nq, d_db.data_ptr(), (b - a) * per, a * per, gathered.data_ptr(), top2.data_ptr(),
                  match.data_ptr(), 50, 0.75, s)
    torch.cuda.synchronize()
    ref = MO.knn2(q, db, threads=8)
    got = top2.cpu().numpy().view(np.uint32)
    bd, sd = (ref[:, 0] >> 23).astype(np.int64), (ref[:, 1] >> 23).astype(np.int64)
    acc = (bd <= 50) & (bd.astype(np.float32) < np.float32(0.75) * sd.astype(np.float32))
    exp = np.where(acc, (ref[:, 0] & 0x7FFFFF).astype(np.int64), -1)
    ok[rank] = int(np.array_equal(got, ref) and np.array_equal(match.cpu().numpy(), exp) and acc.sum() > 100)
    sh.close()
    dist.destroy_process_group()


def test_knn2_sharded_nccl():
    import torch
    import torch.multiprocessing as mp
    world = min(torch.cuda.device_count(), 4)
    if world < 2:
        pytest.skip("needs >= 2 GPUs")
    ctx = mp.get_context("spawn")
    ok = ctx.Array("i", [0] * world)
    port = 29600 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, world, port, ok)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(300)
        assert p.exitcode == 0
    assert list(ok) == [1] * world
