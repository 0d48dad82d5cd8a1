"""N>1 path on CPU (gloo, world_size 2): the keyframe-sharded all-pairs 2-NN exchange of SURVEY §8e.  Each rank
searches its contiguous shard (product's shard_range) with GLOBAL indices using the CPU oracle, the packed [nq,2]
keys are all-gathered exactly as ncclAllGather lays them out, and the min / second-min merge (test-side numpy
restatement of merge_top2_kernel) must equal one search over the whole database.  The CUDA kernels themselves are
covered by tests/test_gpu_matcher.py::test_knn2_sharded_merge_single_gpu."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def merge_top2(gathered):
    """gathered [world, nq, 2] uint32 -> [nq, 2]: two smallest keys per query."""
    w, nq, _ = gathered.shape
    flat = np.sort(gathered.transpose(1, 0, 2).reshape(nq, 2 * w), axis=1)
    return flat[:, :2]


def _worker(rank, world, port, ok):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import match_oracle as MO
    from orb_slam_b200 import shard_range
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    rng = np.random.default_rng(77)  # same data on every rank
    nkf, per = 11, 40                # 11 keyframes do not divide evenly over 2 ranks
    db = rng.integers(0, 256, (nkf * per, 32), dtype=np.uint8)
    q = db[rng.integers(0, len(db), 64)].copy()
    q ^= np.packbits(rng.random((64, 256)) < 0.1, axis=1)
    db[300] = db[100]; q[0] = db[100]            # duplicate across the shard boundary: lower global index wins
    a, b = shard_range(nkf, rank, world)
    local = MO.knn2(q, db[a * per:b * per], idx_base=a * per)
    t = torch.from_numpy(local.view(np.int32).copy())
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t)
    gathered = np.stack([o.numpy().view(np.uint32) for o in out])
    merged = merge_top2(gathered)
    ref = MO.knn2(q, db)
    ok[rank] = int(np.array_equal(merged, ref) and merged[0, 0] == 100 and merged[0, 1] == 300)
    dist.destroy_process_group()


def test_sharded_knn2_gloo_world2():
    from oracle import oracle as O
    O.lib()  # build once before forking workers
    ctx = mp.get_context("spawn")
    ok = ctx.Array("i", [0, 0])
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ok)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    assert list(ok) == [1, 1]


def test_shard_range_covers():
    from orb_slam_b200 import shard_range
    for n in (1, 7, 5000):
        for w in (1, 2, 4, 8):
            r = [shard_range(n, k, w) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[k][1] == r[k + 1][0] for k in range(w - 1))
