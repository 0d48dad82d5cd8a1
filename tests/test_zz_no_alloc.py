"""orbm_knn2_device_static (include/orbm.h): the all-pairs 2-NN with its scratch in a file-scope device array, for
callers that may not allocate.  No counterpart in the reference.  Checked through the C ABI
against the existing entry point orbm_knn2_device: packed keys are integers, so the comparison is bit-exact.  Two
calls in a row with different shapes: the array keeps the previous call's contents, the second result must not
depend on them.  Collected last (file name) so that a failure here hides nothing."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(m, torch, nq, ndb, seed, idx_base=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    db = torch.randint(0, 256, (max(ndb, 1), 32), dtype=torch.uint8, device="cuda", generator=g)
    q = torch.randint(0, 256, (nq, 32), dtype=torch.uint8, device="cuda", generator=g)
    if ndb and nq:                                        # planted near-duplicates and exact ties
        pick = torch.randint(0, ndb, (nq,), device="cuda", generator=g)
        near = db[pick].clone()
        near[:, 0] ^= 0x11
        q[::3] = near[::3]
        if ndb > 8:
            db[ndb // 2] = db[ndb // 4]                   # duplicate row: the lower index must win
    ref = torch.full((nq, 2), 0x5A5A5A5A, dtype=torch.int32, device="cuda")      # packed u32 keys, viewed below
    out = torch.full((nq, 2), 0x3C3C3C3C, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    m.knn2_device(q.data_ptr(), nq, db.data_ptr(), ndb, idx_base, ref.data_ptr(), st)
    m.knn2_device_static(q.data_ptr(), nq, db.data_ptr(), ndb, idx_base, out.data_ptr(), st)
    torch.cuda.synchronize()
    return ref.cpu().numpy().view(np.uint32), out.cpu().numpy().view(np.uint32)


def test_static_scratch_knn2_equals_allocating_entry_point_twice():
    import torch
    from orb_slam_b200 import Matcher
    m = Matcher(0)
    # two calls in a row: a large one (4 queries per thread, many chunks) that fills the scratch, then smaller ones
    # of other shapes (2 and 1 queries per thread, single-tile chunks) that must not see its leftovers
    for nq, ndb, seed, base in ((3000, 200_000, 1, 0), (700, 50_000, 2, 1000), (333, 1000, 3, 0), (1, 1, 4, 7)):
        ref, out = _pair(m, torch, nq, ndb, seed, base)
        assert np.array_equal(ref, out), (nq, ndb)
        assert (ref[:, 0] != 0xFFFFFFFF).all()
    # the same first shape again after the small ones
    ref, out = _pair(m, torch, 3000, 200_000, 1, 0)
    assert np.array_equal(ref, out)
    # empty database: all keys "none"
    ref, out = _pair(m, torch, 5, 0, 5)
    assert np.array_equal(ref, out) and (out == 0xFFFFFFFF).all()


def test_static_scratch_rejects_oversized_calls_without_launching():
    import torch
    from orb_slam_b200 import Matcher, OrbmError
    m = Matcher(0)
    nq = 600_000                                          # 256 + nchunks * nq * 8 > 4 MiB whatever the chunk plan
    q = torch.zeros((nq, 32), dtype=torch.uint8, device="cuda")
    db = torch.zeros((1024, 32), dtype=torch.uint8, device="cuda")
    out = torch.zeros((nq, 2), dtype=torch.int32, device="cuda")
    before = m.kernel_launches()
    with pytest.raises(OrbmError) as e:
        m.knn2_device_static(q.data_ptr(), nq, db.data_ptr(), 1024, 0, out.data_ptr(), None)
    assert e.value.code == -3 and "scratch" in str(e.value)
    assert m.kernel_launches() == before
    assert int(out.abs().sum()) == 0
