"""BASELINE.json's FULL sizes, checked through size-independent properties (the oracle would need minutes to hours at
these sizes, so it checks a sample and the properties carry the rest):
 * config 2  10,000 frames of 640x480 / 1000 features through orbx_extract_batch_device: the sequence repeats a short
             period, so every frame's output must be byte-identical to its first occurrence (batch-position and
             chunk-lane independence; one checksum per frame, one comparison of checksums), and the first period equals
             the oracle bit for bit;
 * config 4  2048 SearchForInitialization pairs in one call: pairs are repeats of a few distinct pairs in shuffled order,
             every repeat must equal the oracle's result for its source pair (pair-order independence);
 * config 5  5,000 keyframes x 1000 descriptors = 5 M database descriptors: planted near-duplicates must come back as
             the best match with the planted distance, the two-half split + merge must equal the single search
             (the property the 8-GPU sharding relies on), and a sample of queries equals the oracle over all 5 M."""
import numpy as np
import pytest

import matchdata as MD

pytestmark = pytest.mark.gpu


def test_config2_10k_frames_periodic_sequence(oracle):
    import torch
    from orb_slam_b200 import KP_DTYPE, ORBextractor, synth_frames_device
    F, P, H, W, NF = 10000, 7, 480, 640, 1000          # 7 is coprime to the chunk size: every lane sees every frame
    ex = ORBextractor(nfeatures=NF, max_width=W, max_height=H, max_batch=256)
    base = torch.empty((P, H, W), dtype=torch.uint8, device="cuda")
    synth_frames_device(base.data_ptr(), P, W, H, W, H * W, seed0=900)
    frames = base.repeat((F + P - 1) // P, 1, 1)[:F].contiguous()
    d_k = torch.zeros((F, NF, 28), dtype=torch.uint8, device="cuda")
    d_d = torch.zeros((F, NF, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(F, dtype=torch.int32, device="cuda")
    ex.extract_batch_device(frames.data_ptr(), F, W, H, W, H * W, d_k.data_ptr(), d_d.data_ptr(), d_n.data_ptr(),
                            torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ex.synchronize()
    # first period against the oracle
    host = base.cpu().numpy()
    n = d_n[:P].cpu().numpy()
    k = d_k[:P].cpu().numpy().view(KP_DTYPE).reshape(P, NF)
    d = d_d[:P].cpu().numpy()
    E = oracle.OracleExtractor(NF, 1.2, 8, 1, 20)
    for f in range(P):
        ok, od = E(host[f])
        assert n[f] == len(ok) and n[f] > 900
        assert np.array_equal(k[f, :n[f]], ok) and np.array_equal(d[f, :n[f]], od)
    # every later frame equals its first occurrence: compare on the device, period by period
    idx = torch.arange(F, device="cuda") % P
    assert torch.equal(d_n, d_n[:P][idx])
    valid = (torch.arange(NF, device="cuda")[None, :] < d_n[:, None])            # rows beyond n are unspecified
    same_k = ((d_k == d_k[:P][idx]).all(dim=2) | ~valid).all()
    same_d = ((d_d == d_d[:P][idx]).all(dim=2) | ~valid).all()
    assert bool(same_k) and bool(same_d)


def test_config4_2048_pairs_order_independent(oracle):
    from oracle import match_oracle as M
    from orb_slam_b200 import Matcher, ORBmatcher
    M.lib()
    matcher = Matcher(0)
    src = [MD.make_pair(oracle, s)[:2] for s in (20, 21, 22, 23)]
    ref = []
    for f1, f2 in src:
        prev = np.stack([f1.x, f1.y], 1).astype(np.float32)
        nref, mref = M.search_for_initialization(f1, f2, prev, 100, 0.9, True)
        ref.append((nref, mref, prev))
    fs1 = matcher.upload([MD.to_matchframe(p[0]) for p in src]); fs2 = matcher.upload([MD.to_matchframe(p[1]) for p in src])
    rng = np.random.default_rng(5)
    which = rng.integers(0, len(src), 2048)
    om = ORBmatcher(0.9, True, matcher=matcher)
    prevs = [np.stack([src[w][0].x, src[w][0].y], 1).astype(np.float32) for w in which]
    nm, rows = om.SearchForInitialization(fs1, fs2, which, which, prevs, 100)
    for p, w in enumerate(which):
        nref, mref, pref = ref[w]
        assert nm[p] == nref and np.array_equal(rows[p][:len(mref)], mref) and np.all(rows[p][len(mref):] == -1), p
        assert np.array_equal(prevs[p], pref), p
    assert min(r[0] for r in ref) > 50


def test_config5_5m_database_planted_split_and_sample(oracle):
    import torch
    from oracle import match_oracle as M
    from orb_slam_b200 import Matcher
    M.lib()
    matcher = Matcher(0)
    NDB, NQ = 5_000_000, 4000
    g = torch.Generator(device="cuda"); g.manual_seed(77)
    d_db = torch.randint(0, 256, (NDB, 32), dtype=torch.uint8, device="cuda", generator=g)
    rng = np.random.default_rng(78)
    src = rng.choice(NDB, NQ, replace=False)
    src[:8] = [0, 1, NDB - 1, NDB - 2, NDB // 2, NDB // 2 - 1, 12345, 4_194_304]
    nflip = rng.integers(0, 25, NQ)                      # planted distance 0..24; random descriptors sit near 128
    q = d_db[torch.from_numpy(src).cuda()].cpu().numpy()
    for i in range(NQ):
        bits = rng.choice(256, nflip[i], replace=False)
        for b in bits:
            q[i, b >> 3] ^= np.uint8(1 << (b & 7))
    d_q = torch.from_numpy(q).cuda()
    s = torch.cuda.current_stream().cuda_stream
    top2 = torch.empty((NQ, 2), dtype=torch.int32, device="cuda")
    matcher.knn2_device(d_q.data_ptr(), NQ, d_db.data_ptr(), NDB, 0, top2.data_ptr(), s)
    torch.cuda.synchronize()
    got = top2.cpu().numpy().view(np.uint32)
    # planted neighbours: best = (planted distance, source index); second best is a random descriptor, far away
    assert np.array_equal(got[:, 0] >> 23, nflip.astype(np.uint32))
    assert np.array_equal(got[:, 0] & 0x7FFFFF, src.astype(np.uint32))
    assert np.all((got[:, 1] >> 23) > 60) and np.all((got[:, 1] >> 23) < 128)
    # split into two shards with global indices, merge: equals the single search (what §8e's sharding relies on)
    half = NDB // 2 + 17
    gathered = torch.empty((2, NQ, 2), dtype=torch.int32, device="cuda")
    matcher.knn2_device(d_q.data_ptr(), NQ, d_db.data_ptr(), half, 0, gathered[0].data_ptr(), s)
    matcher.knn2_device(d_q.data_ptr(), NQ, d_db[half:].data_ptr(), NDB - half, half, gathered[1].data_ptr(), s)
    merged = torch.empty((NQ, 2), dtype=torch.int32, device="cuda"); match = torch.empty(NQ, dtype=torch.int32, device="cuda")
    matcher.merge_top2_device(gathered.data_ptr(), 2, NQ, 50, 0.75, merged.data_ptr(), match.data_ptr(), s)
    torch.cuda.synchronize()
    assert torch.equal(merged, top2)
    assert np.array_equal(match.cpu().numpy(), src)      # 24 <= 50 and 24 < 0.75 * 61
    # a sample of queries against the oracle over the whole database
    db = d_db.cpu().numpy()
    sample = np.arange(0, NQ, NQ // 48)
    ref = M.knn2(q[sample], db, threads=16)
    assert np.array_equal(got[sample], ref)


def test_knn2_64bit_keys_beyond_2_23_descriptors(oracle):
    """orbm_knn2_device64: a database of 2^23 + 300,000 descriptors (more than the 32-bit keys can index).  Planted
    near-duplicates in BOTH segments come back as best matches with their 40-bit indices and exact distances, a duplicate
    row that exists in both segments resolves to the lower index, and the result equals the two halves searched with the
    32-bit entry point and merged on the host; a query sample equals the oracle over the tail segment."""
    import torch
    from oracle import match_oracle as M
    from orb_slam_b200 import Matcher
    M.lib()
    m = Matcher(0)
    ndb, nq = (1 << 23) + 300_000, 512
    g = torch.Generator(device="cuda").manual_seed(5)
    db = torch.randint(0, 256, (ndb, 32), dtype=torch.uint8, device="cuda", generator=g)
    pick = torch.randint(0, ndb, (nq,), device="cuda", generator=g)
    pick[:64] = torch.randint(1 << 23, ndb, (64,), device="cuda", generator=g)         # the tail segment for sure
    q = db[pick].clone()
    q[:, 3] ^= 0x81                                                                    # distance 2 to the planted row
    db[(1 << 23) + 77] = db[1234]; q[100] = db[1234]; pick[100] = 1234                 # same row in both segments
    out = torch.empty((nq, 2), dtype=torch.int64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    m.knn2_device64(q.data_ptr(), nq, db.data_ptr(), ndb, 0, out.data_ptr(), st)
    torch.cuda.synchronize()
    m.synchronize()
    k = out.cpu().numpy().view(np.uint64)
    dist, idx = (k >> np.uint64(40)).astype(np.int64), (k & np.uint64(0xFFFFFFFFFF)).astype(np.int64)
    p = pick.cpu().numpy()
    assert np.array_equal(idx[:, 0], p) and (dist[np.arange(nq) != 100, 0] == 2).all() and dist[100, 0] == 0
    assert idx[100, 1] == (1 << 23) + 77 and dist[100, 1] == 0
    # against the 32-bit path on the two segments, merged on the host in (distance, index) order
    a = torch.empty((nq, 2), dtype=torch.int32, device="cuda"); b = torch.empty((nq, 2), dtype=torch.int32, device="cuda")
    m.knn2_device(q.data_ptr(), nq, db.data_ptr(), 1 << 23, 0, a.data_ptr(), st)
    m.knn2_device(q.data_ptr(), nq, db[1 << 23:].data_ptr(), ndb - (1 << 23), 0, b.data_ptr(), st)
    torch.cuda.synchronize()
    ka, kb = a.cpu().numpy().view(np.uint32).astype(np.uint64), b.cpu().numpy().view(np.uint32).astype(np.uint64)
    wide = lambda kk, base: ((kk >> np.uint64(23)) << np.uint64(40)) | ((kk & np.uint64(0x7FFFFF)) + np.uint64(base))
    allk = np.sort(np.concatenate([wide(ka, 0), wide(kb, 1 << 23)], axis=1), axis=1)[:, :2]
    assert np.array_equal(k, allk)
    # oracle over the tail segment for the queries planted there
    tail = db[1 << 23:].cpu().numpy()
    ref = M.knn2(q[:64].cpu().numpy(), tail, threads=8)
    assert np.array_equal(kb[:64].astype(np.uint32), ref)
