"""Parity of the CUDA matcher (through the C ABI, include/orbm.h) with the CPU oracle.  Everything is integer / index
work, so every comparison is bit-exact (match indices, counts, updated vbPrevMatched, packed 2-NN keys)."""
import numpy as np
import pytest

import matchdata as MD

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def M(oracle):
    from oracle import match_oracle
    match_oracle.lib()
    return match_oracle


def _matcher_with(env):
    """A matcher handle created under the given run-time switches (they are read once, at orbm_create)."""
    import os
    from orb_slam_b200 import Matcher
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return Matcher(0)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


# every test of this file runs under the default commit choice (fixed point for <= 512 problems), under the in-order
# replay (commit_kernel) and under the fixed point forced; the 2-NN tests under each distance-arithmetic variant of the
# POPC kernels (tensor-core kernel off) and with the tensor-core kernel forced for every size
@pytest.fixture(scope="module", params=["auto", "serial+knn0", "parallel+knn1", "knn-tc"])
def matcher(request):
    env = {"auto": {}, "serial+knn0": {"ORBM_COMMIT": "serial", "ORBM_KNN_VARIANT": "0", "ORBM_KNN_TC": "0"},
           "parallel+knn1": {"ORBM_COMMIT": "parallel", "ORBM_KNN_VARIANT": "1", "ORBM_KNN_TC": "0"},
           "knn-tc": {"ORBM_KNN_TC": "1"}}[request.param]
    return _matcher_with(env)


@pytest.fixture(scope="module")
def pairs(oracle):
    """6 frame pairs: 640x480/1000 kp, the 2000-kp initialisation extractor, a small odd size."""
    out = [MD.make_pair(oracle, s) for s in (0, 1, 2)]
    out.append(MD.make_pair(oracle, 3, 640, 480, 2000))
    out.append(MD.make_pair(oracle, 4, 333, 251, 700))
    out.append(MD.make_pair(oracle, 5, 640, 480, 1000, shift=(0, 0)))
    return out


def _rows_equal(row, ref):
    return np.array_equal(row[: len(ref)], ref) and np.all(row[len(ref):] == -1)


def test_descriptor_distance(M, matcher):
    rng = np.random.default_rng(0)
    a = rng.integers(0, 256, (5000, 32), dtype=np.uint8); b = rng.integers(0, 256, (5000, 32), dtype=np.uint8)
    b[:10] = a[:10]; b[10] = ~a[10]
    got = matcher.descriptor_distance(a, b)
    ref = np.unpackbits(a ^ b, axis=1).sum(1)
    assert np.array_equal(got, ref) and got[10] == 256 and got[0] == 0
    assert all(int(got[i]) == M.descriptor_distance(a[i], b[i]) for i in range(0, 5000, 97))


def test_grid_and_area(M, matcher, pairs):
    frames = [p[1] for p in pairs]
    fs = matcher.upload([MD.to_matchframe(f) for f in frames])
    rng = np.random.default_rng(1)
    for k, f in enumerate(frames):
        cs, items = fs.grid(k)
        ocs, oitems = M.grid_csr(f)
        assert np.array_equal(cs, ocs) and np.array_equal(items, oitems)
        for _ in range(40):
            x, y, r = float(rng.uniform(-30, f.max_x + 30)), float(rng.uniform(-30, f.max_y + 30)), float(rng.uniform(0.5, 120))
            mode = int(rng.integers(0, 3))
            lv = (-1, -1) if mode == 0 else ((1, 1) if mode == 1 else (int(rng.integers(-1, 3)), int(rng.integers(3, 7))))
            assert np.array_equal(fs.features_in_area(k, x, y, r, *lv), M.features_in_area(f, x, y, r, *lv))


@pytest.mark.parametrize("ori", [True, False])
def test_search_for_initialization(M, matcher, pairs, ori):
    from orb_slam_b200 import ORBmatcher
    om = ORBmatcher(0.9, ori, matcher=matcher)
    s1 = om.upload([MD.to_matchframe(p[0]) for p in pairs]); s2 = om.upload([MD.to_matchframe(p[1]) for p in pairs])
    idx = list(range(len(pairs)))
    prev = [np.stack([p[0].x, p[0].y], 1).astype(np.float32) for p in pairs]
    oprev = [a.copy() for a in prev]
    nm, rows = om.SearchForInitialization(s1, s2, idx, idx, prev, 100)
    total = 0
    for k, (f1, f2, _) in enumerate(pairs):
        on, om12 = M.search_for_initialization(f1, f2, oprev[k], 100, 0.9, ori)
        assert nm[k] == on and _rows_equal(rows[k], om12), f"pair {k}"
        assert np.array_equal(prev[k], oprev[k])
        total += on
    assert total > 300
    # second call continues from the updated vbPrevMatched with a smaller window (as the tracker would)
    nm2, rows2 = om.SearchForInitialization(s1, s2, idx, idx, prev, 20)
    for k, (f1, f2, _) in enumerate(pairs):
        on, om12 = M.search_for_initialization(f1, f2, oprev[k], 20, 0.9, ori)
        assert nm2[k] == on and _rows_equal(rows2[k], om12) and np.array_equal(prev[k], oprev[k])


def test_window_search_and_projection_frames(M, matcher, pairs):
    from orb_slam_b200 import ORBmatcher
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.6, 10 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.2, 20 + k)) for k, p in enumerate(pairs)]
    s1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
    idx = list(range(len(pairs)))
    for (win, ratio, ori, lo, hi) in [(200, 0.8, True, 0, 2**31 - 1), (100, 0.9, False, 1, 3), (50, 0.7, True, 0, 0)]:
        om = ORBmatcher(ratio, ori, matcher=matcher)
        nm, rows = om.WindowSearch(s1, s2, idx, idx, win, lo, hi)
        for k in idx:
            on, orow = M.window_search(fr1[k], fr2[k], win, ratio, ori, lo, hi)
            assert nm[k] == on and _rows_equal(rows[k], orow), (win, k)
        assert nm.sum() > 100
    om = ORBmatcher(0.9, True, matcher=matcher)
    for win in (15, 50):
        uv = [np.stack([f.x + np.float32(p[2][0] * -1.0) + 1.5, f.y - np.float32(p[2][1]) - 0.75], 1).astype(np.float32)
              for f, p in zip(fr1, pairs)]
        nm, rows = om.SearchByProjectionFrames(s1, s2, idx, idx, uv, win)
        for k in idx:
            on, orow = M.search_by_projection_ff(fr1[k], fr2[k], uv[k], win, 0.9)
            assert nm[k] == on and _rows_equal(rows[k], orow), (win, k)
        assert nm.sum() > 100


def _points_from(f1, shift, rng, kind, scale_factors):
    from orb_slam_b200 import ProjectedPoints
    n = f1.n
    u = (f1.x - np.float32(shift[0]) + rng.uniform(-2, 2, n)).astype(np.float32)
    v = (f1.y - np.float32(shift[1]) + rng.uniform(-2, 2, n)).astype(np.float32)
    valid = (rng.random(n) < 0.8).astype(np.uint8)
    if kind == "local_map":
        cos = rng.uniform(0.99, 1.0, n).astype(np.float32)
        return ProjectedPoints.for_local_map(u, v, cos, f1.octave, f1.desc, scale_factors, th=float(rng.choice([1.0, 5.0])),
                                             valid=valid)
    return ProjectedPoints.for_last_frame(u, v, f1.octave, f1.desc, f1.angle, scale_factors, 15.0, valid=valid)


@pytest.mark.parametrize("variant,ratio,ori,th", [("local_map", 0.8, False, 100), ("best", 0.9, True, 100),
                                                  ("best", 0.9, False, 50)])
def test_search_by_projection(M, matcher, pairs, variant, ratio, ori, th):
    from orb_slam_b200 import ORBmatcher
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    rng = np.random.default_rng(5)
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.15, 30 + k)) for k, p in enumerate(pairs)]
    s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
    pts = [_points_from(p[0], p[2], rng, variant, sf) for p in pairs]
    om = ORBmatcher(ratio, ori, matcher=matcher)
    idx = list(range(len(pairs)))
    nm, rows = om.SearchByProjection(s2, idx, pts, variant, th)
    for k in idx:
        on, orow = M.search_by_projection(fr2[k], pts[k], variant, th, ratio, ori)
        assert nm[k] == on and _rows_equal(rows[k], orow), k
    assert nm.sum() > 200


@pytest.mark.parametrize("variant,ratio,ori", [("kf_f", 0.75, True), ("kf_f", 0.6, False), ("kf_kf", 0.75, True)])
def test_search_by_bow(M, matcher, pairs, variant, ratio, ori):
    from orb_slam_b200 import ORBmatcher
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.7, 40 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.8, 50 + k)) for k, p in enumerate(pairs)]
    s1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
    om = ORBmatcher(ratio, ori, matcher=matcher)
    # also cross pairs (k, k+1): mostly non-matching content
    i1 = list(range(len(pairs))) + [0, 1]
    i2 = list(range(len(pairs))) + [1, 2]
    nm, rows = om.SearchByBoW(s1, s2, i1, i2, variant)
    for p, (a, b) in enumerate(zip(i1, i2)):
        on, orow = M.search_by_bow(fr1[a], fr2[b], variant, ratio, ori)
        assert nm[p] == on and _rows_equal(rows[p], orow), (variant, p)
    assert nm[: len(pairs)].sum() > 100


@pytest.mark.parametrize("ori", [True, False])
def test_search_for_triangulation(M, matcher, pairs, ori):
    from orb_slam_b200 import ORBmatcher
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.3, 60 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.3, 70 + k)) for k, p in enumerate(pairs)]
    s1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    sigma2 = (sf * sf).astype(np.float32)
    # fundamental matrix of a pure image translation (dx, dy): the epipolar line of x1 passes through x1 - shift
    F12 = []
    for p in pairs:
        dx, dy = p[2]
        t = np.array([dx, dy, 0.0]) + 1e-3
        F12.append(np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) * np.float32(0.05))
    om = ORBmatcher(0.6, ori, matcher=matcher)
    idx = list(range(len(pairs)))
    nm, rows = om.SearchForTriangulation(s1, s2, idx, idx, np.stack(F12), sigma2)
    for k in idx:
        on, orow = M.search_for_triangulation(fr1[k], fr2[k], F12[k], sigma2, ori)
        assert nm[k] == on and _rows_equal(rows[k], orow), k
    assert nm.sum() > 50


def test_knn2(M, matcher):
    rng = np.random.default_rng(11)
    for nq, ndb in [(1000, 20000), (300, 5000), (37, 129), (600, 1), (5, 4097)]:
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        q = db[rng.integers(0, ndb, nq)].copy()
        flip = rng.random((nq, 256)) < 0.1
        q ^= np.packbits(flip, axis=1)
        if ndb > 100:
            db[77] = db[33]; q[0] = db[33]  # duplicate: lowest index wins, duplicate is the second best
        got = matcher.knn2(q, db, idx_base=1000)
        ref = M.knn2(q, db, idx_base=1000, threads=8)
        assert np.array_equal(got, ref), (nq, ndb)
    assert matcher.last_kernel_ms() > 0


def test_knn2_tensor_core_kernel_edge_cases(M):
    """knn2_tc_kernel (forced with ORBM_KNN_TC=1) against the oracle on the shapes and contents its schedule and its
    filter have to survive: query counts around the 128 / 256 rows of a CTA / a pair, databases around the 256-descriptor
    tile and smaller than one, one cluster range crossing several query-tile pairs (many queries, few tiles) and many
    clusters sharing one pair (few queries, many tiles), a database of identical descriptors (every distance ties: the
    lowest indices must win), a database far from every query (all distances > 128: the dot products are negative and
    the integer-max shortcut must not be trusted), exact duplicates, an index base."""
    tc = _matcher_with({"ORBM_KNN_TC": "1"})
    pop = _matcher_with({"ORBM_KNN_TC": "0"})
    rng = np.random.default_rng(2024)
    shapes = [(1, 1), (1, 255), (127, 256), (128, 257), (129, 511), (255, 512), (256, 513), (257, 1000), (385, 3000),
              (5000, 300), (3000, 40), (60, 70000), (2, 150000), (1000, 20011)]
    for nq, ndb in shapes:
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        q = db[rng.integers(0, ndb, nq)].copy()
        q ^= np.packbits(rng.random((nq, 256)) < 0.07, axis=1)
        q[::5] = db[rng.integers(0, ndb, len(q[::5]))]           # exact copies: distance 0
        ref = M.knn2(q, db, idx_base=37, threads=8)
        got = tc.knn2(q, db, idx_base=37)
        assert np.array_equal(got, ref), (nq, ndb)
        assert np.array_equal(pop.knn2(q, db, idx_base=37), ref), (nq, ndb)
    # every database descriptor identical: all distances tie, indices 0 and 1 win for every query
    db = np.tile(rng.integers(0, 256, (1, 32), dtype=np.uint8), (3000, 1))
    q = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    got = tc.knn2(q, db)
    assert np.array_equal(got, M.knn2(q, db, threads=8))
    assert np.all((got[:, 0] & 0x7FFFFF) == 0) and np.all((got[:, 1] & 0x7FFFFF) == 1)
    # a database on the far side of EVERY query (queries: upper 128 bits clear; database: upper 128 bits set but for a
    # few): every distance is above 128, every dot product negative
    q = rng.integers(0, 256, (500, 32), dtype=np.uint8); q[:, 16:] = 0
    far = rng.integers(0, 256, (20000, 32), dtype=np.uint8)
    far[:, 16:] = 0xFF ^ np.packbits(rng.random((20000, 128)) < 0.04, axis=1)
    far[123] = ~q[7]                                             # distance exactly 256 to query 7
    ref = M.knn2(q, far, threads=8)
    assert np.array_equal(tc.knn2(q, far), ref)
    assert (ref[:, 0] >> 23).min() > 128
    # mixed: half of the queries only have far neighbours, the others a near one late in the database - lanes of one
    # warp with and without a usable threshold (a vote under that per-lane condition once hung the kernel)
    mixed = far.copy(); mixed[15000:15250] = q[:250]
    assert np.array_equal(tc.knn2(q, mixed), M.knn2(q, mixed, threads=8))
    rng.shuffle(q)
    assert np.array_equal(tc.knn2(q, mixed), M.knn2(q, mixed, threads=8))


def test_knn2_under_cuda_graph_capture(M):
    """The device entry points only enqueue work, so they can be recorded into a CUDA graph once their scratch has been
    reserved by a first call: the tensor-core kernel (a cluster launch + a memset), the POPC kernels and the
    static-scratch entry point are captured in one graph, the graph is replayed on new data in the same buffers and
    every result must equal the oracle."""
    import torch
    rng = np.random.default_rng(77)
    nq, ndb = 700, 30000
    tc, pop = _matcher_with({"ORBM_KNN_TC": "1"}), _matcher_with({"ORBM_KNN_TC": "0"})
    d_q = torch.zeros((nq, 32), dtype=torch.uint8, device="cuda"); d_db = torch.zeros((ndb, 32), dtype=torch.uint8, device="cuda")
    outs = [torch.zeros((nq, 2), dtype=torch.int32, device="cuda") for _ in range(3)]
    st = torch.cuda.Stream()

    def enqueue(sp):
        tc.knn2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, 5, outs[0].data_ptr(), sp)
        pop.knn2_device(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, 5, outs[1].data_ptr(), sp)
        pop.knn2_device_static(d_q.data_ptr(), nq, d_db.data_ptr(), ndb, 5, outs[2].data_ptr(), sp)

    with torch.cuda.stream(st):
        enqueue(st.cuda_stream)                                # warm-up: the handles reserve their scratch here
    st.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=st):
        enqueue(torch.cuda.current_stream().cuda_stream)
    for seed in (1, 2):
        db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
        q = db[rng.integers(0, ndb, nq)] ^ np.packbits(rng.random((nq, 256)) < 0.08, axis=1)
        d_db.copy_(torch.from_numpy(db)); d_q.copy_(torch.from_numpy(q))
        for o in outs:
            o.zero_()
        torch.cuda.synchronize()
        g.replay()
        torch.cuda.synchronize()
        ref = M.knn2(q, db, idx_base=5, threads=8)
        for k, o in enumerate(outs):
            assert np.array_equal(o.cpu().numpy().view(np.uint32), ref), (seed, k)
    tc.synchronize(); pop.synchronize()                        # raise on a deferred kernel error


def test_knn2_sharded_merge_single_gpu(M, matcher):
    """The multi-GPU merge path on one device: 4 database shards searched separately, results laid out like the
    all-gather receive buffer, merged by orbm_merge_top2_device == one search over the whole database."""
    import torch
    rng = np.random.default_rng(12)
    ndb, nq, shards = 40000, 1000, 4
    db = rng.integers(0, 256, (ndb, 32), dtype=np.uint8)
    q = db[rng.integers(0, ndb, nq)].copy()
    q ^= np.packbits(rng.random((nq, 256)) < 0.08, axis=1)
    d_db = torch.from_numpy(db).cuda(); d_q = torch.from_numpy(q).cuda()
    gathered = torch.empty((shards, nq, 2), dtype=torch.int32, device="cuda")
    per = ndb // shards
    s = torch.cuda.current_stream().cuda_stream
    for r in range(shards):
        matcher.knn2_device(d_q.data_ptr(), nq, d_db[r * per:].data_ptr(), per, r * per, gathered[r].data_ptr(), s)
    top2 = torch.empty((nq, 2), dtype=torch.int32, device="cuda"); match = torch.empty(nq, dtype=torch.int32, device="cuda")
    matcher.merge_top2_device(gathered.data_ptr(), shards, nq, 50, 0.75, top2.data_ptr(), match.data_ptr(), s)
    torch.cuda.synchronize()
    ref = M.knn2(q, db, threads=8)
    got = top2.cpu().numpy().view(np.uint32)
    assert np.array_equal(got, ref)
    bd, sd = (ref[:, 0] >> 23).astype(np.int64), (ref[:, 1] >> 23).astype(np.int64)
    ok = (bd <= 50) & (bd.astype(np.float32) < np.float32(0.75) * sd.astype(np.float32))
    exp = np.where(ok, (ref[:, 0] & 0x7FFFFF).astype(np.int64), -1)
    assert np.array_equal(match.cpu().numpy(), exp) and ok.sum() > 500


def test_distinctive_descriptors(M, matcher):
    """N4: batched MapPoint::ComputeDistinctiveDescriptors equals the oracle (index and chosen descriptor)."""
    rng = np.random.default_rng(21)
    sizes = np.concatenate([[1, 2, 3, 0, 32, 33, 64, 65, 300, 0], rng.integers(1, 40, 3000)])
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, (len(sizes), 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    flips = rng.random((len(desc), 256)) < rng.uniform(0.0, 0.2, (len(desc), 1))
    desc = np.packbits(np.unpackbits(desc, axis=1) ^ flips, axis=1)
    best, out = matcher.distinctive_descriptors(desc, start)
    ref = M.distinctive_descriptors(desc, start)
    assert np.array_equal(best, ref)
    has = ref >= 0
    assert np.array_equal(out[has], desc[start[:-1][has] + ref[has]]) and not out[~has].any()
    # ties everywhere: identical observations -> index 0
    same = np.repeat(base[:5], 9, axis=0)
    b2, _ = matcher.distinctive_descriptors(same, np.arange(0, 46, 9))
    assert list(b2) == [0] * 5
    b3, o3 = matcher.distinctive_descriptors(np.zeros((0, 32), np.uint8), [0, 0, 0])
    assert list(b3) == [-1, -1]


def test_cuda_equals_reference_generated_golden(matcher):
    """The CUDA matcher against vectors produced by the reference's OWN Frame / KeyFrame / MapPoint / ORBmatcher code
    (tests/golden/make_ref_match_golden.py): grid, GetFeaturesInArea, SearchForInitialization incl. the updated
    vbPrevMatched, WindowSearch, SearchByBoW (both overloads)."""
    import refgolden as RG
    from orb_slam_b200 import ORBmatcher
    g = RG.load()
    for s in range(int(g["nscenes"])):
        fr = [RG.frame(g, s, 0), RG.frame(g, s, 1)]
        fs = matcher.upload(fr)
        for f in (0, 1):
            p = f"s{s}_f{f}_"
            cs, it = fs.grid(f)
            assert np.array_equal(cs, g[p + "cell_start"]) and np.array_equal(it, g[p + "grid_items"])
            st = g[p + "area_start"]
            for k, ((x, y, r), (l0, l1)) in enumerate(zip(g[p + "area_q"], g[p + "area_lv"])):
                got = fs.features_in_area(f, float(x), float(y), float(r), int(l0), int(l1))
                assert np.array_equal(got, g[p + "area_items"][st[k]:st[k + 1]]), (s, f, k)
        prev0 = np.stack([fr[0].x, fr[0].y], 1)
        for k, (win, ratio, ori) in enumerate(RG.SFI):
            om = ORBmatcher(ratio, bool(ori), matcher=matcher)
            prev = [prev0.copy()]
            nm, rows = om.SearchForInitialization(fs, fs, [0], [1], prev, win)
            ref = g[f"s{s}_sfi{k}_m"]
            assert nm[0] == int(g[f"s{s}_sfi{k}_n"]) and _rows_equal(rows[0], ref)
            assert np.array_equal(prev[0], g[f"s{s}_sfi{k}_prev"])
        for k, (win, l0, l1, ratio, ori) in enumerate(RG.WIN):
            om = ORBmatcher(ratio, bool(ori), matcher=matcher)
            nm, rows = om.WindowSearch(fs, fs, [0], [1], win, l0, l1)
            assert nm[0] == int(g[f"s{s}_win{k}_n"]) and _rows_equal(rows[0], g[f"s{s}_win{k}_m"])
        for k, (variant, ratio, ori) in enumerate(RG.BOW):
            om = ORBmatcher(ratio, bool(ori), matcher=matcher)
            nm, rows = om.SearchByBoW(fs, fs, [0], [1], variant)
            assert nm[0] == int(g[f"s{s}_bow{k}_n"]) and _rows_equal(rows[0], g[f"s{s}_bow{k}_m"])


def test_cuda_equals_reference_generated_projection_golden(matcher):
    """The CUDA orbm_search_by_projection against vectors produced by the reference's OWN SearchByProjection overloads
    (local map :49-125, last frame :1507-1620, relocalisation :1622-1746, Sim3 :286-407) run on its own Frame / KeyFrame /
    MapPoint objects (tests/golden/make_ref_proj_golden.py), all cases of a scene in ONE batched call per variant."""
    import refgolden as RG
    from orb_slam_b200 import ORBmatcher
    cases = list(RG.proj_cases())
    assert len(cases) == 20
    fs = matcher.upload([c["target"] for c in cases])
    groups = {}
    for i, c in enumerate(cases):
        groups.setdefault((c["variant"], c["th_dist"], c["ratio"], c["ori"]), []).append(i)
    for (variant, th_dist, ratio, ori), idx in groups.items():
        om = ORBmatcher(ratio, ori, matcher=matcher)
        nm, rows = om.SearchByProjection(fs, idx, [cases[i]["points"] for i in idx], variant, th_dist)
        for k, i in enumerate(idx):
            assert nm[k] == cases[i]["ref_n"] and _rows_equal(rows[k], cases[i]["ref_m"]), cases[i]["name"]


def _clustered_frame(rng, n, base_desc, nflip, w=640, h=480):
    """n keypoints packed into a 30 x 30 px patch on one octave with near-identical descriptors: every query wants the
    same few targets, so the accept / taken / steal logic forms dependency chains as long as the frame."""
    from types import SimpleNamespace
    x = rng.uniform(300, 330, n).astype(np.float32); y = rng.uniform(200, 230, n).astype(np.float32)
    desc = np.repeat(base_desc[None], n, 0).copy()
    for i in range(n):
        for b in rng.choice(256, rng.integers(0, nflip + 1), replace=False):
            desc[i, b >> 3] ^= np.uint8(1 << (b & 7))
    ang = rng.uniform(0, 360, n).astype(np.float32)
    return SimpleNamespace(x=x, y=y, angle=ang, octave=np.zeros(n, np.int32), desc=desc, valid=np.zeros(n, np.uint8),
                           min_x=0.0, min_y=0.0, max_x=float(w), max_y=float(h), featvec=None, n=n)


@pytest.mark.parametrize("seed,n", [(0, 300), (1, 700), (2, 64)])
def test_commit_fixed_point_under_long_dependency_chains(M, matcher, seed, n):
    """The parallel (fixed-point) commit against the oracle's sequential loop where it has to work hardest: greedy
    best-only assignment (SearchByProjection 'best', th = 256: every query takes the best FREE target, so query q's
    result depends on all earlier ones), SearchForInitialization with ratio 1 (steals), WindowSearch."""
    from orb_slam_b200 import ORBmatcher
    from orb_slam_b200.matcher import ProjectedPoints
    rng = np.random.default_rng(100 + seed)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    f1, f2 = _clustered_frame(rng, n, base, 24), _clustered_frame(rng, n, base, 24)
    fs = matcher.upload([MD.to_matchframe(f1), MD.to_matchframe(f2)])
    sf = np.ones(8, np.float32)
    # greedy best-only assignment: chains of length ~n
    for ori in (False, True):
        pts = ProjectedPoints(f1.x, f1.y, np.full(n, 60.0, np.float32), np.zeros(n, np.int32), np.zeros(n, np.int32),
                              f1.desc, f1.angle, None)
        om = ORBmatcher(0.9, ori, matcher=matcher)
        nm, rows = om.SearchByProjection(fs, [1], [pts], "best", 256)
        on, orow = M.search_by_projection(f2, pts, "best", 256, 0.9, ori)
        assert nm[0] == on and _rows_equal(rows[0], orow)
        if not ori:
            assert on == n                       # everybody ends up with somebody
    # local-map variant (ratio only when best and second share a level: they always do here)
    pts = ProjectedPoints(f1.x, f1.y, np.full(n, 60.0, np.float32), np.full(n, -1, np.int32), np.zeros(n, np.int32),
                          f1.desc, None, None)
    om = ORBmatcher(0.95, True, matcher=matcher)
    nm, rows = om.SearchByProjection(fs, [1], [pts], "local_map", 100)
    on, orow = M.search_by_projection(f2, pts, "local_map", 100, 0.95, True)
    assert nm[0] == on and _rows_equal(rows[0], orow)
    # SearchForInitialization with steals, WindowSearch with taken targets
    for ratio, ori in ((1.0, False), (0.97, True)):
        om = ORBmatcher(ratio, ori, matcher=matcher)
        prev = [np.stack([f1.x, f1.y], 1).astype(np.float32)]
        oprev = prev[0].copy()
        nm, rows = om.SearchForInitialization(fs, fs, [0], [1], prev, 100)
        on, orow = M.search_for_initialization(f1, f2, oprev, 100, ratio, ori)
        assert nm[0] == on and _rows_equal(rows[0], orow) and np.array_equal(prev[0], oprev)
        g1 = MD.with_valid(f1, np.ones(n, np.uint8))
        fsv = matcher.upload([MD.to_matchframe(g1), MD.to_matchframe(f2)])
        nm, rows = om.WindowSearch(fsv, fsv, [0], [1], 100, 0, 7)
        on, orow = M.window_search(g1, f2, 100, ratio, ori, 0, 7)
        assert nm[0] == on and _rows_equal(rows[0], orow)


@pytest.mark.parametrize("seed,n,ori", [(0, 300, False), (1, 500, True)])
def test_triangulation_fixed_point_under_long_dependency_chains(M, matcher, seed, n, ori):
    """SearchForTriangulation through the fixed-point commit where every query competes for the same targets: all
    keypoints in three vocabulary nodes, near-identical descriptors (every candidate within TH_LOW), an epipolar test
    that passes for roughly half of the pairs."""
    from orb_slam_b200 import ORBmatcher
    rng = np.random.default_rng(200 + seed)
    base = rng.integers(0, 256, 32, dtype=np.uint8)
    f1, f2 = _clustered_frame(rng, n, base, 20), _clustered_frame(rng, n, base, 20)
    f1.featvec = MD.synth_featvec(f1.desc, nnodes=3, seed=1)
    f2.featvec = MD.synth_featvec(f2.desc, nnodes=3, seed=1)
    f1.valid = MD.random_valid(n, 0.1, 5 + seed); f2.valid = MD.random_valid(n, 0.1, 6 + seed)
    fs = matcher.upload([MD.to_matchframe(f1), MD.to_matchframe(f2)])
    sigma2 = np.full(8, 4.0, np.float32)
    t = np.array([7.0, -3.0, 0.0]) + 1e-3
    F12 = np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) * np.float32(0.05)
    om = ORBmatcher(0.6, ori, matcher=matcher)
    nm, rows = om.SearchForTriangulation(fs, fs, [0], [1], F12[None], sigma2)
    on, orow = M.search_for_triangulation(f1, f2, F12, sigma2, ori)
    assert nm[0] == on and _rows_equal(rows[0], orow) and on > 20


def _shuffled_batch(npairs, nprob, seed):
    rng = np.random.default_rng(seed)
    which = np.concatenate([np.arange(npairs), rng.integers(0, npairs, nprob - npairs)])
    rng.shuffle(which)
    return [int(w) for w in which]


def test_large_batches_every_kind(M, matcher, pairs):
    """>= 600 problems per call and kind (the default path for such jobs is the in-order replay commit_kernel; the
    fixture also forces each form): the problems repeat the 6 pairs in shuffled order and every one must equal the
    oracle's result for its source pair."""
    from orb_slam_b200 import ORBmatcher
    NP = 640
    which = _shuffled_batch(len(pairs), NP, 3)
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.6, 110 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.25, 120 + k)) for k, p in enumerate(pairs)]
    s1 = matcher.upload([MD.to_matchframe(f) for f in fr1]); s2 = matcher.upload([MD.to_matchframe(f) for f in fr2])

    def check(nm, rows, ref):
        for p, w in enumerate(which):
            assert nm[p] == ref[w][0] and _rows_equal(rows[p], ref[w][1]), (p, w)

    # WindowSearch
    om = ORBmatcher(0.8, True, matcher=matcher)
    nm, rows = om.WindowSearch(s1, s2, which, which, 100, 0, 2**31 - 1)
    check(nm, rows, [M.window_search(a, b, 100, 0.8, True, 0, 2**31 - 1) for a, b in zip(fr1, fr2)])
    # SearchByProjection(F1, F2, window)
    om = ORBmatcher(0.9, True, matcher=matcher)
    uv = [np.stack([f.x - np.float32(p[2][0]) + 1.5, f.y - np.float32(p[2][1]) - 0.75], 1).astype(np.float32)
          for f, p in zip(fr1, pairs)]
    nm, rows = om.SearchByProjectionFrames(s1, s2, which, which, [uv[w] for w in which], 15)
    check(nm, rows, [M.search_by_projection_ff(a, b, u, 15, 0.9) for a, b, u in zip(fr1, fr2, uv)])
    # SearchByProjection, both variants
    rng = np.random.default_rng(7)
    for variant, ratio, ori, th in (("local_map", 0.8, False, 100), ("best", 0.9, True, 100)):
        pts = [_points_from(p[0], p[2], rng, variant, sf) for p in pairs]
        om = ORBmatcher(ratio, ori, matcher=matcher)
        nm, rows = om.SearchByProjection(s2, which, [pts[w] for w in which], variant, th)
        check(nm, rows, [M.search_by_projection(b, q, variant, th, ratio, ori) for b, q in zip(fr2, pts)])
    # SearchByBoW, both overloads
    for variant in ("kf_f", "kf_kf"):
        om = ORBmatcher(0.75, True, matcher=matcher)
        nm, rows = om.SearchByBoW(s1, s2, which, which, variant)
        check(nm, rows, [M.search_by_bow(a, b, variant, 0.75, True) for a, b in zip(fr1, fr2)])
    # SearchForTriangulation
    sigma2 = (sf * sf).astype(np.float32)
    F12 = []
    for p in pairs:
        t = np.array([p[2][0], p[2][1], 0.0]) + 1e-3
        F12.append(np.array([[0, -t[2], t[1]], [t[2], 0, -t[0]], [-t[1], t[0], 0]], np.float32) * np.float32(0.05))
    g1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.3, 130 + k)) for k, p in enumerate(pairs)]
    g2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.3, 140 + k)) for k, p in enumerate(pairs)]
    t1 = matcher.upload([MD.to_matchframe(f) for f in g1]); t2 = matcher.upload([MD.to_matchframe(f) for f in g2])
    om = ORBmatcher(0.6, True, matcher=matcher)
    nm, rows = om.SearchForTriangulation(t1, t2, which, which, np.stack([F12[w] for w in which]), sigma2)
    check(nm, rows, [M.search_for_triangulation(a, b, f, sigma2, True) for a, b, f in zip(g1, g2, F12)])
    # SearchForInitialization (prev_xy updated in place per problem)
    om = ORBmatcher(0.9, True, matcher=matcher)
    prev = [np.stack([pairs[w][0].x, pairs[w][0].y], 1).astype(np.float32) for w in which]
    nm, rows = om.SearchForInitialization(s1, s2, which, which, prev, 100)
    ref = []
    for a, b in zip(fr1, fr2):
        op = np.stack([a.x, a.y], 1).astype(np.float32)
        n, m = M.search_for_initialization(a, b, op, 100, 0.9, True)
        ref.append((n, m, op))
    for p, w in enumerate(which):
        assert nm[p] == ref[w][0] and _rows_equal(rows[p], ref[w][1]) and np.array_equal(prev[p], ref[w][2]), (p, w)


def test_oversized_frames_take_the_global_state_paths(M, matcher, oracle):
    """Frames of > 2048 keypoints: the replay keeps its per-target state in global memory (kCommitSmemTargets), and the
    fixed point does not fit shared memory above ~4000 queries and falls back to the replay."""
    from orb_slam_b200 import ORBmatcher
    f1, f2, shift = MD.make_pair(oracle, 9, 1241, 376, 4500)
    assert f1.n > 4100 and f2.n > 4100
    a = MD.with_valid(f1, MD.random_valid(f1.n, 0.7, 1)); b = MD.with_valid(f2, MD.random_valid(f2.n, 0.2, 2))
    fs = matcher.upload([MD.to_matchframe(a), MD.to_matchframe(b)])
    om = ORBmatcher(0.85, True, matcher=matcher)
    nm, rows = om.WindowSearch(fs, fs, [0], [1], 60, 0, 2**31 - 1)
    on, orow = M.window_search(a, b, 60, 0.85, True, 0, 2**31 - 1)
    assert nm[0] == on and _rows_equal(rows[0], orow) and on > 300
    prev = [np.stack([a.x, a.y], 1).astype(np.float32)]
    oprev = prev[0].copy()
    nm, rows = om.SearchForInitialization(fs, fs, [0], [1], prev, 50)
    on, orow = M.search_for_initialization(a, b, oprev, 50, 0.85, True)
    assert nm[0] == on and _rows_equal(rows[0], orow) and np.array_equal(prev[0], oprev)
    nm, rows = om.SearchByBoW(fs, fs, [0], [1], "kf_f")
    on, orow = M.search_by_bow(a, b, "kf_f", 0.85, True)
    assert nm[0] == on and _rows_equal(rows[0], orow)


def test_cuda_equals_reference_generated_fuse_golden(matcher):
    """The CUDA matcher against vectors produced by the reference's OWN Fuse (both overloads, src/ORBmatcher.cc:1016-1265),
    SearchBySim3 (:1267-1505) and SearchByProjection(F1, F2, window) (:519-594) (tests/golden/make_ref_fuse_golden.py):
    ORBM_PROJ_BEST_INDEPENDENT, orbm_search_by_sim3, orbm_search_by_projection_frames."""
    import refgolden as RG
    import refscene as RS
    from orb_slam_b200 import ORBmatcher
    om = ORBmatcher(0.9, True, matcher=matcher)
    cases = list(RG.fuse_cases())
    fus = [c for c in cases if c["kind"] in ("fuse", "fuse_sim3")]
    fs = matcher.upload([c["target"] for c in fus])
    nm, rows = om.Fuse(fs, list(range(len(fus))), [c["points"] for c in fus], 50)       # all 8 calls in one batch
    for k, c in enumerate(fus):
        assert nm[k] == c["ref_n"], c["name"]
        best = rows[k, :c["points"].n]
        assert np.all(rows[k, c["points"].n:] == -1)
        for key, val in RS.fuse_expectation(c, best).items():
            assert np.array_equal(val, c[key]), (c["name"], key)
    sim = [c for c in cases if c["kind"] == "sim3"]
    fs = matcher.upload([f for c in sim for f in (c["f1"], c["f2"])])
    i1 = list(range(0, 2 * len(sim), 2)); i2 = [i + 1 for i in i1]
    nf, rows = om.SearchBySim3(fs, i1, i2, [c["p12"] for c in sim], [c["p21"] for c in sim], 100)
    for k, c in enumerate(sim):
        m12 = rows[k, :c["f1"].n]
        assert nf[k] == c["ref_n"] and np.array_equal(np.where(c["pre"] >= 0, c["pre"], m12), c["ref_m"]), c["name"]
    for c in (c for c in cases if c["kind"] == "proj_ff"):
        fs = matcher.upload([c["f1"], c["f2"]])
        o2 = ORBmatcher(c["ratio"], True, matcher=matcher)
        nm, rows = o2.SearchByProjectionFrames(fs, fs, [0], [1], [c["uv"]], c["window"])
        assert nm[0] == c["ref_n"] and _rows_equal(rows[0], c["ref_m"]), c["name"]


def test_fuse_and_sim3_scoring_large_batches(M, matcher, pairs):
    """ORBM_PROJ_BEST_INDEPENDENT / orbm_search_by_sim3 against the oracle on 1000-keypoint frames, 6 problems (fixed
    point commit) and 640 problems (replay commit)."""
    from orb_slam_b200 import ORBmatcher, ProjectedPoints
    sf = np.cumprod(np.concatenate([[1.0], np.full(7, np.float32(1.2))]).astype(np.float32)).astype(np.float32)
    rng = np.random.default_rng(77)
    fr1 = [MD.with_valid(p[0], MD.random_valid(p[0].n, 0.5, 210 + k)) for k, p in enumerate(pairs)]
    fr2 = [MD.with_valid(p[1], MD.random_valid(p[1].n, 0.5, 220 + k)) for k, p in enumerate(pairs)]
    fs = matcher.upload([MD.to_matchframe(f) for f in fr1 + fr2])
    P = len(pairs)

    def proj(src, shift, sign):
        n = src.n
        u = (src.x - sign * np.float32(shift[0]) + rng.uniform(-2, 2, n)).astype(np.float32)
        v = (src.y - sign * np.float32(shift[1]) + rng.uniform(-2, 2, n)).astype(np.float32)
        level = np.clip(src.octave + rng.integers(0, 2, n), 0, 7)
        return ProjectedPoints.for_fuse(u, v, level, src.desc, sf, float(rng.choice([2.5, 4.0, 7.5])),
                                        valid=(rng.random(n) < 0.85).astype(np.uint8))
    p12 = [proj(fr1[k], pairs[k][2], 1) for k in range(P)]
    p21 = [proj(fr2[k], pairs[k][2], -1) for k in range(P)]
    om = ORBmatcher(0.9, True, matcher=matcher)
    ref_f = [M.search_by_projection_independent(fr2[k], p12[k], 50) for k in range(P)]
    ref_s = [M.search_by_sim3(fr1[k], fr2[k], p12[k], p21[k], 100) for k in range(P)]
    assert sum(r[0] for r in ref_f) > 600 and sum(r[0] for r in ref_s) > 300
    for which in (list(range(P)), _shuffled_batch(P, 640, 5)):
        nm, rows = om.Fuse(fs, [P + w for w in which], [p12[w] for w in which], 50)
        for p, w in enumerate(which):
            assert nm[p] == ref_f[w][0] and _rows_equal(rows[p], ref_f[w][1]), (p, w)
        nf, rows = om.SearchBySim3(fs, which, [P + w for w in which], [p12[w] for w in which], [p21[w] for w in which], 100)
        for p, w in enumerate(which):
            assert nf[p] == ref_s[w][0] and _rows_equal(rows[p], ref_s[w][1]), (p, w)
