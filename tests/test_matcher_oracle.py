"""CPU tests of the matcher oracle (oracle/match_oracle.cc): known-answer tests derived from the reference source and
agreement with the independent Python restatement tests/pyref_matcher.py (the reference ships no golden
vectors for this path)."""
from types import SimpleNamespace

import numpy as np
import pytest

import matchdata as MD
import pyref_matcher as PY


@pytest.fixture(scope="module")
def M(oracle):
    from oracle import match_oracle
    match_oracle.lib()
    return match_oracle


def test_descriptor_distance_kat(M):
    rng = np.random.default_rng(0)
    z, o = np.zeros(32, np.uint8), np.full(32, 255, np.uint8)
    assert M.descriptor_distance(z, z) == 0 and M.descriptor_distance(z, o) == 256
    for _ in range(200):
        a, b = rng.integers(0, 256, 32, dtype=np.uint8), rng.integers(0, 256, 32, dtype=np.uint8)
        assert M.descriptor_distance(a, b) == sum(bin(int(x) ^ int(y)).count("1") for x, y in zip(a, b))


def test_three_maxima_kat(M):
    h = [0] * 30
    assert M.three_maxima(h) == (-1, -1, -1)
    h[3], h[7], h[1] = 50, 40, 4        # third < 10% of max -> dropped
    assert M.three_maxima(h) == (3, 7, -1)
    h[1] = 5                            # exactly 10%: 5 < 0.1f*50 is false -> kept
    assert M.three_maxima(h) == (3, 7, 1)
    h = [0] * 30; h[2] = h[9] = h[20] = 10   # ties: strict > keeps scan order
    assert M.three_maxima(h) == (2, 9, 20)
    h = [0] * 30; h[5] = 100; h[6] = 9
    assert M.three_maxima(h) == (5, -1, -1)
    rng = np.random.default_rng(1)
    for _ in range(200):
        s = rng.integers(0, 40, 30)
        assert M.three_maxima(s) == PY.three_maxima(s.tolist())


def _mini(xs, ys, desc, octave=None, angle=None, valid=None):
    n = len(xs)
    return SimpleNamespace(x=np.asarray(xs, np.float32), y=np.asarray(ys, np.float32),
                           angle=np.zeros(n, np.float32) if angle is None else np.asarray(angle, np.float32),
                           octave=np.zeros(n, np.int32) if octave is None else np.asarray(octave, np.int32),
                           desc=np.asarray(desc, np.uint8).reshape(n, 32), valid=valid, min_x=0.0, min_y=0.0,
                           max_x=640.0, max_y=480.0, featvec=None, n=n)


def _d(nbits):
    d = np.zeros(32, np.uint8)
    for b in range(nbits):
        d[b // 8] |= 1 << (b % 8)
    return d


def test_init_semantics_handmade(M):
    """Tie -> first visited candidate wins; steal only when strictly better; TH_LOW boundary (<= 50)."""
    z = _d(0)
    # two identical candidates at distance 10 -> best == second -> ratio test fails (10 < 0.9*10 false)
    f1 = _mini([100], [100], [z])
    f2 = _mini([101, 102], [100, 100], [_d(10), _d(10)])
    prev = np.array([[100, 100]], np.float32)
    n, m = M.search_for_initialization(f1, f2, prev, 100, 0.9, False)
    assert n == 0 and m[0] == -1
    # distance exactly 50 accepted, 51 rejected
    for nb, ok in ((50, True), (51, False)):
        f2 = _mini([101], [100], [_d(nb)])
        prev = np.array([[100, 100]], np.float32)
        n, m = M.search_for_initialization(f1, f2, prev, 100, 0.9, False)
        assert (n == 1) == ok
    # steal: query 1 is strictly closer to the target than query 0 -> takes it over, query 0 loses its match
    f1 = _mini([100, 100], [100, 100], [_d(8), _d(2)])
    f2 = _mini([101], [100], [z])
    prev = np.array([[100, 100], [100, 100]], np.float32)
    n, m = M.search_for_initialization(f1, f2, prev, 100, 0.9, False)
    assert n == 1 and list(m) == [-1, 0]
    # equal distance: `vMatchedDistance[i2] <= dist` skips the candidate -> the first owner keeps it
    f1 = _mini([100, 100], [100, 100], [_d(8), _d(8)])
    prev = np.array([[100, 100], [100, 100]], np.float32)
    n, m = M.search_for_initialization(f1, f2, prev, 100, 0.9, False)
    assert n == 1 and list(m) == [0, -1]
    # level > 0 keypoints of F1 are never queried; level > 0 keypoints of F2 are never candidates
    f1 = _mini([100], [100], [z], octave=[1])
    n, m = M.search_for_initialization(f1, f2, np.array([[100, 100]], np.float32), 100, 0.9, False)
    assert n == 0


def test_grid_quirks(M):
    """round() binning vs floor/ceil queries; keypoints whose rounded cell is 64 / 48 are dropped (Appendix A.14)."""
    f = _mini([639.9, 5.0, 14.9, 15.1], [10, 479.9, 10, 10], [_d(0)] * 4)
    cs, items = M.grid_csr(f)
    assert 0 not in items and 1 not in items           # cell index 64 / 48 -> dropped
    assert list(items) == [2, 3] or list(items) == [3, 2]
    # 14.9 -> round(1.49) = 1, 15.1 -> round(1.51) = 2: different cells, ix-major order
    assert list(items) == [2, 3]
    got = M.features_in_area(f, 15.0, 10.0, 1.0)
    assert list(got) == [2, 3]
    assert list(M.features_in_area(f, 700.0, 10.0, 5.0)) == []


def test_oracle_vs_python_restatement(oracle, M):
    for seed in (0, 1):
        f1, f2, _ = MD.make_pair(oracle, seed, 320, 240, 500)
        g2 = PY.build_grid(f2)
        cs, items = M.grid_csr(f2)
        flat = [k for ix in range(64) for iy in range(48) for k in g2[ix][iy]]
        assert flat == list(items)
        for ori in (False, True):
            p1 = np.stack([f1.x, f1.y], 1).astype(np.float32).copy()
            p2 = p1.copy()
            n1, m1 = M.search_for_initialization(f1, f2, p1, 100, 0.9, ori)
            n2, m2 = PY.search_for_initialization(f1, f2, p2, 100, 0.9, ori)
            assert n1 == n2 and np.array_equal(m1, m2) and np.array_equal(p1, p2)
            assert n1 > 20
        kf = MD.with_valid(f1, MD.random_valid(f1.n, 0.7, seed))
        for ori in (False, True):
            n1, m1 = M.search_by_bow(kf, f2, "kf_f", 0.75, ori)
            n2, m2 = PY.search_by_bow_kf_f(kf, f2, 0.75, ori)
            assert n1 == n2 and np.array_equal(m1, m2)


def test_features_in_area_vs_python(oracle, M):
    f1, f2, _ = MD.make_pair(oracle, 0, 320, 240, 500)
    g2 = PY.build_grid(f2)
    rng = np.random.default_rng(7)
    for _ in range(200):
        x, y, r = float(rng.uniform(-20, 340)), float(rng.uniform(-20, 260)), float(rng.uniform(1, 80))
        mode = int(rng.integers(0, 3))
        minl, maxl = (-1, -1) if mode == 0 else ((2, 2) if mode == 1 else (int(rng.integers(-1, 3)), int(rng.integers(3, 6))))
        assert list(M.features_in_area(f2, x, y, r, minl, maxl)) == PY.features_in_area(f2, g2, x, y, r, minl, maxl)


def test_knn2_oracle(M):
    rng = np.random.default_rng(3)
    db = rng.integers(0, 256, (300, 32), dtype=np.uint8)
    q = db[rng.integers(0, 300, 40)].copy()
    q[:, :2] ^= rng.integers(0, 256, (40, 2), dtype=np.uint8)
    db[17] = db[5]  # duplicate descriptor: the lower index must win, the duplicate becomes the second best
    q[0] = db[5]
    got = M.knn2(q, db)
    assert np.array_equal(got, PY.knn2(q, db))
    assert got[0, 0] == 5 and got[0, 1] == 17
    assert np.array_equal(M.knn2(q, db, threads=3), got)
    assert np.array_equal(M.knn2(q, db[:1])[:, 1], np.full(40, 0xFFFFFFFF, np.uint32))


# ---- N4: MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) ------------------------------------------
def _np_distinctive(desc, start):
    """Independent numpy restatement: full distance matrix, np.sort per row, first minimum of the medians."""
    out = []
    for p in range(len(start) - 1):
        d = desc[start[p]:start[p + 1]]
        n = len(d)
        if n == 0:
            out.append(-1)
            continue
        D = np.unpackbits(d[:, None, :] ^ d[None, :, :], axis=2).sum(2)
        med = np.sort(D, axis=1)[:, int(0.5 * (n - 1))]
        out.append(int(np.argmin(med)))
    return np.asarray(out, np.int32)


def test_distinctive_descriptors_hand_cases():
    from oracle import match_oracle as M
    z = np.zeros((1, 32), np.uint8)
    a = z.copy(); a[0, 0] = 0x0F           # 4 bits from z
    b = z.copy(); b[0, :2] = 0xFF          # 16 bits from z
    # one observation: itself
    assert list(M.distinctive_descriptors(z, [0, 1])) == [0]
    # two observations: medians are row[(int)(0.5*1)] = row[0] = 0 for both -> first wins
    assert list(M.distinctive_descriptors(np.vstack([a, b]), [0, 2])) == [0]
    # three: rows sorted {0,4,16}, {0,4,12}, {0,12,16}; medians 4, 4, 12 -> index 0 (first of the tie)
    assert list(M.distinctive_descriptors(np.vstack([z, a, b]), [0, 3])) == [0]
    # three with the central one last: {0,12,16}, {0,4,16}... order (b, z, a): medians 12, 4, 4 -> index 1
    assert list(M.distinctive_descriptors(np.vstack([b, z, a]), [0, 3])) == [1]
    # empty point in a batch
    assert list(M.distinctive_descriptors(np.vstack([z, a, b]), [0, 0, 3, 3])) == [-1, 0, -1]


def test_distinctive_descriptors_random_vs_numpy():
    from oracle import match_oracle as M
    rng = np.random.default_rng(11)
    sizes = [1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100, 0, 7]
    start = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = rng.integers(0, 256, (len(sizes), 32), dtype=np.uint8)
    desc = np.repeat(base, sizes, axis=0)
    flips = rng.random((len(desc), 256)) < rng.uniform(0.02, 0.2, (len(desc), 1))
    desc = np.packbits(np.unpackbits(desc, axis=1) ^ flips, axis=1)
    assert np.array_equal(M.distinctive_descriptors(desc, start), _np_distinctive(desc, start))
