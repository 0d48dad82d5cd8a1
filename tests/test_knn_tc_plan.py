"""Host logic of the tensor-core 2-NN (knn2_tc_kernel, csrc/orbm_knn_tc.inl), without a GPU: the work split computed by
knn2_launch (through orbm_debug_knn2_tc_plan) is walked exactly as the kernel walks it - cluster c takes the tile steps
[c * per, (c + 1) * per) of the pair-major linearisation, one segment per query-tile pair it touches, and writes its
result for that pair to row `c - first cluster of the pair` of the partial-result buffer - and must cover every
(pair, tile) once, never write the same (row, pair) twice and stay inside the buffer the host allocated."""
import ctypes

import numpy as np


def _plan(lib, nq, ndb, sms):
    out = (ctypes.c_int64 * 5)()
    rc = lib.orbm_debug_knn2_tc_plan(int(nq), ctypes.c_int64(int(ndb)), int(sms), out)
    assert rc == 0
    return [int(v) for v in out]


def _walk(npairs, ntiles, clusters, per, pieces):
    total = npairs * ntiles
    seen = np.zeros(total, dtype=np.int32)
    rows = set()
    for c in range(clusters):
        lin, lin1 = c * per, min(total, (c + 1) * per)
        assert lin < lin1                                   # no idle cluster is launched
        while lin < lin1:
            pair = lin // ntiles
            t0 = lin - pair * ntiles
            nseg = min(ntiles - t0, lin1 - lin)
            piece = c - (pair * ntiles) // per
            assert 0 <= piece < pieces
            assert (piece, pair) not in rows                # one writer per row of a pair
            rows.add((piece, pair))
            seen[lin:lin + nseg] += 1
            lin += nseg
    assert seen.min() == 1 and seen.max() == 1              # every tile step exactly once


def test_tensor_core_work_split_covers_every_tile_once():
    from orb_slam_b200 import _lib
    lib = _lib.lib()
    lib.orbm_debug_knn2_tc_plan.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
    lib.orbm_debug_knn2_tc_plan.restype = ctypes.c_int
    rng = np.random.default_rng(5)
    shapes = [(1, 1), (128, 256), (129, 257), (256, 255), (257, 1), (8000, 5_000_000), (100_000, 300), (3, 8_388_608),
              (200_000, 40_000), (5000, 1000)]
    shapes += [(int(np.exp(rng.uniform(0, np.log(300_000)))), int(np.exp(rng.uniform(0, np.log(8_388_608))))) for _ in range(300)]
    for nq, ndb in shapes:
        for sms in (148, 132, 2, 7):
            npairs, ntiles, clusters, per, pieces = _plan(lib, nq, ndb, sms)
            assert npairs == ((nq + 127) // 128 + 1) // 2 and ntiles == (ndb + 255) // 256
            assert 1 <= clusters <= max(1, sms // 2) and clusters * per >= npairs * ntiles
            if npairs * ntiles <= 2_000_000:               # the walk itself is O(tile steps)
                _walk(npairs, ntiles, clusters, per, pieces)
            else:                                           # closed form of the same bound
                assert pieces >= -(-ntiles // per) + 1


def test_tensor_core_plan_rejects_nonsense():
    from orb_slam_b200 import _lib
    lib = _lib.lib()
    lib.orbm_debug_knn2_tc_plan.argtypes = [ctypes.c_int, ctypes.c_int64, ctypes.c_int, ctypes.POINTER(ctypes.c_int64)]
    out = (ctypes.c_int64 * 5)()
    assert lib.orbm_debug_knn2_tc_plan(0, 10, 148, out) != 0
    assert lib.orbm_debug_knn2_tc_plan(10, 0, 148, out) != 0
    assert lib.orbm_debug_knn2_tc_plan(10, 10, 1, out) != 0
