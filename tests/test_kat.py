"""Known-answer tests derived from the reference source: tables of the constructor, grid plan,
pattern checksum, fastAtan2 spot values."""
import zlib

import numpy as np


def test_pattern_crc(oracle):
    p = oracle.pattern()
    assert p.min() == -13 and p.max() == 12
    assert zlib.crc32(p.astype(np.int8).tobytes()) == 0xD1A39030
    # first two pairs of reference src/ORBextractor.cc:199-200
    assert list(p[:8]) == [8, -3, 9, 5, 4, 2, 7, -12]


def test_ctor_tables(oracle):
    f, s, i, u = oracle.OracleExtractor(1000, 1.2, 8, 1, 20).tables()
    assert list(f) == [217, 181, 151, 126, 105, 87, 73, 60]            # src/ORBextractor.cc:476-487
    assert list(u) == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]  # :495-510 (749-px disc)
    assert sum(2 * int(x) + 1 for x in u[1:]) * 2 + 2 * int(u[0]) + 1 == 749
    assert s[1] == np.float32(1.2) and abs(float(s[7]) - 3.5831816) < 1e-6
    f2 = oracle.OracleExtractor(2000, 1.2, 8, 1, 20).tables()[0]
    assert list(f2) == [434, 362, 302, 251, 209, 175, 145, 122]


PLAN_640 = [(640, 480, 5, 6, 122, 75, 8), (533, 400, 5, 6, 101, 62, 7), (444, 333, 4, 5, 103, 61, 8),
            (370, 278, 4, 5, 85, 50, 7), (309, 231, 3, 4, 93, 50, 9), (257, 193, 3, 4, 75, 41, 8),
            (214, 161, 3, 4, 61, 33, 7), (179, 134, 3, 4, 49, 26, 5)]
PLAN_KITTI = [(1241, 376, 5, 16, 242, 22, 6), (1034, 313, 4, 13, 251, 22, 7), (862, 261, 4, 13, 208, 18, 6),
              (718, 218, 3, 9, 229, 21, 10), (598, 181, 3, 9, 189, 17, 8), (499, 151, 3, 9, 156, 14, 7),
              (416, 126, 2, 6, 192, 16, 13), (346, 105, 2, 6, 157, 13, 11)]


def _plan(oracle, w, h, nf):
    E = oracle.OracleExtractor(nf, 1.2, 8, 1, 20)
    E(oracle.synth(h, w, 1))
    out = []
    for l in range(8):
        g = E.level_grid(l)
        out.append((*E.level_size(l), g["cols"], g["rows"], g["cellW"], g["cellH"], g["quota"]))
    return out


def test_grid_plan(oracle):
    """SURVEY §0.3 P6 table (reference :462-487,527-547,785-786), incl. the wide-image rows>cols quirk."""
    assert _plan(oracle, 640, 480, 1000) == PLAN_640
    assert _plan(oracle, 1241, 376, 2000) == PLAN_KITTI


def test_fast_atan2_spot(oracle):
    assert oracle.fast_atan2(0.0, 0.0) == 0.0
    assert oracle.fast_atan2(0.0, -5.0) == 180.0
    assert abs(oracle.fast_atan2(-1.0, 5.0) - 348.6897888) < 1e-4
    assert abs(oracle.fast_atan2(1.0, 1.0) - 45.0) < 0.01


def test_synth_saturates_quotas(oracle):
    """The synthetic frame must exercise quota redistribution: every level fills its quota."""
    E = oracle.OracleExtractor(1000, 1.2, 8, 1, 20)
    k, d = E(oracle.synth(480, 640, 0))
    assert len(k) == 1000 and d.shape == (1000, 32)
    assert [int((k["octave"] == l).sum()) for l in range(8)] == [217, 181, 151, 126, 105, 87, 73, 60]
    assert np.all(k["class_id"] == -1) and np.all((k["angle"] >= 0) & (k["angle"] < 360))
    assert list(np.unique(k["size"])) == [31.0, 37.0, 44.0, 53.0, 64.0, 77.0, 92.0, 111.0]
