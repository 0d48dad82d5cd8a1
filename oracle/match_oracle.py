"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of the matcher restatement (oracle/match_oracle.cc, built into oracle/liborb_oracle.so).  Frames are
duck-typed: any object with x, y, angle, octave, desc, valid, min_x, min_y, max_x, max_y (and featvec for the BoW
routines).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle as _O

GRID_COLS, GRID_ROWS = 64, 48


class _Frame(C.Structure):
    _fields_ = [("n", C.c_int32), ("x", C.c_void_p), ("y", C.c_void_p), ("angle", C.c_void_p), ("octave", C.c_void_p),
                ("desc", C.c_void_p), ("valid", C.c_void_p), ("min_x", C.c_float), ("min_y", C.c_float),
                ("inv_w", C.c_float), ("inv_h", C.c_float)]


class _FeatVec(C.Structure):
    _fields_ = [("nnodes", C.c_int32), ("node_id", C.c_void_p), ("start", C.c_void_p), ("items", C.c_void_p)]


_bound = {}


def lib(fast=False):
    L = _O.lib(fast)
    if fast in _bound:
        return L
    vp, f32 = C.c_void_p, C.c_float
    FP, VP = C.POINTER(_Frame), C.POINTER(_FeatVec)
    L.mo_descriptor_distance.argtypes = [vp, vp]
    L.mo_grid_csr.argtypes = [FP, vp, vp]
    L.mo_features_in_area.argtypes = [FP, f32, f32, f32, C.c_int, C.c_int, vp]
    L.mo_three_maxima.argtypes = [vp, C.c_int, vp]
    L.mo_search_for_initialization.argtypes = [FP, FP, vp, C.c_int, f32, C.c_int, vp]
    L.mo_window_search.argtypes = [FP, FP, C.c_int, f32, C.c_int, C.c_int, C.c_int, vp]
    L.mo_search_by_projection_ff.argtypes = [FP, FP, vp, vp, C.c_int, f32, vp]
    L.mo_search_by_projection.argtypes = [FP, C.c_int, vp, vp, vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, f32, C.c_int, vp]
    L.mo_search_by_projection_independent.argtypes = [FP, C.c_int, vp, vp, vp, vp, vp, vp, vp, C.c_int, vp]
    L.mo_search_by_bow_kf_f.argtypes = [FP, FP, VP, VP, f32, C.c_int, vp]
    L.mo_search_by_bow_kf_kf.argtypes = [FP, FP, VP, VP, f32, C.c_int, vp]
    L.mo_search_for_triangulation.argtypes = [FP, FP, VP, VP, vp, vp, C.c_int, vp]
    L.mo_knn2.argtypes = [vp, C.c_int, vp, C.c_int64, C.c_int64, vp]
    if hasattr(L, "mo_knn2_mt"):
        L.mo_knn2_mt.argtypes = [vp, C.c_int, vp, C.c_int64, C.c_int64, vp, C.c_int]
    if hasattr(L, "mo_search_for_initialization_many"):
        L.mo_search_for_initialization_many.argtypes = [FP, FP, C.c_int, vp, C.c_int, C.c_int, f32, C.c_int, vp, vp,
                                                        C.c_int]
    _bound[fast] = True
    return L


class _Held:
    """A C struct plus the numpy arrays it points into."""

    def __init__(self, s, keep):
        self.s, self.keep = s, keep


def frame_struct(f) -> _Held:
    x = np.ascontiguousarray(f.x, np.float32); y = np.ascontiguousarray(f.y, np.float32)
    a = np.ascontiguousarray(f.angle, np.float32); o = np.ascontiguousarray(f.octave, np.int32)
    d = np.ascontiguousarray(f.desc, np.uint8)
    v = None if getattr(f, "valid", None) is None else np.ascontiguousarray(f.valid, np.uint8)
    iw = np.float32(GRID_COLS) / np.float32(np.float32(f.max_x) - np.float32(f.min_x))
    ih = np.float32(GRID_ROWS) / np.float32(np.float32(f.max_y) - np.float32(f.min_y))
    s = _Frame(len(x), x.ctypes.data, y.ctypes.data, a.ctypes.data, o.ctypes.data, d.ctypes.data,
               None if v is None else v.ctypes.data, f.min_x, f.min_y, float(iw), float(ih))
    return _Held(s, (x, y, a, o, d, v))


def featvec_struct(f) -> _Held:
    nid, st, it = (np.ascontiguousarray(a, np.int32) for a in f.featvec)
    return _Held(_FeatVec(len(nid), nid.ctypes.data, st.ctypes.data, it.ctypes.data), (nid, st, it))


def descriptor_distance(a, b) -> int:
    a = np.ascontiguousarray(a, np.uint8); b = np.ascontiguousarray(b, np.uint8)
    return int(lib().mo_descriptor_distance(a.ctypes.data, b.ctypes.data))


def distinctive_descriptors(desc, start):
    """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) for a batch: best index per point."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); start = np.ascontiguousarray(start, np.int32)
    best = np.zeros(len(start) - 1, np.int32)
    L = lib()
    L.mo_distinctive_descriptors.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    L.mo_distinctive_descriptors(desc.ctypes.data, start.ctypes.data, len(best), best.ctypes.data)
    return best


def grid_csr(f):
    h = frame_struct(f)
    cs = np.zeros(GRID_COLS * GRID_ROWS + 1, np.int32); items = np.zeros(max(h.s.n, 1), np.int32)
    lib().mo_grid_csr(C.byref(h.s), cs.ctypes.data, items.ctypes.data)
    return cs, items[: cs[-1]]


def features_in_area(f, x, y, r, min_level=-1, max_level=-1):
    h = frame_struct(f)
    out = np.zeros(max(h.s.n, 1), np.int32)
    n = lib().mo_features_in_area(C.byref(h.s), x, y, r, min_level, max_level, out.ctypes.data)
    return out[:n].copy()


def three_maxima(sizes):
    s = np.ascontiguousarray(sizes, np.int32); out = np.zeros(3, np.int32)
    lib().mo_three_maxima(s.ctypes.data, len(s), out.ctypes.data)
    return tuple(int(v) for v in out)


def search_for_initialization(f1, f2, prev_xy, window, ratio, check_ori):
    """prev_xy [n1,2] float32 is updated in place; returns (nmatches, vnMatches12)."""
    h1, h2 = frame_struct(f1), frame_struct(f2)
    assert prev_xy.dtype == np.float32 and prev_xy.flags.c_contiguous
    m = np.empty(max(h1.s.n, 1), np.int32)
    n = lib().mo_search_for_initialization(C.byref(h1.s), C.byref(h2.s), prev_xy.ctypes.data, window, ratio,
                                           int(check_ori), m.ctypes.data)
    return n, m[: h1.s.n]


def window_search(f1, f2, window, ratio, check_ori, min_level=0, max_level=2**31 - 1):
    h1, h2 = frame_struct(f1), frame_struct(f2)
    m = np.empty(max(h2.s.n, 1), np.int32)
    n = lib().mo_window_search(C.byref(h1.s), C.byref(h2.s), window, ratio, int(check_ori), min_level, max_level,
                               m.ctypes.data)
    return n, m[: h2.s.n]


def search_by_projection_ff(f1, f2, uv, window, ratio):
    h1, h2 = frame_struct(f1), frame_struct(f2)
    u = np.ascontiguousarray(uv[:, 0], np.float32); v = np.ascontiguousarray(uv[:, 1], np.float32)
    m = np.empty(max(h2.s.n, 1), np.int32)
    n = lib().mo_search_by_projection_ff(C.byref(h1.s), C.byref(h2.s), u.ctypes.data, v.ctypes.data, window, ratio,
                                         m.ctypes.data)
    return n, m[: h2.s.n]


def search_by_projection(f, pts, variant, th_dist, ratio, check_ori):
    """pts: object with u, v, radius, min_level, max_level, desc, angle, valid (ProjectedPoints-like)."""
    h = frame_struct(f)
    g = lambda a, t: None if a is None else np.ascontiguousarray(a, t)
    u, v, r = g(pts.u, np.float32), g(pts.v, np.float32), g(pts.radius, np.float32)
    mn, mx = g(pts.min_level, np.int32), g(pts.max_level, np.int32)
    d, a, val = g(pts.desc, np.uint8), g(pts.angle, np.float32), g(pts.valid, np.uint8)
    m = np.empty(max(h.s.n, 1), np.int32)
    p = lambda z: None if z is None else z.ctypes.data
    n = lib().mo_search_by_projection(C.byref(h.s), len(u), p(u), p(v), p(r), p(mn), p(mx), p(d), p(a), p(val),
                                      0 if variant == "local_map" else 1, th_dist, ratio, int(check_ori), m.ctypes.data)
    return n, m[: h.s.n]


def search_by_projection_independent(f, pts, th_dist):
    """Fuse / SearchBySim3 best-only loop: per POINT the best keypoint of f within th_dist (-1 none) -> (n, best)."""
    h = frame_struct(f)
    g = lambda a, t: None if a is None else np.ascontiguousarray(a, t)
    u, v, r = g(pts.u, np.float32), g(pts.v, np.float32), g(pts.radius, np.float32)
    mn, mx = g(pts.min_level, np.int32), g(pts.max_level, np.int32)
    d, val = g(pts.desc, np.uint8), g(pts.valid, np.uint8)
    m = np.empty(max(len(u), 1), np.int32)
    p = lambda z: None if z is None else z.ctypes.data
    n = lib().mo_search_by_projection_independent(C.byref(h.s), len(u), p(u), p(v), p(r), p(mn), p(mx), p(d), p(val),
                                                  th_dist, m.ctypes.data)
    return n, m[: len(u)]


def search_by_sim3(f1, f2, pts12, pts21, th_dist):
    """SearchBySim3's two scans + agreement (src/ORBmatcher.cc:1487-1502) -> (nFound, match12[i1] = i2 or -1)."""
    _, m1 = search_by_projection_independent(f2, pts12, th_dist)
    _, m2 = search_by_projection_independent(f1, pts21, th_dist)
    out = np.full(len(m1), -1, np.int32)
    for i1, i2 in enumerate(m1):
        if i2 >= 0 and m2[i2] == i1:
            out[i1] = i2
    return int((out >= 0).sum()), out


def search_by_bow(f1, f2, variant, ratio, check_ori):
    h1, h2, v1, v2 = frame_struct(f1), frame_struct(f2), featvec_struct(f1), featvec_struct(f2)
    nout = h2.s.n if variant == "kf_f" else h1.s.n
    m = np.empty(max(nout, 1), np.int32)
    fn = lib().mo_search_by_bow_kf_f if variant == "kf_f" else lib().mo_search_by_bow_kf_kf
    n = fn(C.byref(h1.s), C.byref(h2.s), C.byref(v1.s), C.byref(v2.s), ratio, int(check_ori), m.ctypes.data)
    return n, m[:nout]


def search_for_triangulation(f1, f2, F12, sigma2, check_ori):
    h1, h2, v1, v2 = frame_struct(f1), frame_struct(f2), featvec_struct(f1), featvec_struct(f2)
    F = np.ascontiguousarray(F12, np.float32).reshape(9); s2 = np.ascontiguousarray(sigma2, np.float32)
    m = np.empty(max(h1.s.n, 1), np.int32)
    n = lib().mo_search_for_triangulation(C.byref(h1.s), C.byref(h2.s), C.byref(v1.s), C.byref(v2.s), F.ctypes.data,
                                          s2.ctypes.data, int(check_ori), m.ctypes.data)
    return n, m[: h1.s.n]


def knn2(q, db, idx_base=0, fast=False, threads=1):
    q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32); db = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
    out = np.empty((len(q), 2), np.uint32)
    L = lib(fast)
    if threads > 1:
        L.mo_knn2_mt(q.ctypes.data, len(q), db.ctypes.data, len(db), idx_base, out.ctypes.data, threads)
    else:
        L.mo_knn2(q.ctypes.data, len(q), db.ctypes.data, len(db), idx_base, out.ctypes.data)
    return out
