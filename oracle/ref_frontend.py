"""ORACLE — TEST INFRASTRUCTURE ONLY.

The reference's OWN front-end objects: Frame / KeyFrame / MapPoint / ORBmatcher / ORBVocabulary compiled from
/root/reference (oracle/ref_orbmatcher.cc, oracle/_ref/libref_orbslam.so), OpenCV routines served by cv2
(oracle/ref_extractor.py).  Frames are built by Frame's own constructor from an IMAGE: extraction, undistortion, image
bounds and grid binning are all the reference's code.  Authoring container only; tests skip without it.
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os

import numpy as np

from . import ref_extractor as RE

LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_orbslam.so")
_P, _I, _F = C.c_void_p, C.c_int, C.c_float


def available() -> bool:
    if not os.path.exists(LIB):
        return False
    try:
        import cv2  # noqa: F401
    except Exception:
        return False
    return True


DROPIN_LIB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref", "libref_dropin.so")


@contextlib.contextmanager
def library(path):
    """Run the same wrappers against another build of them: DROPIN_LIB = the reference's Frame / KeyFrame / MapPoint
    sources compiled against include/shim/ORBextractor.h + ORBmatcher.h and liborbx.so (oracle/Makefile ref_dropin) -
    the CUDA path behind the reference's own classes.  Objects remember the library they were created with."""
    global LIB
    old, LIB = LIB, path
    try:
        yield
    finally:
        LIB = old


def _lib():
    L = RE._lib(LIB)
    if not getattr(L, "_m_ready", False):
        sig = [
            ("ref_m_world_create", "restype", _P),
            ("ref_m_world_create", "argtypes", [C.c_char_p]),
            ("ref_m_frame_create", "restype", _P),
            ("ref_m_frame_create", "argtypes", [_P, _P, _P, _I, _I, _P, _P, _I]),
            ("ref_m_frame_n", "argtypes", [_P]),
            ("ref_m_frame_bounds", "argtypes", [_P, _P, _P]),
            ("ref_m_frame_read", "argtypes", [_P, _P, _P, _P, _P]),
            ("ref_m_frame_grid", "argtypes", [_P, _P, _P]),
            ("ref_m_features_in_area", "argtypes", [_P, _F, _F, _F, _I, _I, _P]),
            ("ref_m_search_for_initialization", "argtypes", [_P, _P, _P, _P, _I, _F, _I]),
            ("ref_m_frame_set_valid", "argtypes", [_P, _P]),
            ("ref_m_window_search", "argtypes", [_P, _P, _I, _I, _I, _F, _I, _P]),
            ("ref_m_frame_compute_bow", "argtypes", [_P]),
            ("ref_m_frame_bow", "argtypes", [_P] + [_P] * 7),
            ("ref_m_keyframe_create", "restype", _P),
            ("ref_m_keyframe_create", "argtypes", [_P]),
            ("ref_m_search_by_bow_kf_f", "argtypes", [_P, _P, _P, _F, _I, _P]),
            ("ref_m_search_by_bow_kf_kf", "argtypes", [_P, _P, _P, _I, _F, _I, _P]),
            ("ref_m_distinctive_desc", "argtypes", [_P, _P, _P, _I, _P, _P]),
            ("ref_m_search_by_projection_local", "argtypes", [_P, _P, _I, _P, _P, _P, _P, _P, _P, _P, _F, _F, _P]),
            ("ref_m_search_by_projection_last", "argtypes", [_P, _P, _P, _F, _F, _I, _P]),
            ("ref_m_search_for_triangulation", "argtypes", [_P, _P, _P, _I, _F, _I, _P]),
            ("ref_m_frame_set_point_geometry", "argtypes", [_P, _P, _P, _P, _P]),
            ("ref_m_frame_points_take_descriptors", "argtypes", [_P, _P]),
            ("ref_m_search_by_projection_reloc", "argtypes", [_P, _P, _P, _P, _F, _I, _F, _I, _P]),
            ("ref_m_search_by_projection_sim3", "argtypes", [_P, _P, _P, _P, _P, _I, _I, _F, _P]),
            ("ref_m_is_in_frustum", "argtypes", [_P, _P, _F] + [_P] * 6),
            ("ref_m_kfdb_add", "argtypes", [_P, _P]),
            ("ref_m_detect_reloc", "argtypes", [_P, _P, _P, _I, _P, _P, _P, _P]),
            ("ref_m_detect_loop", "argtypes", [_P, _P, _P, _I, _P, _F, _P, _P, _P, _P]),
            ("ref_m_frame_id", "argtypes", [_P]),
            ("ref_m_keyframe_id", "argtypes", [_P]),
            ("ref_m_fuse", "argtypes", [_P, _P, _P, _P, _P, _I, _F, _P, _P, _P]),
            ("ref_m_fuse_sim3", "argtypes", [_P, _P, _P, _P, _I, _F, _P, _P]),
            ("ref_m_search_by_sim3", "argtypes", [_P, _P, _P, _P, _P, _F, _P]),
            ("ref_m_search_by_projection_ff", "argtypes", [_P, _P, _P, _I, _F, _I, _P]),
        ]
        for name, kind, val in sig:          # the drop-in build leaves out the routines its shim does not cover
            if hasattr(L, name):
                setattr(getattr(L, name), kind, val)
        L._m_ready = True
    return L


class World:
    """One object graph: vocabulary (optional text file), Map, KeyFrameDatabase, and the extractor frames are made with."""

    def __init__(self, voc_path=None, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20):
        self._L = _lib()
        self._w = self._L.ref_m_world_create(os.fsencode(voc_path) if voc_path else b"")
        if not self._w:
            raise ValueError("vocabulary load failed")
        self.extractor = RE.RefExtractor(nfeatures, scaleFactor, nlevels, scoreType, fastTh, lib_path=LIB)

    def frame(self, image, K, D, reset_static=False):
        return RefFrame(self, image, K, D, reset_static)

    def distinctive(self, keyframes, idx):
        """MapPoint::ComputeDistinctiveDescriptors for a point observed at keypoint idx[k] of keyframes[k] ->
        (mDescriptor, order): `order` = the sequence in which the reference's std::map<KeyFrame*,..> visits them."""
        n = len(keyframes)
        arr = (C.c_void_p * n)(*[k for k in keyframes])
        ix = np.ascontiguousarray(idx, np.int32); out = np.zeros(32, np.uint8); order = np.zeros(n, np.int32)
        self._L.ref_m_distinctive_desc(self._w, arr, ix.ctypes.data, n, out.ctypes.data, order.ctypes.data)
        return out, order


class RefFrame:
    def __init__(self, world: World, image, K, D, reset_static):
        self.world, self._L = world, world._L
        img = np.ascontiguousarray(image, np.uint8)
        K = np.ascontiguousarray(K, np.float32).reshape(9); D = np.ascontiguousarray(D, np.float32).reshape(5)
        self._f = self._L.ref_m_frame_create(world._w, world.extractor._h, img.ctypes.data, img.shape[1], img.shape[0],
                                             K.ctypes.data, D.ctypes.data, int(reset_static))
        self.n = self._L.ref_m_frame_n(self._f)
        b = np.zeros(4, np.int32); inv = np.zeros(2, np.float32)
        self._L.ref_m_frame_bounds(self._f, b.ctypes.data, inv.ctypes.data)
        self.bounds, self.inv = tuple(int(v) for v in b), (float(inv[0]), float(inv[1]))
        n1 = max(self.n, 1)
        self.kps = np.zeros(n1, RE.KP_DTYPE); self.xu = np.zeros(n1, np.float32); self.yu = np.zeros(n1, np.float32)
        self.desc = np.zeros((n1, 32), np.uint8)
        self._L.ref_m_frame_read(self._f, self.kps.ctypes.data, self.xu.ctypes.data, self.yu.ctypes.data,
                                 self.desc.ctypes.data)
        self.kps, self.xu, self.yu, self.desc = self.kps[:self.n], self.xu[:self.n], self.yu[:self.n], self.desc[:self.n]
        self._kf = None

    def grid(self):
        cs = np.zeros(64 * 48 + 1, np.int32); items = np.zeros(max(self.n, 1), np.int32)
        self._L.ref_m_frame_grid(self._f, cs.ctypes.data, items.ctypes.data)
        return cs, items[:cs[-1]]

    def features_in_area(self, x, y, r, minl=-1, maxl=-1):
        out = np.zeros(max(self.n, 1), np.int32)
        n = self._L.ref_m_features_in_area(self._f, x, y, r, minl, maxl, out.ctypes.data)
        return out[:n].copy()

    def set_valid(self, valid):
        v = np.ascontiguousarray(valid, np.uint8)
        self._L.ref_m_frame_set_valid(self._f, v.ctypes.data)
        self._kf = None

    def compute_bow(self):
        self._L.ref_m_frame_compute_bow(self._f)
        n1 = max(self.n, 1)
        w = np.zeros(n1, np.int32); v = np.zeros(n1, np.float64); nw = np.zeros(1, np.int32)
        nd = np.zeros(n1, np.int32); st = np.zeros(n1 + 1, np.int32); it = np.zeros(n1, np.int32); nn = np.zeros(1, np.int32)
        self._L.ref_m_frame_bow(self._f, w.ctypes.data, v.ctypes.data, nw.ctypes.data, nd.ctypes.data, st.ctypes.data,
                                it.ctypes.data, nn.ctypes.data)
        nw, nn = int(nw[0]), int(nn[0])
        return (w[:nw].copy(), v[:nw].copy()), (nd[:nn].copy(), st[:nn + 1].copy(), it[:st[nn]].copy())

    def keyframe(self):
        """KeyFrame(F, map, db): call after set_valid / compute_bow."""
        if self._kf is None:
            self._kf = self._L.ref_m_keyframe_create(self._f)
        return self._kf

    def set_point_geometry(self, pos, min_dist, max_dist, normal=None):
        """World position, scale-invariance distances and normal of the MapPoints set_valid created (per keypoint)."""
        g = lambda a: np.ascontiguousarray(a, np.float32)
        pos, mn, mx = g(pos), g(min_dist), g(max_dist)
        nr = g(normal) if normal is not None else None
        self._L.ref_m_frame_set_point_geometry(self._f, pos.ctypes.data, mn.ctypes.data, mx.ctypes.data,
                                               nr.ctypes.data if nr is not None else None)

    def points_take_descriptors(self):
        """Each MapPoint of this frame observes its own keypoint in the frame's keyframe -> GetDescriptor() = that row."""
        self._L.ref_m_frame_points_take_descriptors(self._f, self.keyframe())


def search_for_initialization(f1: RefFrame, f2: RefFrame, prev_xy, window, ratio, check_ori):
    prev = np.ascontiguousarray(prev_xy, np.float32).copy()
    m = np.zeros(max(f1.n, 1), np.int32)
    n = _lib().ref_m_search_for_initialization(f1._f, f2._f, prev.ctypes.data, m.ctypes.data, window, ratio, int(check_ori))
    return n, m[:f1.n], prev


def window_search(f1: RefFrame, f2: RefFrame, window, minl, maxl, ratio, check_ori):
    m = np.zeros(max(f2.n, 1), np.int32)
    n = _lib().ref_m_window_search(f1._f, f2._f, window, minl, maxl, ratio, int(check_ori), m.ctypes.data)
    return n, m[:f2.n]


def search_by_bow(f1: RefFrame, f2: RefFrame, variant, ratio, check_ori):
    L = _lib()
    if variant == "kf_f":
        m = np.zeros(max(f2.n, 1), np.int32)
        n = L.ref_m_search_by_bow_kf_f(f1.keyframe(), f1._f, f2._f, ratio, int(check_ori), m.ctypes.data)
        return n, m[:f2.n]
    m = np.zeros(max(f1.n, 1), np.int32)
    n = L.ref_m_search_by_bow_kf_kf(f1.keyframe(), f2.keyframe(), f2._f, f1.n, ratio, int(check_ori), m.ctypes.data)
    return n, m[:f1.n]


def search_by_projection_local(f: RefFrame, desc_frame: RefFrame, u, v, level, view_cos, in_view, bad, desc_idx, th,
                               ratio):
    """SearchByProjection(F, vpMapPoints, th): point k carries descriptor desc_idx[k] of desc_frame."""
    g = lambda a, t: np.ascontiguousarray(a, t)
    u, v, vc = g(u, np.float32), g(v, np.float32), g(view_cos, np.float32)
    lv, di = g(level, np.int32), g(desc_idx, np.int32)
    iv, bd = g(in_view, np.uint8), g(bad, np.uint8)
    m = np.zeros(max(f.n, 1), np.int32)
    n = _lib().ref_m_search_by_projection_local(f._f, desc_frame._f, len(u), u.ctypes.data, v.ctypes.data, lv.ctypes.data,
                                                vc.ctypes.data, iv.ctypes.data, bd.ctypes.data, di.ctypes.data, th, ratio,
                                                m.ctypes.data)
    return n, m[:f.n]


def search_by_projection_last(cur: RefFrame, last: RefFrame, pos, th, ratio, check_ori):
    """SearchByProjection(CurrentFrame, LastFrame, th) with the current pose at identity; pos [last.n, 3]."""
    pos = np.ascontiguousarray(pos, np.float32)
    m = np.zeros(max(cur.n, 1), np.int32)
    n = _lib().ref_m_search_by_projection_last(cur._f, last._f, pos.ctypes.data, th, ratio, int(check_ori), m.ctypes.data)
    return n, m[:cur.n]


def search_for_triangulation(f1: RefFrame, f2: RefFrame, F12, ratio, check_ori):
    """SearchForTriangulation between the keyframes made from f1 and f2 (after set_valid / compute_bow)."""
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    m = np.zeros(max(f1.n, 1), np.int32)
    n = _lib().ref_m_search_for_triangulation(f1.keyframe(), f2.keyframe(), F.ctypes.data, f1.n, ratio, int(check_ori),
                                              m.ctypes.data)
    return n, m[:f1.n]


def search_by_projection_reloc(cur: RefFrame, kf_frame: RefFrame, already, th, orb_dist, ratio, check_ori):
    """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), current pose = identity; pKF = the keyframe of
    kf_frame (set_valid, set_point_geometry, points_take_descriptors done)."""
    a = np.ascontiguousarray(already, np.uint8)
    m = np.zeros(max(cur.n, 1), np.int32)
    n = _lib().ref_m_search_by_projection_reloc(cur._f, kf_frame.keyframe(), kf_frame._f, a.ctypes.data, th, int(orb_dist),
                                                ratio, int(check_ori), m.ctypes.data)
    return n, m[:cur.n]


def search_by_projection_sim3(kf_frame: RefFrame, src: RefFrame, sel, pre, th, ratio):
    """SearchByProjection(pKF, Scw = identity, vpPoints, vpMatched, th): vpPoints = src's MapPoints sel[k];
    pre[k] >= 0 pre-assigns point k to keyframe keypoint pre[k].  Returns (n, matched[idx] = k or -1)."""
    sel = np.ascontiguousarray(sel, np.int32); pre = np.ascontiguousarray(pre, np.int32)
    m = np.zeros(max(kf_frame.n, 1), np.int32)
    n = _lib().ref_m_search_by_projection_sim3(kf_frame.keyframe(), kf_frame._f, src._f, sel.ctypes.data, pre.ctypes.data,
                                               len(sel), int(th), ratio, m.ctypes.data)
    return n, m[:kf_frame.n]


def is_in_frustum(view: RefFrame, src: RefFrame, cos_limit=0.5):
    """Frame::isInFrustum of `view` (pose identity) for the MapPoint on every keypoint of `src` ->
    (ret, in_view, u, v, level, view_cos), zeros where src has no MapPoint."""
    n = max(src.n, 1)
    ret = np.zeros(n, np.uint8); iv = np.zeros(n, np.uint8); u = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
    lv = np.zeros(n, np.int32); vc = np.zeros(n, np.float32)
    _lib().ref_m_is_in_frustum(view._f, src._f, cos_limit, ret.ctypes.data, iv.ctypes.data, u.ctypes.data, v.ctypes.data,
                               lv.ctypes.data, vc.ctypes.data)
    k = src.n
    return ret[:k], iv[:k], u[:k], v[:k], lv[:k], vc[:k]


def fuse(kf_frame: RefFrame, src: RefFrame, sel, th):
    """Fuse(pKF, vpMapPoints, th): pKF = kf_frame's keyframe, vpMapPoints = src's MapPoints sel[k] ->
    (nFused, added[k], went_bad[k], repl[k]) - see oracle/ref_orbmatcher.cc::ref_m_fuse."""
    sel = np.ascontiguousarray(sel, np.int32)
    n = len(sel)
    added = np.zeros(n, np.int32); bad = np.zeros(n, np.uint8); repl = np.zeros(n, np.int32)
    r = _lib().ref_m_fuse(kf_frame.keyframe(), kf_frame._f, src._f, src.keyframe(), sel.ctypes.data, n, th,
                          added.ctypes.data, bad.ctypes.data, repl.ctypes.data)
    return r, added, bad, repl


def fuse_sim3(kf_frame: RefFrame, src: RefFrame, sel, th):
    """Fuse(pKF, Scw = identity, vpPoints, th) -> (nFused, added[k], target_went_bad[idx])."""
    sel = np.ascontiguousarray(sel, np.int32)
    added = np.zeros(len(sel), np.int32); tbad = np.zeros(max(kf_frame.n, 1), np.uint8)
    r = _lib().ref_m_fuse_sim3(kf_frame.keyframe(), kf_frame._f, src._f, sel.ctypes.data, len(sel), th,
                               added.ctypes.data, tbad.ctypes.data)
    return r, added, tbad[:kf_frame.n]


def search_by_sim3(f1: RefFrame, f2: RefFrame, pre12, th):
    """SearchBySim3(pKF1, pKF2, vpMatches12, 1, I, 0, th) -> (nFound, out12[i1] = keypoint of KF2 or -1)."""
    pre = np.ascontiguousarray(pre12, np.int32)
    out = np.zeros(max(f1.n, 1), np.int32)
    n = _lib().ref_m_search_by_sim3(f1.keyframe(), f1._f, f2.keyframe(), f2._f, pre.ctypes.data, th, out.ctypes.data)
    return n, out[:f1.n]


def search_by_projection_ff(f1: RefFrame, f2: RefFrame, pos, window, ratio, check_ori):
    """SearchByProjection(F1, F2, windowSize, matches2), F2 pose = identity; pos [f1.n, 3] -> (n, matches21)."""
    pos = np.ascontiguousarray(pos, np.float32)
    m = np.zeros(max(f2.n, 1), np.int32)
    n = _lib().ref_m_search_by_projection_ff(f1._f, f2._f, pos.ctypes.data, int(window), ratio, int(check_ori), m.ctypes.data)
    return n, m[:f2.n]


def kfdb_add(world: World, frames):
    """KeyFrameDatabase::add for the keyframes of `frames` (after compute_bow)."""
    for f in frames:
        _lib().ref_m_kfdb_add(world._w, f.keyframe())


def _kf_array(frames):
    return (C.c_void_p * len(frames))(*[f.keyframe() for f in frames])


def detect_relocalisation_candidates(world: World, query: RefFrame, frames):
    """KeyFrameDatabase::DetectRelocalisationCandidates(F) -> (ncand, touched, words, score, cand) per keyframe of `frames`."""
    n = len(frames)
    t = np.zeros(n, np.uint8); w = np.zeros(n, np.int32); s = np.zeros(n, np.float32); c = np.zeros(n, np.uint8)
    assert _lib().ref_m_frame_id(query._f) != 0
    r = _lib().ref_m_detect_reloc(world._w, query._f, _kf_array(frames), n, t.ctypes.data, w.ctypes.data, s.ctypes.data, c.ctypes.data)
    return r, t, w, s, c


def detect_loop_candidates(world: World, query: RefFrame, frames, connected, min_score):
    """KeyFrameDatabase::DetectLoopCandidates(pKF, minScore); connected[k] != 0 -> frames[k] is a covisibility neighbour."""
    n = len(frames)
    con = np.ascontiguousarray(connected, np.uint8)
    t = np.zeros(n, np.uint8); w = np.zeros(n, np.int32); s = np.zeros(n, np.float32); c = np.zeros(n, np.uint8)
    assert _lib().ref_m_keyframe_id(query.keyframe()) != 0
    r = _lib().ref_m_detect_loop(world._w, query.keyframe(), _kf_array(frames), n, con.ctypes.data, min_score, t.ctypes.data,
                                 w.ctypes.data, s.ctypes.data, c.ctypes.data)
    return r, t, w, s, c
