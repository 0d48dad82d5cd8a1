"""Host-side mirror of ORB_SLAM::ORBmatcher (reference include/ORBmatcher.h:37-107) on top of the C ABI of
include/orbm.h.  Same constructor (nnratio, checkOri), same method names and argument meaning; Frame / KeyFrame objects
are replaced by :class:`MatchFrame` (flat numpy arrays) uploaded into a device-resident :class:`FrameSet`.
numpy only owns memory: every distance and every match decision comes from the CUDA kernels.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from ._lib import lib

TH_HIGH, TH_LOW, HISTO_LENGTH = 100, 50, 30      # reference src/ORBmatcher.cc:40-42
FRAME_GRID_COLS, FRAME_GRID_ROWS = 64, 48        # reference include/Frame.h:35-36
KEY_NONE = 0xFFFFFFFF


class OrbmError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"orbm error {code}: {msg}")
        self.code = code


class _Frame(C.Structure):
    _fields_ = [("n", C.c_int32), ("x", C.c_void_p), ("y", C.c_void_p), ("angle", C.c_void_p), ("octave", C.c_void_p),
                ("desc", C.c_void_p), ("valid", C.c_void_p), ("min_x", C.c_float), ("min_y", C.c_float),
                ("inv_w", C.c_float), ("inv_h", C.c_float)]


class _FeatVec(C.Structure):
    _fields_ = [("nnodes", C.c_int32), ("node_id", C.c_void_p), ("start", C.c_void_p), ("items", C.c_void_p)]


class Camera(C.Structure):
    """orbm_camera: the CV_32F intrinsics / distortion the reference stores (src/Tracking.cc:52-71) + image size."""
    _fields_ = [(n, C.c_float) for n in ("fx", "fy", "cx", "cy", "k1", "k2", "p1", "p2", "k3")] + \
               [("width", C.c_int32), ("height", C.c_int32)]


class FrustumView(C.Structure):
    """orbm_frustum_view: a Frame's pose (mRcw row major, mtcw, mOw), intrinsics, image bounds and scale pyramid."""
    _fields_ = [("Rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("Ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("min_x", C.c_int32), ("max_x", C.c_int32), ("min_y", C.c_int32), ("max_y", C.c_int32),
                ("nlevels", C.c_int32), ("scale_factors", C.c_float * 16)]

    @classmethod
    def make(cls, Rcw, tcw, Ow, fx, fy, cx, cy, bounds, scale_factors):
        sf = np.zeros(16, np.float32); sf[:len(scale_factors)] = scale_factors
        return cls((C.c_float * 9)(*np.asarray(Rcw, np.float32).ravel()), (C.c_float * 3)(*np.asarray(tcw, np.float32)),
                   (C.c_float * 3)(*np.asarray(Ow, np.float32)), fx, fy, cx, cy, *[int(b) for b in bounds],
                   len(scale_factors), (C.c_float * 16)(*sf))


class _Points(C.Structure):
    _fields_ = [("n", C.c_int32), ("u", C.c_void_p), ("v", C.c_void_p), ("radius", C.c_void_p),
                ("min_level", C.c_void_p), ("max_level", C.c_void_p), ("desc", C.c_void_p), ("angle", C.c_void_p),
                ("valid", C.c_void_p)]


def _bind(L):
    if getattr(L, "_orbm_bound", False):
        return L
    vp, i32, f32 = C.c_void_p, C.c_int32, C.c_float
    L.orbm_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.orbm_destroy.argtypes = [vp]; L.orbm_destroy.restype = None
    L.orbm_last_error.argtypes = [vp]; L.orbm_last_error.restype = C.c_char_p
    L.orbm_kernel_launches.argtypes = [vp]; L.orbm_kernel_launches.restype = C.c_int64
    L.orbm_last_kernel_ms.argtypes = [vp]; L.orbm_last_kernel_ms.restype = C.c_float
    L.orbm_descriptor_distance.argtypes = [vp, vp, vp, C.c_int, vp]
    L.orbm_frameset_create.argtypes = [vp, vp, vp, C.c_int, C.POINTER(vp)]
    L.orbm_frameset_destroy.argtypes = [vp]; L.orbm_frameset_destroy.restype = None
    L.orbm_frameset_set_valid.argtypes = [vp, C.c_int, vp]
    L.orbm_frameset_grid.argtypes = [vp, C.c_int, vp, vp]
    L.orbm_frameset_fill_valid.argtypes = [vp, C.c_int]
    L.orbm_image_bounds.argtypes = [C.POINTER(Camera), vp, vp]
    L.orbm_frameset_from_extractor.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.POINTER(Camera), vp, C.POINTER(vp)]
    L.orbm_frameset_counts.argtypes = [vp, vp]
    L.orbm_frameset_keypoints.argtypes = [vp, C.c_int, vp, vp, vp, vp, vp]
    L.orbm_features_in_area.argtypes = [vp, C.c_int, f32, f32, f32, C.c_int, C.c_int, vp, C.c_int, C.POINTER(C.c_int)]
    L.orbm_vocabulary_create.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp, C.POINTER(vp)]
    L.orbm_vocabulary_load_text.argtypes = [vp, C.c_char_p, C.POINTER(vp)]
    L.orbm_vocabulary_destroy.argtypes = [vp]; L.orbm_vocabulary_destroy.restype = None
    L.orbm_vocabulary_info.argtypes = [vp, vp]
    L.orbm_vocabulary_transform.argtypes = [vp, vp, C.c_int, C.c_int, vp, vp, vp, vp, vp, vp, vp]
    L.orbm_frameset_compute_bow.argtypes = [vp, vp, C.c_int]
    L.orbm_frameset_bow_sizes.argtypes = [vp, vp, vp, vp]
    L.orbm_frameset_bowvec.argtypes = [vp, C.c_int, vp, vp]
    L.orbm_frameset_featvec.argtypes = [vp, C.c_int, vp, vp, vp]
    L.orbm_bow_score_l1.argtypes = [vp, vp, vp, vp, C.c_int, vp]
    L.orbm_is_in_frustum.argtypes = [vp, vp, C.c_int, vp, vp, vp, vp, C.c_int, f32, vp, vp, vp, vp, vp]
    L.orbm_distinctive_descriptors.argtypes = [vp, vp, vp, C.c_int, vp, vp]
    L.orbm_search_for_initialization.argtypes = [vp, vp, vp, vp, vp, C.c_int, vp, C.c_int, C.c_int, f32, C.c_int, vp, vp]
    L.orbm_window_search.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, f32, C.c_int, C.c_int, C.c_int, vp, C.c_int, vp]
    L.orbm_search_by_projection_frames.argtypes = [vp, vp, vp, vp, vp, C.c_int, vp, C.c_int, C.c_int, f32, vp, vp]
    L.orbm_search_by_projection.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, f32, C.c_int, vp, C.c_int, vp]
    L.orbm_bowdb_create.argtypes = [vp, vp, C.POINTER(vp)]
    L.orbm_bowdb_destroy.argtypes = [vp]
    L.orbm_bowdb_destroy.restype = None
    L.orbm_bowdb_query.argtypes = [vp, vp, vp, C.c_int, vp, vp, vp, vp, vp]
    L.orbm_search_by_sim3.argtypes = [vp, vp, vp, vp, vp, vp, C.c_int, C.c_int, vp, C.c_int, vp]
    L.orbm_search_by_bow.argtypes = [vp, vp, vp, vp, vp, C.c_int, C.c_int, f32, C.c_int, vp, C.c_int, vp]
    L.orbm_search_for_triangulation.argtypes = [vp, vp, vp, vp, vp, C.c_int, vp, vp, C.c_int, C.c_int, vp, C.c_int, vp]
    L.orbm_knn2.argtypes = [vp, vp, C.c_int, vp, C.c_int64, C.c_int64, vp]
    L.orbm_knn2_device.argtypes = [vp, vp, C.c_int, vp, C.c_int64, C.c_int64, vp, vp]
    L.orbm_knn2_device64.argtypes = [vp, vp, C.c_int, vp, C.c_int64, C.c_int64, vp, vp]
    L.orbm_synchronize.argtypes = [vp]
    L.orbm_comm_release.argtypes = [vp, vp]
    L.orbm_comm_release.restype = None
    L.orbm_knn2_device_static.argtypes = [vp, vp, C.c_int, vp, C.c_int64, C.c_int64, vp, vp]
    L.orbm_merge_top2_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, f32, vp, vp, vp]
    L.orbm_knn2_sharded.argtypes = [vp, vp, C.c_int, vp, C.c_int64, C.c_int64, C.c_int, vp, vp, C.c_int, f32, vp, vp, vp]
    L._orbm_bound = True
    return L


@dataclass
class MatchFrame:
    """What the matcher reads from a Frame / KeyFrame (reference include/Frame.h, include/KeyFrame.h)."""
    x: np.ndarray            # mvKeysUn[i].pt.x
    y: np.ndarray
    angle: np.ndarray        # mvKeysUn[i].angle
    octave: np.ndarray
    desc: np.ndarray         # mDescriptors [n,32] uint8
    valid: np.ndarray | None = None   # per routine: keypoint has a good MapPoint
    min_x: float = 0.0       # mnMinX, mnMinY, mnMaxX, mnMaxY (Frame::ComputeImageBounds)
    min_y: float = 0.0
    max_x: float = 640.0
    max_y: float = 480.0
    featvec: tuple | None = None      # (node_id[nn], start[nn+1], items[]) = DBoW2::FeatureVector flattened
    _keep: list = field(default_factory=list, repr=False)

    def __post_init__(self):
        self.x = np.ascontiguousarray(self.x, np.float32); self.y = np.ascontiguousarray(self.y, np.float32)
        self.angle = np.ascontiguousarray(self.angle, np.float32)
        self.octave = np.ascontiguousarray(self.octave, np.int32)
        self.desc = np.ascontiguousarray(self.desc, np.uint8).reshape(-1, 32)
        if self.valid is not None:
            self.valid = np.ascontiguousarray(self.valid, np.uint8)
        if self.featvec is not None:
            self.featvec = tuple(np.ascontiguousarray(a, np.int32) for a in self.featvec)

    @property
    def n(self):
        return len(self.x)

    @classmethod
    def from_keypoints(cls, kps, desc, width, height, **kw):
        """From the extractor's output (no undistortion: k1 == 0, src/Frame.cc:290-294; bounds :338-344)."""
        return cls(kps["x"], kps["y"], kps["angle"], kps["octave"], desc, max_x=float(width), max_y=float(height), **kw)

    def grid_inv(self):
        # mfGridElementWidthInv = (float)FRAME_GRID_COLS / (float)(mnMaxX - mnMinX)   (src/Frame.cc:75-76)
        iw = np.float32(FRAME_GRID_COLS) / np.float32(np.float32(self.max_x) - np.float32(self.min_x))
        ih = np.float32(FRAME_GRID_ROWS) / np.float32(np.float32(self.max_y) - np.float32(self.min_y))
        return float(iw), float(ih)

    def c_struct(self):
        iw, ih = self.grid_inv()
        v = self.valid.ctypes.data if self.valid is not None and self.n else None
        p = lambda a: a.ctypes.data if self.n else None
        return _Frame(self.n, p(self.x), p(self.y), p(self.angle), p(self.octave), p(self.desc), v, self.min_x,
                      self.min_y, iw, ih)

    def c_featvec(self):
        nid, st, it = self.featvec
        return _FeatVec(len(nid), nid.ctypes.data if len(nid) else None, st.ctypes.data, it.ctypes.data if len(it) else None)


class Matcher:
    """Owns the orbm handle (device, stream, workspaces).  One per host thread."""

    def __init__(self, device=0):
        self._L = _bind(lib())
        self._h = C.c_void_p()
        rc = self._L.orbm_create(device, C.byref(self._h))
        if rc != 0:
            self._h = None
            raise OrbmError(rc, (self._L.orbm_last_error(None) or b"").decode())

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbm_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise OrbmError(rc, (self._L.orbm_last_error(self._h) or b"").decode())

    def kernel_launches(self):
        return int(self._L.orbm_kernel_launches(self._h))

    def last_kernel_ms(self):
        return float(self._L.orbm_last_kernel_ms(self._h))

    def upload(self, frames):
        return FrameSet(self, frames)

    # M1
    def descriptor_distance(self, a, b):
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32); b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        assert a.shape == b.shape
        out = np.empty(len(a), np.int32)
        self._check(self._L.orbm_descriptor_distance(self._h, a.ctypes.data, b.ctypes.data, len(a), out.ctypes.data))
        return out

    # M8
    def is_in_frustum(self, views, pos, normal, min_dist, max_dist, cos_limit=0.5):
        """Frame::isInFrustum (src/Frame.cc:136-197) for every (view, map point) pair ->
        (in_view u8, u, v, level, view_cos), each [nviews, n]."""
        views = list(views)
        arr = (FrustumView * len(views))(*views)
        pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3); nrm = np.ascontiguousarray(normal, np.float32)
        mn = np.ascontiguousarray(min_dist, np.float32); mx = np.ascontiguousarray(max_dist, np.float32)
        n, nv = len(pos), len(views)
        if nrm.size != 3 * n or len(mn) != n or len(mx) != n:
            raise ValueError("per-point arrays disagree in length")
        iv = np.zeros((nv, n), np.uint8); u = np.zeros((nv, n), np.float32); v = np.zeros((nv, n), np.float32)
        lv = np.zeros((nv, n), np.int32); vc = np.zeros((nv, n), np.float32)
        self._check(self._L.orbm_is_in_frustum(self._h, arr, nv, pos.ctypes.data, nrm.ctypes.data, mn.ctypes.data,
                                               mx.ctypes.data, n, cos_limit, iv.ctypes.data, u.ctypes.data,
                                               v.ctypes.data, lv.ctypes.data, vc.ctypes.data))
        return iv, u, v, lv, vc

    def distinctive_descriptors(self, desc, start):
        """MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:185-250) for a batch of map points: point p
        observes desc[start[p]:start[p+1]].  -> (best index per point, chosen descriptor per point [npoints, 32])."""
        d = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32); st = np.ascontiguousarray(start, np.int32)
        npts = len(st) - 1
        if npts < 0 or (npts and int(st[-1]) != len(d)):
            raise ValueError("start must have npoints + 1 entries ending at len(desc)")
        best = np.zeros(max(npts, 0), np.int32); out = np.zeros((max(npts, 0), 32), np.uint8)
        self._check(self._L.orbm_distinctive_descriptors(self._h, d.ctypes.data if len(d) else None, st.ctypes.data,
                                                         npts, best.ctypes.data, out.ctypes.data))
        return best, out

    def knn2(self, q, db, idx_base=0):
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32); db = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
        out = np.empty((len(q), 2), np.uint32)
        self._check(self._L.orbm_knn2(self._h, q.ctypes.data, len(q), db.ctypes.data, len(db), idx_base, out.ctypes.data))
        return out

    def knn2_device(self, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream=None):
        self._check(self._L.orbm_knn2_device(self._h, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream))

    def knn2_device64(self, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream=None):
        """64-bit keys (dist << 40 | index) for databases beyond 2^23 descriptors; top2: [nq, 2] uint64 on the device."""
        self._check(self._L.orbm_knn2_device64(self._h, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream))

    def synchronize(self):
        """orbm_synchronize: waits for the handle's stream, raises on a deferred knn2 / NCCL error."""
        self._check(self._L.orbm_synchronize(self._h))

    def knn2_device_static(self, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream=None):
        """orbm_knn2_device without any device allocation (static scratch, include/orbm.h)."""
        self._check(self._L.orbm_knn2_device_static(self._h, q_ptr, nq, db_ptr, ndb, idx_base, top2_ptr, stream))

    def merge_top2_device(self, gathered_ptr, nshards, nq, th_dist, ratio, top2_ptr, match_ptr, stream=None):
        self._check(self._L.orbm_merge_top2_device(self._h, gathered_ptr, nshards, nq, th_dist, ratio, top2_ptr,
                                                   match_ptr, stream))


class ORBVocabulary:
    """ORB_SLAM::ORBVocabulary = DBoW2::TemplatedVocabulary<FORB::TDescriptor, FORB> (reference
    include/ORBVocabulary.h:31-32) resident on the device.  Node and word ids are the reference's
    (loadFromTextFile numbering, Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1337-1424)."""

    def __init__(self, matcher: Matcher, handle):
        self.m, self._L, self._v = matcher, matcher._L, handle
        info = np.zeros(6, np.int32)
        matcher._check(self._L.orbm_vocabulary_info(self._v, info.ctypes.data))
        self.k, self.L, self.scoring, self.weighting, self.nwords, self.nnodes = (int(x) for x in info)

    @classmethod
    def loadFromTextFile(cls, matcher: Matcher, filename):
        v = C.c_void_p()
        matcher._check(matcher._L.orbm_vocabulary_load_text(matcher._h, os.fsencode(filename), C.byref(v)))
        return cls(matcher, v)

    @classmethod
    def from_arrays(cls, matcher: Matcher, k, L, scoring, weighting, parent, desc, weight, is_leaf):
        parent = _i32(parent); desc = np.ascontiguousarray(desc, np.uint8)
        weight = np.ascontiguousarray(weight, np.float64); is_leaf = np.ascontiguousarray(is_leaf, np.uint8)
        if not (len(parent) == len(weight) == len(is_leaf) and desc.size == 32 * len(parent)):
            raise ValueError("vocabulary arrays disagree in length")
        v = C.c_void_p()
        matcher._check(matcher._L.orbm_vocabulary_create(matcher._h, k, L, scoring, weighting, len(parent),
                                                         parent.ctypes.data, desc.ctypes.data, weight.ctypes.data,
                                                         is_leaf.ctypes.data, C.byref(v)))
        return cls(matcher, v)

    def size(self):
        return self.nwords

    def getBranchingFactor(self):
        return self.k

    def getDepthLevels(self):
        return self.L

    def transform(self, features, levelsup=4):
        """transform(features, BowVector, FeatureVector, levelsup) (TemplatedVocabulary.h:1126-1194) of one
        descriptor set [n, 32] -> ((word_id, weight), (node_id, start, items)), ids ascending as in the std::maps."""
        d = np.ascontiguousarray(features, np.uint8).reshape(-1, 32)
        n = len(d)
        bw = np.zeros(max(n, 1), np.int32); bv = np.zeros(max(n, 1), np.float64)
        fn = np.zeros(max(n, 1), np.int32); fs = np.zeros(n + 1, np.int32); fi = np.zeros(max(n, 1), np.int32)
        nb = np.zeros(1, np.int32); nn = np.zeros(1, np.int32)
        self.m._check(self._L.orbm_vocabulary_transform(self._v, d.ctypes.data, n, levelsup, bw.ctypes.data,
                                                        bv.ctypes.data, nb.ctypes.data, fn.ctypes.data,
                                                        fs.ctypes.data, fi.ctypes.data, nn.ctypes.data))
        nb, nn = int(nb[0]), int(nn[0])
        return (bw[:nb].copy(), bv[:nb].copy()), (fn[:nn].copy(), fs[:nn + 1].copy(), fi[:fs[nn]].copy())

    def score(self, set1, set2, idx1, idx2):
        """TemplatedVocabulary::score for L1_NORM vocabularies (ScoringObject.cpp:23-67) between the BowVectors of
        frame pairs of two sets on which compute_bow has run."""
        if self.scoring != 0:
            raise ValueError("only L1_NORM scoring (the ORB_SLAM vocabulary's) is implemented")
        i1, i2 = _i32(idx1), _i32(idx2)
        out = np.zeros(len(i1), np.float64)
        self.m._check(self._L.orbm_bow_score_l1(set1._s, set2._s, i1.ctypes.data, i2.ctypes.data, len(i1),
                                                out.ctypes.data))
        return out

    def close(self):
        if getattr(self, "_v", None) and getattr(self.m, "_h", None):
            self._L.orbm_vocabulary_destroy(self._v)
        self._v = None

    __del__ = close


def image_bounds(camera: Camera):
    """Frame::ComputeImageBounds + grid element sizes (src/Frame.cc:320-349, :75-76), computed by the library (host
    side, no device needed): ((mnMinX, mnMaxX, mnMinY, mnMaxY), (mfGridElementWidthInv, mfGridElementHeightInv))."""
    L = _bind(lib())
    b = np.zeros(4, np.int32); inv = np.zeros(2, np.float32)
    if L.orbm_image_bounds(C.byref(camera), b.ctypes.data, inv.ctypes.data) != 0:
        raise ValueError("bad camera")
    return tuple(int(v) for v in b), (float(inv[0]), float(inv[1]))


class FrameSet:
    """Frames resident on the device with their 64x48 grids (reference src/Frame.cc:108-122)."""

    def __init__(self, matcher: Matcher, frames):
        self.m = matcher
        self.frames = list(frames)
        self._L = matcher._L
        arr = (_Frame * len(self.frames))(*[f.c_struct() for f in self.frames])
        fvs = None
        if all(f.featvec is not None for f in self.frames):
            fvs = (_FeatVec * len(self.frames))(*[f.c_featvec() for f in self.frames])
        self._s = C.c_void_p()
        matcher._check(self._L.orbm_frameset_create(matcher._h, arr, fvs, len(self.frames), C.byref(self._s)))
        self.max_n = max(f.n for f in self.frames)
        self.nframes = len(self.frames)

    @classmethod
    def from_extractor(cls, matcher: Matcher, kps_ptr, desc_ptr, n_ptr, nframes, kp_stride, camera: Camera,
                       stream=None):
        """N1: the Frame constructor's glue on the device (src/Frame.cc:59-127) — builds the frame set
        from orbx_extract_batch_device's DEVICE outputs (undistort + grid), no host round trip of the features."""
        self = cls.__new__(cls)
        self.m, self._L, self.frames = matcher, matcher._L, None
        self._s = C.c_void_p()
        matcher._check(self._L.orbm_frameset_from_extractor(matcher._h, kps_ptr, desc_ptr, n_ptr, nframes, kp_stride,
                                                            C.byref(camera), stream, C.byref(self._s)))
        self.nframes = nframes
        self.counts = np.zeros(nframes, np.int32)
        matcher._check(self._L.orbm_frameset_counts(self._s, self.counts.ctypes.data))
        self.max_n = int(self.counts.max()) if nframes else 0
        return self

    def frame_n(self, frame):
        return self.frames[frame].n if self.frames is not None else int(self.counts[frame])

    def keypoints(self, frame):
        """(x, y, angle, octave, desc) of one resident frame (undistorted coordinates)."""
        n = self.frame_n(frame)
        x = np.zeros(n, np.float32); y = np.zeros(n, np.float32); a = np.zeros(n, np.float32)
        o = np.zeros(n, np.int32); d = np.zeros((n, 32), np.uint8)
        self.m._check(self._L.orbm_frameset_keypoints(self._s, frame, x.ctypes.data, y.ctypes.data, a.ctypes.data,
                                                      o.ctypes.data, d.ctypes.data))
        return x, y, a, o, d

    def compute_bow(self, vocabulary: "ORBVocabulary", levelsup=4):
        """N2: Frame::ComputeBoW (src/Frame.cc:279-286) for every resident frame, on the device; the set
        can then take part in SearchByBoW."""
        self.m._check(self._L.orbm_frameset_compute_bow(self._s, vocabulary._v, levelsup))
        self._has_fv = True

    def _sizes(self):
        nf = len(self.frames) if self.frames is not None else len(self.counts)
        w = np.zeros(nf, np.int32); n = np.zeros(nf, np.int32); i = np.zeros(nf, np.int32)
        self.m._check(self._L.orbm_frameset_bow_sizes(self._s, w.ctypes.data, n.ctypes.data, i.ctypes.data))
        return w, n, i

    def bow_vector(self, frame):
        """mBowVec of one frame: (word_id[], weight[]) in ascending word id."""
        nw = int(self._sizes()[0][frame])
        w = np.zeros(max(nw, 1), np.int32); v = np.zeros(max(nw, 1), np.float64)
        self.m._check(self._L.orbm_frameset_bowvec(self._s, frame, w.ctypes.data, v.ctypes.data))
        return w[:nw], v[:nw]

    def feature_vector(self, frame):
        """mFeatVec of one frame: (node_id[], start[], items[])."""
        _, nn, ni = self._sizes()
        nn, ni = int(nn[frame]), int(ni[frame])
        a = np.zeros(max(nn, 1), np.int32); s = np.zeros(nn + 1, np.int32); it = np.zeros(max(ni, 1), np.int32)
        self.m._check(self._L.orbm_frameset_featvec(self._s, frame, a.ctypes.data, s.ctypes.data, it.ctypes.data))
        return a[:nn], s, it[:ni]

    def close(self):
        if getattr(self, "_s", None) and getattr(self.m, "_h", None):
            self._L.orbm_frameset_destroy(self._s)
        self._s = None

    __del__ = close

    def set_valid(self, frame, valid):
        v = None if valid is None else np.ascontiguousarray(valid, np.uint8)
        self.m._check(self._L.orbm_frameset_set_valid(self._s, frame, None if v is None else v.ctypes.data))

    def fill_valid(self, value=True):
        self.m._check(self._L.orbm_frameset_fill_valid(self._s, int(bool(value))))

    def grid(self, frame):
        cs = np.zeros(FRAME_GRID_COLS * FRAME_GRID_ROWS + 1, np.int32)
        items = np.zeros(max(self.frame_n(frame), 1), np.int32)
        self.m._check(self._L.orbm_frameset_grid(self._s, frame, cs.ctypes.data, items.ctypes.data))
        return cs, items[: cs[-1]]

    def features_in_area(self, frame, x, y, r, min_level=-1, max_level=-1):
        cap = max(self.frame_n(frame), 1)
        out = np.zeros(cap, np.int32)
        n = C.c_int(0)
        self.m._check(self._L.orbm_features_in_area(self._s, frame, x, y, r, min_level, max_level, out.ctypes.data, cap,
                                                    C.byref(n)))
        return out[: n.value].copy()


def _i32(a):
    return np.ascontiguousarray(a, np.int32)


@dataclass
class ProjectedPoints:
    """Map points projected into a target frame (the SearchByProjection family)."""
    u: np.ndarray
    v: np.ndarray
    radius: np.ndarray
    min_level: np.ndarray
    max_level: np.ndarray
    desc: np.ndarray
    angle: np.ndarray | None = None
    valid: np.ndarray | None = None

    def __post_init__(self):
        f = lambda a: np.ascontiguousarray(a, np.float32)
        self.u, self.v, self.radius = f(self.u), f(self.v), f(self.radius)
        self.min_level, self.max_level = _i32(self.min_level), _i32(self.max_level)
        self.desc = np.ascontiguousarray(self.desc, np.uint8).reshape(-1, 32)
        if self.angle is not None:
            self.angle = f(self.angle)
        if self.valid is not None:
            self.valid = np.ascontiguousarray(self.valid, np.uint8)

    @property
    def n(self):
        return len(self.u)

    def c_struct(self):
        p = lambda a: a.ctypes.data if a is not None and self.n else None
        return _Points(self.n, p(self.u), p(self.v), p(self.radius), p(self.min_level), p(self.max_level), p(self.desc),
                       p(self.angle), p(self.valid))

    @classmethod
    def for_local_map(cls, proj_x, proj_y, view_cos, level, desc, scale_factors, th=1.0, valid=None):
        """Radius and level window of SearchByProjection(Frame&, vector<MapPoint*>&, th), src/ORBmatcher.cc:53-73,
        RadiusByViewingCos :127-133."""
        view_cos = np.asarray(view_cos, np.float32)
        level = _i32(level)
        r = np.where(view_cos.astype(np.float64) > 0.998, np.float32(2.5), np.float32(4.0)).astype(np.float32)
        if th != 1.0:
            r = (r * np.float32(th)).astype(np.float32)
        radius = (r * np.asarray(scale_factors, np.float32)[level]).astype(np.float32)
        return cls(proj_x, proj_y, radius, level - 1, level, desc, None, valid)

    @classmethod
    def for_last_frame(cls, u, v, octave, desc, angle, scale_factors, th, valid=None):
        """SearchByProjection(CurrentFrame, LastFrame, th), src/ORBmatcher.cc:1545-1550."""
        octave = _i32(octave)
        radius = (np.float32(th) * np.asarray(scale_factors, np.float32)[octave]).astype(np.float32)
        return cls(u, v, radius, octave - 1, octave + 1, desc, angle, valid)


    @classmethod
    def for_relocalisation(cls, u, v, level, desc, angle, scale_factors, th, valid=None):
        """SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist), src/ORBmatcher.cc:1669-1672: radius =
        th * mvScaleFactors[predicted level], levels [pred-1, pred+1]; use variant 'best' with th_dist = ORBdist."""
        level = _i32(level)
        radius = (np.float32(th) * np.asarray(scale_factors, np.float32)[level]).astype(np.float32)
        return cls(u, v, radius, level - 1, level + 1, desc, angle, valid)

    @classmethod
    def for_sim3(cls, u, v, level, desc, scale_factors, th, valid=None):
        """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th), src/ORBmatcher.cc:362-383: radius = th *
        GetScaleFactor(pred) (th is an int there), candidates on levels [pred-1, pred], no orientation check; use
        variant 'best' with th_dist = TH_LOW and checkOri off."""
        level = _i32(level)
        radius = (np.float32(int(th)) * np.asarray(scale_factors, np.float32)[level]).astype(np.float32)
        return cls(u, v, radius, level - 1, level, desc, None, valid)

    @classmethod
    def for_fuse(cls, u, v, level, desc, scale_factors, th, valid=None):
        """Fuse (src/ORBmatcher.cc:1078-1086, :1207-1215) and SearchBySim3 (:1353-1360, :1438-1445): radius = th *
        scale[predicted level] (th is a float there), candidates on levels [pred-1, pred]; use variant 'independent'."""
        level = _i32(level)
        radius = (np.float32(th) * np.asarray(scale_factors, np.float32)[level]).astype(np.float32)
        return cls(u, v, radius, level - 1, level, desc, None, valid)


def predict_scale_level(dist, min_distance, scale_factors):
    """`ratio = dist / minDistance; lower_bound(mvScaleFactors, ratio)` clamped to nLevels-1, all in float
    (src/ORBmatcher.cc:1660-1667, :355-360; src/Frame.cc:178-186)."""
    sf = np.asarray(scale_factors, np.float32)
    ratio = (np.asarray(dist, np.float32) / np.asarray(min_distance, np.float32)).astype(np.float32)
    return np.minimum(np.searchsorted(sf, ratio, side="left"), len(sf) - 1).astype(np.int32)


class ORBmatcher:
    """ORB_SLAM::ORBmatcher(nnratio = 0.6, checkOri = true), include/ORBmatcher.h:41.  Every Search* method takes
    frame sets and lists of frame pairs (batched); a single pair is a list of length one."""

    TH_HIGH, TH_LOW, HISTO_LENGTH = TH_HIGH, TH_LOW, HISTO_LENGTH

    def __init__(self, nnratio=0.6, checkOri=True, *, matcher: Matcher | None = None, device=0):
        self.mfNNratio = float(nnratio)
        self.mbCheckOrientation = bool(checkOri)
        self.m = matcher or Matcher(device)
        self._L = self.m._L

    def upload(self, frames):
        return self.m.upload(frames)

    # static int DescriptorDistance(const cv::Mat&, const cv::Mat&), include/ORBmatcher.h:44
    def DescriptorDistance(self, a, b):
        d = self.m.descriptor_distance(a, b)
        return int(d[0]) if len(d) == 1 else d

    def _pairs(self, idx1, idx2):
        i1, i2 = _i32(np.atleast_1d(idx1)), _i32(np.atleast_1d(idx2))
        assert i1.shape == i2.shape
        return i1, i2

    def SearchForInitialization(self, set1: FrameSet, set2: FrameSet, idx1, idx2, vbPrevMatched, windowSize=10,
                                out=None, nm=None):
        """src/ORBmatcher.cc:598-713.  vbPrevMatched: list (per pair) of [n1,2] float arrays, or one packed float32
        array [pairs, set1.max_n, 2]; updated in place either way.  `out` [pairs, set1.max_n] int32 / `nm` [pairs]
        int32 may be supplied (e.g. views of pinned memory) to avoid a fresh pageable allocation per call.
        Returns (nmatches[np], vnMatches12 rows [np, stride])."""
        i1, i2 = self._pairs(idx1, idx2)
        stride = max(set1.max_n, 1)
        packed = (isinstance(vbPrevMatched, np.ndarray) and vbPrevMatched.dtype == np.float32
                  and vbPrevMatched.shape == (len(i1), stride, 2) and vbPrevMatched.flags.c_contiguous)
        if packed:      # the ABI's own layout [pair][stride][2]: used (and updated) in place, no per-pair copies
            xy = vbPrevMatched
        else:
            xy = np.zeros((len(i1), stride, 2), np.float32)
            for p, a in enumerate(vbPrevMatched):
                xy[p, : len(a)] = a
        if out is None:
            out = np.empty((len(i1), stride), np.int32)
        if nm is None:
            nm = np.empty(len(i1), np.int32)
        assert out.shape == (len(i1), stride) and out.dtype == np.int32 and out.flags.c_contiguous
        assert nm.shape == (len(i1),) and nm.dtype == np.int32
        self.m._check(self._L.orbm_search_for_initialization(self.m._h, set1._s, set2._s, i1.ctypes.data, i2.ctypes.data,
                                                             len(i1), xy.ctypes.data, stride, int(windowSize),
                                                             self.mfNNratio, int(self.mbCheckOrientation),
                                                             out.ctypes.data, nm.ctypes.data))
        if not packed:
            for p, a in enumerate(vbPrevMatched):
                a[...] = xy[p, : len(a)]
        return nm, out

    def WindowSearch(self, set1, set2, idx1, idx2, windowSize, minScaleLevel=0, maxScaleLevel=2**31 - 1):
        """src/ORBmatcher.cc:409-516.  Returns (nmatches, rows[i2] = i1)."""
        i1, i2 = self._pairs(idx1, idx2)
        stride = max(set2.max_n, 1)
        out = np.empty((len(i1), stride), np.int32); nm = np.empty(len(i1), np.int32)
        self.m._check(self._L.orbm_window_search(self.m._h, set1._s, set2._s, i1.ctypes.data, i2.ctypes.data, len(i1),
                                                 int(windowSize), self.mfNNratio, int(self.mbCheckOrientation),
                                                 int(minScaleLevel), int(maxScaleLevel), out.ctypes.data, stride,
                                                 nm.ctypes.data))
        return nm, out

    def SearchByProjectionFrames(self, set1, set2, idx1, idx2, uv, windowSize):
        """SearchByProjection(Frame&, Frame&, windowSize, ...), src/ORBmatcher.cc:519-594; uv: per pair [n1,2]."""
        i1, i2 = self._pairs(idx1, idx2)
        stride = max(set1.max_n, set2.max_n, 1)
        xy = np.zeros((len(i1), stride, 2), np.float32)
        for p, a in enumerate(uv):
            xy[p, : len(a)] = a
        out = np.empty((len(i1), stride), np.int32); nm = np.empty(len(i1), np.int32)
        self.m._check(self._L.orbm_search_by_projection_frames(self.m._h, set1._s, set2._s, i1.ctypes.data,
                                                               i2.ctypes.data, len(i1), xy.ctypes.data, stride,
                                                               int(windowSize), self.mfNNratio, out.ctypes.data,
                                                               nm.ctypes.data))
        return nm, out

    def SearchByProjection(self, fset, idx, points, variant="local_map", th_dist=TH_HIGH):
        """Projected-point scoring: variant 'local_map' (src/ORBmatcher.cc:49-125) or 'best' (:1507-1620, :1622-1746,
        :286-407).  points: list of ProjectedPoints, one per problem.  Returns (nmatches, rows[idx] = point)."""
        ii = _i32(np.atleast_1d(idx))
        arr = (_Points * len(points))(*[p.c_struct() for p in points])
        v = {"local_map": 0, "best": 1, "independent": 2}[variant]
        # 'independent' (Fuse / SearchBySim3 loops) reports per POINT, the others per keypoint of the target frame
        stride = max([p.n for p in points] + [1]) if v == 2 else max(fset.max_n, 1)
        out = np.empty((len(ii), stride), np.int32); nm = np.empty(len(ii), np.int32)
        self.m._check(self._L.orbm_search_by_projection(self.m._h, fset._s, ii.ctypes.data, arr, len(ii), v,
                                                        int(th_dist), self.mfNNratio, int(self.mbCheckOrientation),
                                                        out.ctypes.data, stride, nm.ctypes.data))
        return nm, out

    def Fuse(self, fset, idx, points, th_dist=TH_LOW):
        """Candidate scoring of ORBmatcher::Fuse (both overloads, src/ORBmatcher.cc:1016-1134, :1136-1265): per problem
        (nFused, rows[k] = bestIdx of point k or -1).  What the reference then does with (point, bestIdx) - Replace or
        AddObservation, :1116-1131 - is pointer bookkeeping and stays with the caller."""
        return self.SearchByProjection(fset, idx, points, "independent", th_dist)

    def SearchBySim3(self, fset, idx1, idx2, points12, points21, th_dist=TH_HIGH):
        """ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1267-1505): both scans + agreement -> (nFound, match12 rows)."""
        i1, i2 = self._pairs(idx1, idx2)
        a12 = (_Points * len(points12))(*[p.c_struct() for p in points12])
        a21 = (_Points * len(points21))(*[p.c_struct() for p in points21])
        stride = max(fset.max_n, 1)
        out = np.empty((len(i1), stride), np.int32); nf = np.empty(len(i1), np.int32)
        self.m._check(self._L.orbm_search_by_sim3(self.m._h, fset._s, i1.ctypes.data, i2.ctypes.data, a12, a21, len(i1),
                                                  int(th_dist), out.ctypes.data, stride, nf.ctypes.data))
        return nf, out

    def SearchByBoW(self, set1, set2, idx1, idx2, variant="kf_f"):
        """'kf_f': SearchByBoW(KeyFrame*, Frame&, ...) src/ORBmatcher.cc:155-284, rows[idxF] = idxKF;
        'kf_kf': SearchByBoW(KeyFrame*, KeyFrame*, ...) :715-850, rows[idx1] = idx2."""
        i1, i2 = self._pairs(idx1, idx2)
        stride = max(set1.max_n, set2.max_n, 1)
        out = np.empty((len(i1), stride), np.int32); nm = np.empty(len(i1), np.int32)
        self.m._check(self._L.orbm_search_by_bow(self.m._h, set1._s, set2._s, i1.ctypes.data, i2.ctypes.data, len(i1),
                                                 0 if variant == "kf_f" else 1, self.mfNNratio,
                                                 int(self.mbCheckOrientation), out.ctypes.data, stride, nm.ctypes.data))
        return nm, out

    def SearchForTriangulation(self, set1, set2, idx1, idx2, F12, sigma2):
        """src/ORBmatcher.cc:852-1014; F12: [np,3,3] float32, sigma2: per-level.  rows[idx1] = idx2."""
        i1, i2 = self._pairs(idx1, idx2)
        F = np.ascontiguousarray(F12, np.float32).reshape(len(i1), 9)
        s2 = np.ascontiguousarray(sigma2, np.float32)
        stride = max(set1.max_n, 1)
        out = np.empty((len(i1), stride), np.int32); nm = np.empty(len(i1), np.int32)
        self.m._check(self._L.orbm_search_for_triangulation(self.m._h, set1._s, set2._s, i1.ctypes.data, i2.ctypes.data,
                                                            len(i1), F.ctypes.data, s2.ctypes.data, len(s2),
                                                            int(self.mbCheckOrientation), out.ctypes.data, stride,
                                                            nm.ctypes.data))
        return nm, out


class BowDatabase:
    """KeyFrameDatabase candidate scoring (reference src/KeyFrameDatabase.cc:75-308) over a resident keyframe set with
    BowVectors: inverted file on the device, shared-word counts, the 0.8 * max cut and the L1 scores of the survivors."""

    def __init__(self, matcher: "Matcher", keyframes: "FrameSet"):
        self.m, self.keyframes = matcher, keyframes
        p = C.c_void_p()
        matcher._check(matcher._L.orbm_bowdb_create(matcher._h, keyframes._s, C.byref(p)))
        self._d = p

    def query(self, queries: "FrameSet", qidx, exclude=None):
        """-> (common [nq, nkf], max_common [nq], min_common [nq], score [nq, nkf] (-1 below the cut / excluded))."""
        qi = _i32(np.atleast_1d(qidx))
        nq, nkf = len(qi), self.keyframes.nframes
        common = np.zeros((nq, max(nkf, 1)), np.int32); score = np.full((nq, max(nkf, 1)), -1, np.float32)
        mx = np.zeros(nq, np.int32); mn = np.zeros(nq, np.int32)
        ex = None if exclude is None else np.ascontiguousarray(exclude, np.uint8).reshape(nq, nkf)
        self.m._check(self.m._L.orbm_bowdb_query(self._d, queries._s, qi.ctypes.data, nq, None if ex is None else ex.ctypes.data,
                                                 common.ctypes.data, mx.ctypes.data, mn.ctypes.data, score.ctypes.data))
        return common[:, :nkf], mx, mn, score[:, :nkf]

    def close(self):
        if self._d:
            self.m._L.orbm_bowdb_destroy(self._d)
            self._d = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def shard_range(n_units, rank, world):
    """Contiguous shard [a, b) of n_units keyframes owned by `rank` (database sharded by keyframe)."""
    return n_units * rank // world, n_units * (rank + 1) // world


class ShardedKnn2:
    """All-pairs 2-NN with the database sharded over the ranks of one box (BASELINE config 5): every rank searches
    its shard with global indices, ONE ncclAllGather exchanges the per-shard packed top-2 keys, a merge kernel applies
    min / second-min and the accept rule.  The communicator is created through the C ABI (orbm_comm_*); the 128-byte
    NCCL id travels over whatever control channel the launcher has (torch.distributed here)."""

    def __init__(self, matcher: Matcher, rank: int, world: int, unique_id: bytes):
        self.m, self.rank, self.world = matcher, rank, world
        L = matcher._L
        L.orbm_comm_init_rank.argtypes = [C.c_void_p, C.c_int, C.c_char_p, C.c_int, C.POINTER(C.c_void_p)]
        L.orbm_comm_destroy.argtypes = [C.c_void_p]; L.orbm_comm_destroy.restype = None
        self._comm = C.c_void_p()
        matcher._check(L.orbm_comm_init_rank(matcher._h, world, unique_id, rank, C.byref(self._comm)))

    @staticmethod
    def new_unique_id() -> bytes:
        L = _bind(lib())
        buf = C.create_string_buffer(128)
        L.orbm_comm_unique_id.argtypes = [C.c_char_p]
        rc = L.orbm_comm_unique_id(buf)
        if rc != 0:
            raise OrbmError(rc, "NCCL not available in this process")
        return buf.raw

    @classmethod
    def from_torch_distributed(cls, matcher, device):
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        t = torch.zeros(128, dtype=torch.uint8, device=device)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.new_unique_id()), dtype=torch.uint8))
        dist.broadcast(t, 0)
        return cls(matcher, rank, world, bytes(t.cpu().numpy().tobytes()))

    def close(self):
        if getattr(self, "_comm", None) and getattr(self.m, "_h", None):
            self.m._L.orbm_comm_release(self.m._h, self._comm)
        self._comm = None

    def search(self, q_ptr, nq, db_shard_ptr, ndb_shard, idx_base, gathered_ptr, top2_ptr, match_ptr, th_dist=TH_LOW,
               ratio=0.75, stream=None):
        """Device pointers; asynchronous on `stream`.  gathered: [world, nq, 2] uint32 scratch."""
        self.m._check(self.m._L.orbm_knn2_sharded(self.m._h, q_ptr, nq, db_shard_ptr, ndb_shard, idx_base, self.world,
                                                  self._comm, gathered_ptr, th_dist, ratio, top2_ptr, match_ptr, stream))
