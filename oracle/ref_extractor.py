"""ORACLE — TEST INFRASTRUCTURE ONLY.

The reference's OWN extractor: /root/reference/src/ORBextractor.cc compiled in place (oracle/Makefile target `ref`,
oracle/ref_orbextractor.cc, type-only OpenCV stand-in oracle/ref_shim_cv/) with its five OpenCV library calls bridged
to the real OpenCV installed here (cv2): cv::resize, cv::copyMakeBorder, cv::GaussianBlur, cv::FAST, cv::fastAtan2.
`RefExtractor` mirrors ORB_SLAM::ORBextractor (include/ORBextractor.h:32-77).  Needs oracle/_ref/libref_orbextractor.so
(built where /root/reference exists) and an importable cv2; tests that use it skip otherwise.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libref_orbextractor.so")

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"),
                     ("octave", "<i4"), ("class_id", "<i4")])

_u8p = C.POINTER(C.c_uint8)
_RESIZE = C.CFUNCTYPE(None, _u8p, C.c_int, C.c_int, C.c_long, _u8p, C.c_int, C.c_int, C.c_long, C.c_int)
_BORDER = C.CFUNCTYPE(None, _u8p, C.c_int, C.c_int, C.c_long, _u8p, C.c_long, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int)
_BLUR = C.CFUNCTYPE(None, _u8p, C.c_int, C.c_int, C.c_long, C.c_int, C.c_double, C.c_int)
_FAST = C.CFUNCTYPE(C.c_int, _u8p, C.c_int, C.c_int, C.c_long, C.c_int, C.c_int, C.POINTER(C.c_float), C.c_int)
_ATAN = C.CFUNCTYPE(C.c_float, C.c_float, C.c_float)
_UNDIST = C.CFUNCTYPE(None, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_float), C.c_int)


def available() -> bool:
    if not os.path.exists(REF_LIB):
        return False
    try:
        import cv2  # noqa: F401
    except Exception:
        return False
    return True


def _view(ptr, w, h, step):
    """numpy view of an 8-bit image the C++ side owns (rows `step` bytes apart)."""
    n = (h - 1) * step + w
    flat = np.ctypeslib.as_array(ptr, shape=(n,))
    return np.lib.stride_tricks.as_strided(flat, shape=(h, w), strides=(step, 1))


_state = {}


def _lib(path=None):
    """Loads a library that contains ref_orbextractor.cc and points its OpenCV bridge at cv2."""
    key = path or REF_LIB
    if key in _state:
        return _state[key]["L"]
    import cv2
    cv2.setNumThreads(1)
    detectors = {}

    def resize(src, sw, sh, sstep, dst, dw, dh, dstep, interp):
        out = cv2.resize(np.ascontiguousarray(_view(src, sw, sh, sstep)), (dw, dh), interpolation=interp)
        _view(dst, dw, dh, dstep)[:] = out

    def border(src, w, h, sstep, dst, dstep, top, bottom, left, right, btype):
        # the source may be a view INTO the destination (ComputePyramid :806): copy it out first.  BORDER_ISOLATED is
        # implied: the array handed to cv2 has no parent
        s = np.ascontiguousarray(_view(src, w, h, sstep))
        out = cv2.copyMakeBorder(s, top, bottom, left, right, btype & ~16)
        _view(dst, w + left + right, h + top + bottom, dstep)[:] = out

    def blur(img, w, h, step, ksize, sigma, btype):
        # in place on a view whose parent holds the reflect-101 border the reference built (:806-816); the filter
        # would read those parent pixels, which equal what BORDER_REFLECT_101 synthesises for the isolated view
        v = _view(img, w, h, step)
        v[:] = cv2.GaussianBlur(np.ascontiguousarray(v), (ksize, ksize), sigma, sigmaY=sigma, borderType=btype)

    def fast(img, w, h, step, th, nonmax, xyr, cap):
        key = (th, nonmax)
        if key not in detectors:
            detectors[key] = cv2.FastFeatureDetector_create(th, bool(nonmax), cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
        kps = detectors[key].detect(np.ascontiguousarray(_view(img, w, h, step)))
        n = min(len(kps), cap)
        out = np.ctypeslib.as_array(xyr, shape=(cap * 3,))
        for i in range(n):
            k = kps[i]
            out[3 * i] = k.pt[0]; out[3 * i + 1] = k.pt[1]; out[3 * i + 2] = k.response
        return n

    def atan2(y, x):
        return cv2.fastAtan2(y, x)

    def undistort(xy, n, K9, D, nd):
        pts = np.ctypeslib.as_array(xy, shape=(n, 2))
        K = np.ctypeslib.as_array(K9, shape=(9,)).reshape(3, 3).copy()
        Dv = np.ctypeslib.as_array(D, shape=(nd,)).reshape(nd, 1).copy()
        pts[:] = cv2.undistortPoints(pts.reshape(-1, 1, 2).copy(), K, Dv, None, K).reshape(-1, 2)

    L = C.CDLL(key)
    cbs = (_RESIZE(resize), _BORDER(border), _BLUR(blur), _FAST(fast), _ATAN(atan2))
    und = _UNDIST(undistort)
    L.ref_orb_set_undistort.argtypes = [_UNDIST]
    L.ref_orb_set_undistort(und)
    L.ref_orb_set_bridge.argtypes = [_RESIZE, _BORDER, _BLUR, _FAST, _ATAN, C.c_int]
    L.ref_orb_create.restype = C.c_void_p
    L.ref_orb_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.ref_orb_destroy.argtypes = [C.c_void_p]
    L.ref_orb_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_long, C.c_void_p, C.c_void_p, C.c_int]
    _state[key] = dict(L=L, cbs=cbs, und=und, cv2_version=cv2.__version__)
    set_retain_mode(1, key)
    return L


def set_retain_mode(mode: int, path=None):
    """KeyPointsFilter::retainBest variant: 0 = OpenCV 2.4 (nth_element at begin + n), 1 = OpenCV >= 3.4 (begin + n - 1,
    what the installed cv2 4.13 itself does)."""
    for key, st in _state.items():
        if path is None or key == path:
            st["L"].ref_orb_set_bridge(*st["cbs"], mode)


class RefExtractor:
    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20, lib_path=None):
        self._L = _lib(lib_path)
        self._h = self._L.ref_orb_create(nfeatures, scaleFactor, nlevels, scoreType, fastTh)

    def __call__(self, image):
        img = np.ascontiguousarray(image, np.uint8)
        h, w = img.shape
        cap = 1 << 16
        kps = np.zeros(cap, KP_DTYPE); desc = np.zeros((cap, 32), np.uint8)
        n = self._L.ref_orb_extract(self._h, img.ctypes.data, w, h, w, kps.ctypes.data, desc.ctypes.data, cap)
        if n < 0:
            raise RuntimeError("keypoint capacity too small")
        return kps[:n].copy(), desc[:n].copy()

    def __del__(self):
        try:
            self._L.ref_orb_destroy(self._h)
        except Exception:
            pass


# ---- the same compiled reference with its OpenCV routines served natively (oracle/ref_native_bridge.cc) -------------
NATIVE_LIB = os.path.join(_HERE, "_ref", "libref_orbextractor_native.so")
_native = {}


def native_available() -> bool:
    return os.path.exists(NATIVE_LIB)


def _native_lib():
    if "L" not in _native:
        L = C.CDLL(NATIVE_LIB)
        L.ref_orb_use_native_bridge.argtypes = [C.c_int, C.c_int]
        L.ref_orb_extract_many.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                                           C.c_int, C.c_long, C.c_long, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int,
                                           C.c_int]
        _native["L"] = L
    return _native["L"]


class NativeRefExtractor:
    """ORB_SLAM::ORBextractor (the reference's own src/ORBextractor.cc, -O3) with cv::resize / copyMakeBorder /
    GaussianBlur / FAST / fastAtan2 served by the oracle's cv2-pinned native restatements: no Python in the loop, no
    cv2 needed.  The timed CPU baseline of bench.py (`kind: "reference"`).  retain_mode / blur_mode: 0 = OpenCV 2.4
    behaviour, 1 = OpenCV >= 3.4 (what the cv2-bridged build sees)."""

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20, retain_mode=0, blur_mode=0):
        self._L = _native_lib()
        self.args = (nfeatures, scaleFactor, nlevels, scoreType, fastTh)
        self.modes = (retain_mode, blur_mode)
        self.nfeatures = nfeatures

    def extract_many(self, images: np.ndarray, nthreads: int = 1, cap: int | None = None):
        """images [F,H,W] uint8 -> (kps [F,cap], desc [F,cap,32], n [F])."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.strides[2] == 1
        self._L.ref_orb_use_native_bridge(*self.modes)      # process-wide bridge state: set per call
        F, H, W = images.shape
        cap = cap or self.nfeatures + 64
        kps = np.zeros((F, cap), KP_DTYPE); desc = np.zeros((F, cap, 32), np.uint8); n = np.zeros(F, np.int32)
        rc = self._L.ref_orb_extract_many(*self.args, images.ctypes.data, F, W, H, images.strides[1], images.strides[0],
                                          kps.ctypes.data, desc.ctypes.data, n.ctypes.data, cap, nthreads)
        if rc != 0:
            raise RuntimeError("keypoint capacity too small")
        return kps, desc, n

    def __call__(self, image):
        k, d, n = self.extract_many(np.ascontiguousarray(image, np.uint8)[None], 1)
        return k[0, :n[0]].copy(), d[0, :n[0]].copy()
