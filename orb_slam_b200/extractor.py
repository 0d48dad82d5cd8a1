"""Host-side mirror of ORB_SLAM::ORBextractor (reference include/ORBextractor.h:32-77) on top of the C ABI.

Same constructor arguments, same call semantics: ``keypoints, descriptors = extractor(image)`` stands for
``(*mpORBextractor)(im, cv::Mat(), mvKeys, mDescriptors)`` (reference src/Frame.cc:59).  numpy / torch are used only to
own memory; every result comes from the CUDA kernels behind include/orbx.h.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from ._lib import OrbxParams, f32p, i32p, i64p, lib, u8p

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
STAGES = ("pyramid", "fast", "select", "blur", "describe")


class OrbxError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"orbx error {code}: {msg}")
        self.code = code


class ORBextractor:
    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20, *, max_width=640,
                 max_height=480, max_batch=256, device=0, blur_mode=0, retain_mode=0):
        self._L = lib()
        p = OrbxParams()
        self._L.orbx_default_params(C.byref(p))
        p.nfeatures, p.scale_factor, p.nlevels, p.score_type, p.fast_th = nfeatures, scaleFactor, nlevels, scoreType, fastTh
        p.max_width, p.max_height, p.max_batch, p.device = max_width, max_height, max_batch, device
        p.blur_mode, p.retain_mode = blur_mode, retain_mode
        self.params = p
        self.nfeatures = nfeatures
        self._h = C.c_void_p()
        rc = self._L.orbx_create(C.byref(p), C.byref(self._h))
        if rc != 0:
            self._h = None
            raise OrbxError(rc, (self._L.orbx_last_error(None) or b"").decode())

    @classmethod
    def from_settings(cls, path, **kw):
        """Reads ORBextractor.nFeatures/scaleFactor/nLevels/fastTh/nScoreType (reference src/Tracking.cc:105-113)."""
        L = lib()
        p = OrbxParams()
        L.orbx_default_params(C.byref(p))
        rc = L.orbx_params_from_yaml(str(path).encode(), C.byref(p))
        if rc != 0:
            raise OrbxError(rc, (L.orbx_last_error(None) or b"").decode())
        return cls(p.nfeatures, p.scale_factor, p.nlevels, p.score_type, p.fast_th, **kw)

    def close(self):
        if getattr(self, "_h", None):
            self._L.orbx_destroy(self._h)
            self._h = None

    __del__ = close

    def _check(self, rc):
        if rc != 0:
            raise OrbxError(rc, (self._L.orbx_last_error(self._h) or b"").decode())

    def GetLevels(self):
        return self._L.orbx_get_levels(self._h)

    def GetScaleFactor(self):
        return self._L.orbx_get_scale_factor(self._h)

    # -- operator() -------------------------------------------------------------------------------------------
    def __call__(self, image: np.ndarray, mask=None):
        """image: HxW uint8 (host).  Returns (keypoints[KP_DTYPE], descriptors[N,32] uint8)."""
        if image is None or image.size == 0:
            return np.zeros(0, KP_DTYPE), np.zeros((0, 32), np.uint8)
        if image.dtype != np.uint8 or image.ndim != 2:
            raise TypeError("image must be CV_8UC1 (reference asserts, src/ORBextractor.cc:725)")
        if image.strides[1] != 1:
            image = np.ascontiguousarray(image)
        kps = np.zeros(self.nfeatures, KP_DTYPE)
        desc = np.zeros((self.nfeatures, 32), np.uint8)
        n = C.c_int32(0)
        self._check(self._L.orbx_extract(self._h, image.ctypes.data, image.shape[1], image.shape[0], image.strides[0],
                                         kps.ctypes.data, desc.ctypes.data, C.byref(n)))
        return kps[: n.value].copy(), desc[: n.value].copy()

    def extract_batch(self, images: np.ndarray, out=None):
        """images: [F,H,W] uint8 host array (pinned memory makes the copies asynchronous).
        Returns (kps[F,nfeatures], desc[F,nfeatures,32], n[F])."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.strides[2] == 1
        F, H, W = images.shape
        if out is None:
            out = (np.zeros((F, self.nfeatures), KP_DTYPE), np.zeros((F, self.nfeatures, 32), np.uint8),
                   np.zeros(F, np.int32))
        kps, desc, n = out
        self._check(self._L.orbx_extract_batch(self._h, images.ctypes.data, F, W, H, images.strides[1],
                                               images.strides[0], kps.ctypes.data, desc.ctypes.data, n.ctypes.data))
        return kps, desc, n

    def extract_batch_ptr(self, img_ptr, F, W, H, stride, frame_stride, kps_ptr, desc_ptr, n_ptr):
        """Raw host pointers (e.g. pinned torch tensors): the e2e call bench.py times."""
        self._check(self._L.orbx_extract_batch(self._h, img_ptr, F, W, H, stride, frame_stride, kps_ptr, desc_ptr, n_ptr))

    def extract_batch_device(self, img_ptr, F, W, H, stride, frame_stride, kps_ptr, desc_ptr, n_ptr, stream=None):
        """Device pointers; asynchronous on `stream` (cudaStream_t as int; None = the handle's stream)."""
        self._check(self._L.orbx_extract_batch_device(self._h, img_ptr, F, W, H, stride, frame_stride, kps_ptr,
                                                      desc_ptr, n_ptr, stream))

    def synchronize(self):
        self._check(self._L.orbx_synchronize(self._h))

    # -- measurement ------------------------------------------------------------------------------------------
    def profile_enable(self, on=True):
        self._check(self._L.orbx_profile_enable(self._h, int(on)))

    def profile_read(self):
        ms = np.zeros(len(STAGES), np.float32)
        ln = np.zeros(len(STAGES), np.int64)
        self._check(self._L.orbx_profile_read(self._h, ms.ctypes.data_as(f32p), ln.ctypes.data_as(i64p)))
        return {s: (float(ms[i]), int(ln[i])) for i, s in enumerate(STAGES)}

    def kernel_launches(self):
        return int(self._L.orbx_kernel_launches(self._h))

    # -- parity introspection (tests) ---------------------------------------------------------------------------
    def debug_level_info(self, level):
        v = np.zeros(8, np.int32)
        self._check(self._L.orbx_debug_level_info(self._h, level, v.ctypes.data_as(i32p)))
        return dict(zip(("w", "h", "cols", "rows", "cellW", "cellH", "quota", "nDesired"), map(int, v)))

    def debug_level_image(self, frame, level, which=0):
        i = self.debug_level_info(level)
        out = np.empty((i["h"], i["w"]), np.uint8)
        self._check(self._L.orbx_debug_level_image(self._h, frame, level, which, out.ctypes.data_as(u8p)))
        return out

    def debug_retain_best(self, lists, keep, retain_mode=0, batch=0):
        """retainBest(list, keep) + resize through the device port of std::nth_element; lists: response arrays.
        Returns per list the original positions of the records in their new order (all of them)."""
        start = np.concatenate([[0], np.cumsum([len(a) for a in lists])]).astype(np.int32)
        resp = np.ascontiguousarray(np.concatenate([np.asarray(a, np.float32) for a in lists]) if lists else np.zeros(0, np.float32))
        kp = np.ascontiguousarray(keep, np.int32)
        order = np.zeros(max(len(resp), 1), np.int32)
        self._check(self._L.orbx_debug_retain_best(self._h, resp.ctypes.data, start.ctypes.data, kp.ctypes.data,
                                                   len(lists), retain_mode, batch, order.ctypes.data))
        return [order[start[i]:start[i + 1]].copy() for i in range(len(lists))]

    def debug_level_cells(self, frame, level):
        i = self.debug_level_info(level)
        a = np.zeros(i["cols"] * i["rows"], np.int32)
        b = np.zeros_like(a)
        self._check(self._L.orbx_debug_level_cells(self._h, frame, level, a.ctypes.data_as(i32p), b.ctypes.data_as(i32p)))
        return a, b

    def debug_cell_candidates(self, frame, level, cell, cap=16384):
        xs = np.zeros(cap, np.int32); ys = np.zeros(cap, np.int32); r = np.zeros(cap, np.float32)
        n = C.c_int32(0)
        self._check(self._L.orbx_debug_cell_candidates(self._h, frame, level, cell, xs.ctypes.data_as(i32p),
                                                       ys.ctypes.data_as(i32p), r.ctypes.data_as(f32p), cap, C.byref(n)))
        return xs[: n.value].copy(), ys[: n.value].copy(), r[: n.value].copy()


def synth_frames_device(ptr, nframes, width, height, stride, frame_stride, seed0=0, stream=None):
    """Fill a DEVICE buffer with synthetic frames (seed = seed0 + frame index); CPU twin: oracle/synth.cc."""
    rc = lib().orbx_synth_frames_device(ptr, nframes, width, height, stride, frame_stride, seed0, stream)
    if rc != 0:
        raise OrbxError(rc, "orbx_synth_frames_device failed")
