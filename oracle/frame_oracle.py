"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/frame_oracle.cc: Frame::UndistortKeyPoints / ComputeImageBounds (reference
src/Frame.cc:288-349).  Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this module.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import oracle as _O


class Camera(C.Structure):
    _fields_ = [(n, C.c_float) for n in ("fx", "fy", "cx", "cy", "k1", "k2", "p1", "p2", "k3")]


def undistort_points(cam: Camera, xy: np.ndarray) -> np.ndarray:
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    out = np.empty_like(xy)
    L = _O.lib(False)
    L.fo_undistort_points.argtypes = [C.POINTER(Camera), C.c_void_p, C.c_int, C.c_void_p]
    L.fo_undistort_points(C.byref(cam), xy.ctypes.data, len(xy), out.ctypes.data)
    return out


def image_bounds(cam: Camera, cols: int, rows: int):
    """-> ((mnMinX, mnMaxX, mnMinY, mnMaxY), (mfGridElementWidthInv, mfGridElementHeightInv))"""
    b = np.zeros(4, np.int32); inv = np.zeros(2, np.float32)
    L = _O.lib(False)
    L.fo_image_bounds.argtypes = [C.POINTER(Camera), C.c_int, C.c_int, C.c_void_p, C.c_void_p]
    L.fo_image_bounds(C.byref(cam), cols, rows, b.ctypes.data, inv.ctypes.data)
    return tuple(int(v) for v in b), (float(inv[0]), float(inv[1]))


class View(C.Structure):
    """What Frame holds after UpdatePoseMatrices (src/Frame.cc:129-134) + intrinsics, bounds, scale pyramid."""
    _fields_ = [("Rcw", C.c_float * 9), ("tcw", C.c_float * 3), ("Ow", C.c_float * 3),
                ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float), ("cy", C.c_float),
                ("min_x", C.c_int32), ("max_x", C.c_int32), ("min_y", C.c_int32), ("max_y", C.c_int32),
                ("nlevels", C.c_int32), ("scale", C.c_float * 16)]


def make_view(Rcw, tcw, fx, fy, cx, cy, bounds, scale_factors, Ow=None):
    """Ow defaults to -Rcw^T tcw evaluated as cv::gemm does for that expression (GEMM_1_T, alpha = -1: double
    accumulation in k order, one rounding to float)."""
    Rcw = np.asarray(Rcw, np.float32).reshape(3, 3); tcw = np.asarray(tcw, np.float32).reshape(3)
    if Ow is None:
        Ow = np.zeros(3, np.float32)
        for i in range(3):
            s = np.float64(0)
            for k in range(3):
                s = s + np.float64(Rcw[k, i]) * np.float64(tcw[k])
            Ow[i] = np.float32(s * -1.0)
    sf = np.zeros(16, np.float32); sf[:len(scale_factors)] = scale_factors
    return View((C.c_float * 9)(*Rcw.ravel()), (C.c_float * 3)(*tcw), (C.c_float * 3)(*np.asarray(Ow, np.float32)),
                fx, fy, cx, cy, *[int(b) for b in bounds], len(scale_factors), (C.c_float * 16)(*sf))


def is_in_frustum(view: View, pos, normal, min_dist, max_dist, cos_limit=0.5):
    """Frame::isInFrustum (src/Frame.cc:136-197) for n map points -> (in_view u8, u, v, level, view_cos)."""
    pos = np.ascontiguousarray(pos, np.float32).reshape(-1, 3); normal = np.ascontiguousarray(normal, np.float32)
    mn = np.ascontiguousarray(min_dist, np.float32); mx = np.ascontiguousarray(max_dist, np.float32)
    n = len(pos)
    iv = np.zeros(n, np.uint8); u = np.zeros(n, np.float32); v = np.zeros(n, np.float32)
    lv = np.zeros(n, np.int32); vc = np.zeros(n, np.float32)
    L = _O.lib(False)
    L.fo_is_in_frustum.argtypes = [C.POINTER(View)] + [C.c_void_p] * 4 + [C.c_int, C.c_float] + [C.c_void_p] * 5
    L.fo_is_in_frustum(C.byref(view), pos.ctypes.data, normal.ctypes.data, mn.ctypes.data, mx.ctypes.data, n,
                       cos_limit, iv.ctypes.data, u.ctypes.data, v.ctypes.data, lv.ctypes.data, vc.ctypes.data)
    return iv, u, v, lv, vc
