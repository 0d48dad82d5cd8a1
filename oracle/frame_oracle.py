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
