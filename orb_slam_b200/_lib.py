"""ctypes binding of liborbx.so.  There is deliberately no fallback: if the CUDA library is missing or does not
load, importing the product API raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liborbx.so")

u8p = C.POINTER(C.c_uint8)
i32p = C.POINTER(C.c_int32)
f32p = C.POINTER(C.c_float)
i64p = C.POINTER(C.c_int64)


class OrbxParams(C.Structure):
    _fields_ = [("nfeatures", C.c_int32), ("scale_factor", C.c_float), ("nlevels", C.c_int32),
                ("score_type", C.c_int32), ("fast_th", C.c_int32), ("max_width", C.c_int32),
                ("max_height", C.c_int32), ("max_batch", C.c_int32), ("device", C.c_int32),
                ("blur_mode", C.c_int32), ("retain_mode", C.c_int32)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing: run `python -m orb_slam_b200.build` (no CPU fallback exists)")
    L = C.CDLL(LIB_PATH)
    vp = C.c_void_p
    L.orbx_default_params.argtypes = [C.POINTER(OrbxParams)]
    L.orbx_params_from_yaml.argtypes = [C.c_char_p, C.POINTER(OrbxParams)]
    L.orbx_create.argtypes = [C.POINTER(OrbxParams), C.POINTER(vp)]
    L.orbx_destroy.argtypes = [vp]
    L.orbx_destroy.restype = None
    L.orbx_get_levels.argtypes = [vp]
    L.orbx_get_scale_factor.argtypes = [vp]
    L.orbx_get_scale_factor.restype = C.c_float
    L.orbx_last_error.argtypes = [vp]
    L.orbx_last_error.restype = C.c_char_p
    L.orbx_extract.argtypes = [vp, vp, C.c_int, C.c_int, C.c_size_t, vp, vp, i32p]
    L.orbx_extract_batch.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, vp, vp, vp]
    L.orbx_extract_batch_device.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, vp, vp, vp, vp]
    L.orbx_synchronize.argtypes = [vp]
    L.orbx_profile_enable.argtypes = [vp, C.c_int]
    L.orbx_profile_read.argtypes = [vp, f32p, i64p]
    L.orbx_kernel_launches.argtypes = [vp]
    L.orbx_kernel_launches.restype = C.c_int64
    L.orbx_debug_level_info.argtypes = [vp, C.c_int, i32p]
    L.orbx_debug_level_image.argtypes = [vp, C.c_int, C.c_int, C.c_int, u8p]
    L.orbx_debug_cell_candidates.argtypes = [vp, C.c_int, C.c_int, C.c_int, i32p, i32p, f32p, C.c_int, i32p]
    L.orbx_debug_level_cells.argtypes = [vp, C.c_int, C.c_int, i32p, i32p]
    L.orbx_debug_retain_best.argtypes = [vp, vp, vp, vp, C.c_int, C.c_int, C.c_int, vp]
    L.orbx_host_alloc.argtypes = [C.c_size_t, C.c_int, C.POINTER(vp)]
    L.orbx_host_free.argtypes = [vp]
    L.orbx_host_free.restype = None
    L.orbx_synth_frames_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_size_t, C.c_size_t, C.c_uint64, vp]
    _lib = L
    return L
