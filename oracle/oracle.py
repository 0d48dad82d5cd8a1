"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/liborb_oracle.so (CPU restatement of the reference's ORBextractor / ORBmatcher hot
path, see orb_oracle.cc / match_oracle.cc).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
``--impl reference`` leg may import this module; the product package orb_slam_b200/ never does.
"""
from __future__ import annotations

import ctypes as C
import hashlib
import os
import platform
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

KP_DTYPE = np.dtype(
    [("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"), ("response", "<f4"), ("octave", "<i4"),
     ("class_id", "<i4")]
)
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)


def _ptr(a, t):
    return a.ctypes.data_as(t)


def build(fast: bool = False, force: bool = False) -> str:
    """Compile the oracle with the recipe in oracle/Makefile.  The -march=native variant is rebuilt whenever the
    host CPU differs from the one it was built on (the .so travels from the build container to the GPU box)."""
    name = "liborb_oracle_fast.so" if fast else "liborb_oracle.so"
    path = os.path.join(_HERE, name)
    stamp = path + ".host"
    sig = ""
    if fast:
        try:
            flags = [l for l in open("/proc/cpuinfo") if l.startswith("flags")][0]
        except Exception:
            flags = platform.processor()
        sig = hashlib.sha1(flags.encode()).hexdigest()
        if os.path.exists(path) and (not os.path.exists(stamp) or open(stamp).read() != sig):
            force = True
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".cc") and not f.startswith("ref_")]
    srcs.append(os.path.join(_HERE, "..", "orb_slam_b200", "csrc", "orb_pattern_31.inc"))
    stale = (not os.path.exists(path)) or any(os.path.getmtime(s) > os.path.getmtime(path) for s in srcs)
    if force or stale:
        if fast and os.path.exists(path):
            os.remove(path)
        subprocess.run(["make", "-C", _HERE, name], check=True, capture_output=True)
        if fast:
            open(stamp, "w").write(sig)
    return path


def build_ref(reference: str = "/root/reference") -> str | None:
    """oracle/_ref/libref_dbow2.so + libref_orbextractor.so: the reference's own DBoW2 and src/ORBextractor.cc compiled
    where they lie (Makefile target `ref`).  Only
    possible where the reference tree exists (the authoring container); elsewhere the prebuilt file is used as is."""
    out = os.path.join(_HERE, "_ref", "libref_dbow2.so")
    if os.path.isdir(os.path.join(reference, "Thirdparty", "DBoW2")):
        subprocess.run(["make", "-C", _HERE, "ref", f"REFERENCE={reference}"], check=True, capture_output=True)
    return out if os.path.exists(out) and os.path.exists(os.path.join(_HERE, "_ref", "libref_orbextractor.so")) else None


_libs = {}


def lib(fast: bool = False):
    if fast in _libs:
        return _libs[fast]
    L = C.CDLL(build(fast))
    L.orb_oracle_create.restype = C.c_void_p
    L.orb_oracle_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]
    L.orb_oracle_destroy.argtypes = [C.c_void_p]
    L.orb_oracle_keep_dump.argtypes = [C.c_void_p, C.c_int]
    L.orb_oracle_extract.restype = C.c_int
    L.orb_oracle_extract.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_long, C.c_void_p, _u8p, C.c_int]
    L.orb_oracle_tables.argtypes = [C.c_void_p, _i32p, _f32p, _f32p, _i32p]
    L.orb_oracle_level_size.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p]
    L.orb_oracle_level_image.argtypes = [C.c_void_p, C.c_int, C.c_int, _u8p]
    L.orb_oracle_level_grid.argtypes = [C.c_void_p, C.c_int, _i32p]
    L.orb_oracle_level_cells.argtypes = [C.c_void_p, C.c_int, _i32p, _i32p]
    L.orb_oracle_cell_keys.restype = C.c_int
    L.orb_oracle_cell_keys.argtypes = [C.c_void_p, C.c_int, C.c_int, _i32p, _i32p, _f32p, C.c_int]
    L.orb_oracle_level_kept.restype = C.c_int
    L.orb_oracle_level_kept.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    L.orb_oracle_resize.argtypes = [_u8p, C.c_int, C.c_int, C.c_long, _u8p, C.c_int, C.c_int, C.c_long]
    L.orb_oracle_fast.restype = C.c_int
    L.orb_oracle_fast.argtypes = [_u8p, C.c_int, C.c_int, C.c_long, C.c_int, _i32p, _i32p, _i32p, C.c_int]
    L.orb_oracle_fast_score_map.argtypes = [_u8p, C.c_int, C.c_int, C.c_long, _u8p]
    L.orb_oracle_fast_atan2.restype = C.c_float
    L.orb_oracle_fast_atan2.argtypes = [C.c_float, C.c_float]
    L.orb_oracle_blur7.argtypes = [_u8p, C.c_int, C.c_int, C.c_long, _u8p, C.c_int]
    L.orb_oracle_retain_best.restype = C.c_int
    L.orb_oracle_retain_best.argtypes = [_f32p, _i32p, C.c_int, C.c_int, C.c_int]
    L.orb_oracle_sincos.argtypes = [C.c_float, _f32p, _f32p]
    L.orb_oracle_pattern.restype = C.POINTER(C.c_int8)
    L.orb_oracle_synth.argtypes = [_u8p, C.c_int, C.c_int, C.c_long, C.c_uint64]
    if hasattr(L, "orb_oracle_extract_many"):
        L.orb_oracle_extract_many.restype = C.c_int
        L.orb_oracle_extract_many.argtypes = [C.c_void_p, _u8p, C.c_int, C.c_int, C.c_int, C.c_long, C.c_long,
                                              C.c_void_p, _u8p, _i32p, C.c_int, C.c_int]
    _libs[fast] = L
    return L


# ----------------------------------------------------------------------------------------------------------------
# primitives
# ----------------------------------------------------------------------------------------------------------------
def synth(h: int, w: int, seed: int) -> np.ndarray:
    out = np.empty((h, w), np.uint8)
    lib().orb_oracle_synth(_ptr(out, _u8p), w, h, w, seed)
    return out


def resize(src: np.ndarray, dw: int, dh: int) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orb_oracle_resize(_ptr(src, _u8p), src.shape[1], src.shape[0], src.strides[0], _ptr(dst, _u8p), dw, dh, dw)
    return dst


def fast(img: np.ndarray, threshold: int):
    """FAST-9/16 + NMS on a (sub-)image; returns (xs, ys, scores) in OpenCV's row-major output order."""
    assert img.strides[1] == 1
    cap = img.shape[0] * img.shape[1] // 4 + 16
    xs = np.empty(cap, np.int32); ys = np.empty(cap, np.int32); sc = np.empty(cap, np.int32)
    n = lib().orb_oracle_fast(_ptr(img, _u8p), img.shape[1], img.shape[0], img.strides[0], threshold,
                              _ptr(xs, _i32p), _ptr(ys, _i32p), _ptr(sc, _i32p), cap)
    assert n <= cap
    return xs[:n].copy(), ys[:n].copy(), sc[:n].copy()


def fast_score_map(img: np.ndarray) -> np.ndarray:
    img = np.ascontiguousarray(img)
    out = np.empty(img.shape, np.uint8)
    lib().orb_oracle_fast_score_map(_ptr(img, _u8p), img.shape[1], img.shape[0], img.strides[0], _ptr(out, _u8p))
    return out


def fast_atan2(y: float, x: float) -> float:
    return float(lib().orb_oracle_fast_atan2(y, x))


def blur7(img: np.ndarray, mode: int = 0) -> np.ndarray:
    img = np.ascontiguousarray(img)
    out = np.empty(img.shape, np.uint8)
    lib().orb_oracle_blur7(_ptr(img, _u8p), img.shape[1], img.shape[0], img.strides[0], _ptr(out, _u8p), mode)
    return out


def retain_best(resp: np.ndarray, n_points: int, retain_mode: int = 0):
    """KeyPointsFilter::retainBest(...) + resize(n_points) on responses; returns (responses, original indices)."""
    r = np.ascontiguousarray(resp, np.float32).copy()
    tag = np.arange(len(r), dtype=np.int32)
    n = lib().orb_oracle_retain_best(_ptr(r, _f32p), _ptr(tag, _i32p), len(r), n_points, retain_mode)
    return r[:n], tag[:n]


def nth_element_killer(n: int, nth: int):
    """Responses on which libstdc++'s nth_element (response-greater) degenerates to its depth limit -> (resp, comparisons)."""
    L = lib()
    L.orb_oracle_nth_element_killer.restype = C.c_long
    L.orb_oracle_nth_element_killer.argtypes = [C.c_int, C.c_int, _f32p]
    r = np.zeros(n, np.float32)
    cnt = L.orb_oracle_nth_element_killer(n, nth, _ptr(r, _f32p))
    return r, int(cnt)


def sincos(angle_rad: float):
    c = C.c_float(); s = C.c_float()
    lib().orb_oracle_sincos(angle_rad, C.byref(c), C.byref(s))
    return c.value, s.value


def pattern() -> np.ndarray:
    return np.ctypeslib.as_array(lib().orb_oracle_pattern(), shape=(1024,)).copy()


# ----------------------------------------------------------------------------------------------------------------
# extractor
# ----------------------------------------------------------------------------------------------------------------
class OracleExtractor:
    """Mirror of ORB_SLAM::ORBextractor (include/ORBextractor.h:32-77) on the CPU restatement."""

    HARRIS_SCORE, FAST_SCORE = 0, 1

    def __init__(self, nfeatures=1000, scaleFactor=1.2, nlevels=8, scoreType=1, fastTh=20, blur_mode=0,
                 retain_mode=0, fast_build=False):
        self._L = lib(fast_build)
        self.nfeatures, self.nlevels = nfeatures, nlevels
        self._h = self._L.orb_oracle_create(nfeatures, scaleFactor, nlevels, scoreType, fastTh, blur_mode, retain_mode)

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.orb_oracle_destroy(self._h)
            self._h = None

    def keep_dump(self, on: bool):
        self._L.orb_oracle_keep_dump(self._h, int(on))

    def __call__(self, image: np.ndarray):
        assert image.dtype == np.uint8 and image.ndim == 2 and image.strides[1] == 1
        cap = max(self.nfeatures, 1) + 64
        kps = np.zeros(cap, KP_DTYPE)
        desc = np.zeros((cap, 32), np.uint8)
        n = self._L.orb_oracle_extract(self._h, _ptr(image, _u8p), image.shape[1], image.shape[0], image.strides[0],
                                       kps.ctypes.data, _ptr(desc, _u8p), cap)
        if n < 0:
            raise RuntimeError(f"oracle: degenerate configuration (code {n}); the reference would assert/crash")
        return kps[:n].copy(), desc[:n].copy()

    def extract_many(self, images: np.ndarray, nthreads: int = 1):
        """images [F,H,W] uint8 -> (kps [F,cap], desc [F,cap,32], n [F]); `nthreads` host threads."""
        assert images.dtype == np.uint8 and images.ndim == 3 and images.strides[2] == 1
        F, H, W = images.shape
        cap = self.nfeatures
        kps = np.zeros((F, cap), KP_DTYPE); desc = np.zeros((F, cap, 32), np.uint8); n = np.zeros(F, np.int32)
        rc = self._L.orb_oracle_extract_many(self._h, _ptr(images, _u8p), F, W, H, images.strides[1],
                                             images.strides[0], kps.ctypes.data, _ptr(desc, _u8p), _ptr(n, _i32p),
                                             cap, nthreads)
        if rc != 0:
            raise RuntimeError("oracle: degenerate configuration")
        return kps, desc, n

    def tables(self):
        f = np.zeros(self.nlevels, np.int32); s = np.zeros(self.nlevels, np.float32)
        i = np.zeros(self.nlevels, np.float32); u = np.zeros(16, np.int32)
        self._L.orb_oracle_tables(self._h, _ptr(f, _i32p), _ptr(s, _f32p), _ptr(i, _f32p), _ptr(u, _i32p))
        return f, s, i, u

    def level_size(self, level):
        w = C.c_int32(); h = C.c_int32()
        self._L.orb_oracle_level_size(self._h, level, C.byref(w), C.byref(h))
        return w.value, h.value

    def level_image(self, level, which=0):
        w, h = self.level_size(level)
        out = np.empty((h, w), np.uint8)
        self._L.orb_oracle_level_image(self._h, level, which, _ptr(out, _u8p))
        return out

    def level_grid(self, level):
        g = np.zeros(6, np.int32)
        self._L.orb_oracle_level_grid(self._h, level, _ptr(g, _i32p))
        return dict(cols=int(g[0]), rows=int(g[1]), cellW=int(g[2]), cellH=int(g[3]), quota=int(g[4]), nkept=int(g[5]))

    def level_cells(self, level):
        g = self.level_grid(level)
        n = g["cols"] * g["rows"]
        a = np.zeros(n, np.int32); b = np.zeros(n, np.int32)
        self._L.orb_oracle_level_cells(self._h, level, _ptr(a, _i32p), _ptr(b, _i32p))
        return a, b

    def cell_keys(self, level, cell, cap=8192):
        xs = np.zeros(cap, np.int32); ys = np.zeros(cap, np.int32); r = np.zeros(cap, np.float32)
        n = self._L.orb_oracle_cell_keys(self._h, level, cell, _ptr(xs, _i32p), _ptr(ys, _i32p), _ptr(r, _f32p), cap)
        assert n <= cap
        return xs[:n].copy(), ys[:n].copy(), r[:n].copy()

    def level_kept(self, level):
        kps = np.zeros(4 * self.nfeatures + 64, KP_DTYPE)
        n = self._L.orb_oracle_level_kept(self._h, level, kps.ctypes.data, len(kps))
        return kps[:n].copy()
