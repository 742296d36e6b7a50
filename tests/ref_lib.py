"""ctypes bindings for oracle/_ref/libref_orbextractor.so — the UNMODIFIED reference
src/ORBextractor.cc compiled over oracle/ref_shim (TEST INFRASTRUCTURE ONLY).  The library exists
only where oracle/Makefile could see /root/reference (or where a prebuilt copy travelled)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

import oracle_lib as O

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libref_orbextractor.so")
_LIB = None


def available() -> bool:
    return os.path.exists(REF_SO)


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(REF_SO)
        L.ref_extractor_create.restype = C.c_void_p
        L.ref_extractor_create.argtypes = [C.POINTER(O.Params)]
        L.ref_extractor_destroy.argtypes = [C.c_void_p]
        L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_size_t, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_get_level.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.POINTER(C.c_int),
                                    C.POINTER(C.c_int)]
        L.ref_get_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.ref_extract_match_many.restype = C.c_int64
        _LIB = L
    return _LIB


class RefExtractor:
    """ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:51) through the shim build."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7):
        self.params = O.Params(nfeatures, scale, nlevels, ini_th, min_th)
        self.h = C.c_void_p(lib().ref_extractor_create(C.byref(self.params)))
        self.nlevels = nlevels
        self.cap = nfeatures + 64 * nlevels + 64

    def __del__(self):
        if getattr(self, "h", None):
            lib().ref_extractor_destroy(self.h)
            self.h = None

    def extract(self, img: np.ndarray):
        img = np.ascontiguousarray(img)
        kps = np.zeros(self.cap, O.KP_DTYPE)
        desc = np.zeros((self.cap, 32), np.uint8)
        n = lib().ref_extract(self.h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0],
                              kps.ctypes.data, desc.ctypes.data, self.cap)
        assert n >= 0
        assert lib().ref_arena_overflowed() == 0
        return kps[:n].copy(), desc[:n].copy()

    def level(self, level: int, border: bool = False) -> np.ndarray:
        w, h = C.c_int(), C.c_int()
        assert lib().ref_get_level(self.h, level, int(border), None, 0, C.byref(w), C.byref(h)) == 0
        out = np.empty((h.value, w.value), np.uint8)
        assert lib().ref_get_level(self.h, level, int(border), out.ctypes.data, out.strides[0], None, None) == 0
        return out

    def tables(self):
        n = self.nlevels
        a, b, c, d = (np.empty(n, np.float32) for _ in range(4))
        lib().ref_get_tables(self.h, a.ctypes.data, b.ctypes.data, c.ctypes.data, d.ctypes.data)
        return dict(scale=a, inv_scale=b, sigma2=c, inv_sigma2=d)
