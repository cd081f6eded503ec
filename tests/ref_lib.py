"""ctypes binding of oracle/_ref/libref_orbextractor.so -- TEST INFRASTRUCTURE ONLY.

That library is the REFERENCE's own src/ORBextractor.cpp compiled where it lies under /root/reference
against the cv shim in oracle/ref_shim/ (oracle/Makefile). It exists only in the authoring container
(the GPU box has no /root/reference and only receives the prebuilt .so); tests that need it skip when
it is absent and fall back to the golden vectors made from it (tests/golden/ref_extract.npz,
tools/make_golden_ref.py).
"""
import ctypes as C
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
LIB_PATH = ROOT / "oracle" / "_ref" / "libref_orbextractor.so"
KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])

_lib = None


def available():
    return LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(str(LIB_PATH))
        L.ref_extractor_create.restype = C.c_void_p
        L.ref_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.ref_extractor_destroy.argtypes = [C.c_void_p]
        L.ref_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
        L.ref_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 4
        L.ref_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_get_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.ref_set_allocator_mode.argtypes = [C.c_int]
        L.ref_arena_used.restype = C.c_uint64
        _lib = L
    return _lib


def set_allocator_mode(mode):
    """0 = system allocator (octree ties broken by heap address, as shipped); 1 = monotonic arena
    (pointer order == creation order == the oracle's canonical tie-break)."""
    if lib().ref_set_allocator_mode(mode) != 0:
        raise MemoryError("could not reserve the monotonic arena")


class RefExtractor:
    """ORB_SLAM2::ORBextractor itself (R/include/ORBextractor.h:45-123)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self._h = lib().ref_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_extractor_destroy(self._h)
            self._h = None

    def __call__(self, img, cap=20000):
        img = np.ascontiguousarray(img, np.uint8)
        k = np.empty(cap, KP_DTYPE)
        d = np.empty((cap, 32), np.uint8)
        n = lib().ref_extract(self._h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0],
                              k.ctypes.data, d.ctypes.data, cap)
        if n < 0:
            return self(img, cap=-n)
        return k[:n].copy(), d[:n].copy()

    def tables(self):
        n = self.nlevels
        out = [np.empty(n, np.float32) for _ in range(4)]
        lib().ref_tables(self._h, *[o.ctypes.data for o in out])
        return dict(zip(("scale", "inv_scale", "sigma2", "inv_sigma2"), out))

    def level(self, l):
        w, h = C.c_int(), C.c_int()
        lib().ref_level_size(self._h, l, C.byref(w), C.byref(h))
        out = np.empty((h.value, w.value), np.uint8)
        lib().ref_get_level(self._h, l, out.ctypes.data, w.value)
        return out
