"""ctypes binding of oracle/_ref/libref_voc.so (the reference's own vendored DBoW2) -- TEST INFRASTRUCTURE ONLY."""
import ctypes as C
from pathlib import Path

import numpy as np

LIB_PATH = Path(__file__).resolve().parents[1] / "oracle" / "_ref" / "libref_voc.so"
P = C.c_void_p
_lib = None


def available():
    return LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(str(LIB_PATH))
        L.ref_voc_load.restype = P
        L.ref_voc_load.argtypes = [C.c_char_p]
        L.ref_voc_destroy.argtypes = [P]
        L.ref_voc_words.argtypes = [P]
        L.ref_voc_transform.argtypes = [P, P, C.c_int, C.c_int] + [P] * 7
        _lib = L
    return _lib


def flat_transform(fn, handle, desc, levelsup):
    """Runs a transform with the flattened outputs of ref_voc_transform: (words, values, nodes, off, idx)."""
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    n = len(desc)
    m = max(n, 1)
    words, values = np.empty(m, np.uint32), np.empty(m, np.float64)
    nodes, off, idx = np.empty(m, np.uint32), np.zeros(m + 1, np.int32), np.empty(m, np.uint32)
    nn, ni = C.c_int(0), C.c_int(0)
    p = lambda a: a.ctypes.data_as(P)
    k = fn(handle, p(desc), n, levelsup, p(words), p(values), p(nodes), p(off), p(idx), C.byref(nn), C.byref(ni))
    return words[:k].copy(), values[:k].copy(), nodes[:nn.value].copy(), off[:nn.value + 1].copy(), idx[:ni.value].copy()


class Vocabulary:
    def __init__(self, path):
        self._h = lib().ref_voc_load(str(path).encode())
        assert self._h, "the reference could not load the vocabulary"

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_voc_destroy(self._h)
            self._h = None

    def words(self):
        return lib().ref_voc_words(self._h)

    def transform(self, desc, levelsup):
        return flat_transform(lib().ref_voc_transform, self._h, desc, levelsup)
