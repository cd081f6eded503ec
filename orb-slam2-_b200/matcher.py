"""Python mirror of ORB_SLAM2::ORBmatcher's Hamming searches (R/include/ORBmatcher.h:37-143) over
the C ABI. Compute is CUDA only (liborbx.so)."""
import ctypes as C

import numpy as np

from ._capi import check, lib, ptr


class ORBmatcher:
    TH_LOW = 50          # R/src/ORBmatcher.cpp:37-39
    TH_HIGH = 100
    HISTO_LENGTH = 30

    def __init__(self, nnratio=0.6, checkOri=True, *, device=0):
        self.mfNNratio = float(np.float32(nnratio))
        self.mbCheckOrientation = bool(checkOri)
        self.device = device
        h = C.c_void_p()
        check(lib().orbx_matcher_create(device, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_matcher_destroy(self._h)
            self._h = None

    __del__ = close

    def DescriptorDistance(self, a, b):
        """Hamming distance of 32-byte rows; a, b: (n,32) or (32,) uint8 (R/src/ORBmatcher.cpp:1901-1917)."""
        a = np.ascontiguousarray(a, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(b, np.uint8).reshape(-1, 32)
        if a.shape != b.shape:
            raise ValueError("a and b must hold the same number of descriptors")
        d = np.empty(a.shape[0], np.int32)
        check(lib().orbx_descriptor_distance(self._h, ptr(a), ptr(b), ptr(d), a.shape[0]))
        return d

    def match_bf2(self, q, db, th=None):
        """Brute-force top-2 + ratio test: (idx, best, second, accepted) per query."""
        q = np.ascontiguousarray(q, np.uint8).reshape(-1, 32)
        db = np.ascontiguousarray(db, np.uint8).reshape(-1, 32)
        nq = q.shape[0]
        idx, d1, d2 = (np.empty(nq, np.int32) for _ in range(3))
        acc = np.empty(nq, np.uint8)
        check(lib().orbx_match_bf2(self._h, ptr(q), nq, ptr(db), db.shape[0], self.mfNNratio,
                                   self.TH_LOW if th is None else th, ptr(idx), ptr(d1), ptr(d2), ptr(acc)))
        return idx, d1, d2, acc

    def match_bf2_device(self, d_q, nq, d_db, ndb, idx_base, d_idx, d_d1, d_d2, stream=0):
        check(lib().orbx_match_bf2_device(self._h, d_q, nq, d_db, ndb, idx_base, d_idx, d_d1, d_d2, stream))

    def merge_device(self, d_idx, d_d1, d_d2, nshards, nq, th, o_idx, o_d1, o_d2, o_acc, stream=0):
        check(lib().orbx_match_merge_device(self._h, d_idx, d_d1, d_d2, nshards, nq, self.mfNNratio, th,
                                            o_idx, o_d1, o_d2, o_acc, stream))

    def launch_count(self):
        return int(lib().orbx_matcher_launch_count(self._h))
