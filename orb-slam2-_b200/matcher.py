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

    def match_window(self, kps_un, desc, bounds, q_xy, q_r, q_min_level, q_max_level, q_desc):
        """Candidates of Frame::GetFeaturesInArea for every query, in the reference's order, with their Hamming
        distances: (cand_off[nq+1], cand_idx, cand_dist). bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
        kps_un = np.ascontiguousarray(kps_un)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        q_xy = np.ascontiguousarray(q_xy, np.float32).reshape(-1, 2)
        nq = q_xy.shape[0]
        q_r = np.ascontiguousarray(q_r, np.float32)
        q_min_level = np.ascontiguousarray(q_min_level, np.int32)
        q_max_level = np.ascontiguousarray(q_max_level, np.int32)
        q_desc = np.ascontiguousarray(q_desc, np.uint8).reshape(-1, 32)
        if not (len(q_r) == len(q_min_level) == len(q_max_level) == len(q_desc) == nq) or len(desc) != len(kps_un):
            raise ValueError("query / frame arrays disagree in length")
        off = np.zeros(nq + 1, np.int32)
        cap = max(64, 32 * nq)
        while True:
            idx, dist = np.empty(cap, np.int32), np.empty(cap, np.int32)
            n = C.c_int32(0)
            rc = lib().orbx_match_window(self._h, ptr(kps_un), ptr(desc), len(kps_un), bounds[0], bounds[1], bounds[2],
                                         bounds[3], ptr(q_xy), ptr(q_r), ptr(q_min_level), ptr(q_max_level), ptr(q_desc),
                                         nq, ptr(off), ptr(idx), ptr(dist), cap, C.byref(n))
            if rc == -3 and n.value > cap:
                cap = n.value
                continue
            check(rc)
            return off, idx[:n.value].copy(), dist[:n.value].copy()

    def stereo_match(self, ex_left, ex_right, kps_l, desc_l, kps_r, desc_r, mbf, mb, frame_left=0, frame_right=0):
        """Frame::ComputeStereoMatches on the pyramids the two extractors still hold: (mvuRight, mvDepth, n)."""
        kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
        desc_l = np.ascontiguousarray(desc_l, np.uint8).reshape(-1, 32)
        desc_r = np.ascontiguousarray(desc_r, np.uint8).reshape(-1, 32)
        nl = len(kps_l)
        u, z = np.empty(max(nl, 1), np.float32), np.empty(max(nl, 1), np.float32)
        n = C.c_int32(0)
        check(lib().orbx_stereo_match(self._h, ex_left._h, ex_right._h, frame_left, frame_right, ptr(kps_l), ptr(desc_l), nl,
                                      ptr(kps_r), ptr(desc_r), len(kps_r), mbf, mb, ptr(u), ptr(z), C.byref(n)))
        return u[:nl], z[:nl], n.value

    def stereo_match_batch_device(self, ex, pairs, left_frame0, right_frame0, frame_step, d_kps, d_desc, d_n, cap, mbf, mb,
                                  d_u_right, d_depth, d_n_matched, stream=0):
        """ComputeStereoMatches for `pairs` pairs of the device-resident batch `ex` processed last (raw device
        pointers, asynchronous): pair p = frames left_frame0 + p*frame_step / right_frame0 + p*frame_step."""
        check(lib().orbx_stereo_match_batch_device(self._h, ex._h, pairs, left_frame0, right_frame0, frame_step, d_kps, d_desc,
                                                   d_n, cap, mbf, mb, d_u_right, d_depth, d_n_matched, stream))

    def match_blocks(self, desc_a, desc_b, a_off, a_idx, b_off, b_idx):
        """Distance matrices of the vocabulary nodes two feature vectors share (SearchByBoW / SearchForTriangulation,
        R/src/ORBmatcher.cpp:220-365, :632-772, :785-983): (dist_off[nblocks+1] int64, dist uint16), block k row-major
        [a entries][b entries] at dist[dist_off[k]:dist_off[k+1]]."""
        desc_a = np.ascontiguousarray(desc_a, np.uint8).reshape(-1, 32)
        desc_b = np.ascontiguousarray(desc_b, np.uint8).reshape(-1, 32)
        a_off, a_idx = np.ascontiguousarray(a_off, np.int32), np.ascontiguousarray(a_idx, np.int32)
        b_off, b_idx = np.ascontiguousarray(b_off, np.int32), np.ascontiguousarray(b_idx, np.int32)
        if len(a_off) != len(b_off) or len(a_off) < 1:
            raise ValueError("a_off and b_off must hold nblocks+1 entries each")
        nblocks = len(a_off) - 1
        off = np.zeros(nblocks + 1, np.int64)
        need = int(np.sum(np.diff(a_off).astype(np.int64) * np.diff(b_off).astype(np.int64))) if nblocks else 0
        dist = np.empty(max(need, 1), np.uint16)
        n = C.c_int64(0)
        check(lib().orbx_match_blocks(self._h, ptr(desc_a), len(desc_a), ptr(desc_b), len(desc_b), ptr(a_off), ptr(a_idx),
                                      ptr(b_off), ptr(b_idx), nblocks, ptr(off), ptr(dist), len(dist), C.byref(n)))
        return off, dist[:n.value]

    def distinctive(self, desc, obs_off):
        """MapPoint::ComputeDistinctiveDescriptors (R/src/MapPoint.cpp:306-385) for many points: point p observes rows
        desc[obs_off[p]:obs_off[p+1]]; returns (best position within the group or -1, its median distance)."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        obs_off = np.ascontiguousarray(obs_off, np.int32)
        npoints = len(obs_off) - 1
        if npoints < 0 or (npoints > 0 and obs_off[-1] != len(desc)):
            raise ValueError("obs_off must be a CSR over the descriptor rows")
        best, med = np.empty(max(npoints, 1), np.int32), np.empty(max(npoints, 1), np.int32)
        check(lib().orbx_distinctive(self._h, ptr(desc), ptr(obs_off), npoints, ptr(best), ptr(med)))
        return best[:npoints], med[:npoints]

    def launch_count(self):
        return int(lib().orbx_matcher_launch_count(self._h))


def undistort_points(xy, K4, dist, device=0):
    """cv::undistortPoints(xy, K, dist, None, K) on the device (R/src/Frame.cpp:491, :523): (n, 2) float32."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    K4 = np.ascontiguousarray(K4, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    out = np.empty_like(xy)
    check(lib().orbx_undistort_points(device, ptr(xy), len(xy), ptr(K4), ptr(dist), len(dist), ptr(out)))
    return out


class Frame:
    """Device-resident mirror of what Frame::Frame leaves for the matchers (R/src/Frame.cpp:187-250): mvKeysUn
    (UndistortKeyPoints), mDescriptors and the 64x48 grid (AssignFeaturesToGrid), kept in HBM on the matcher's device."""

    def __init__(self, matcher):
        self.matcher = matcher
        h = C.c_void_p()
        check(lib().orbx_frame_create(matcher._h, C.byref(h)))
        self._h = h
        self.N = 0

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_frame_destroy(self._h)
            self._h = None

    __del__ = close

    def set(self, kps, desc, bounds, K4=None, dist=None):
        """Host arrays: keypoints (KP_DTYPE) and descriptors of one frame; K4 = (fx, fy, cx, cy) and dist = distortion
        coefficients (None or dist[0] == 0: no undistortion); bounds = (mnMinX, mnMaxX, mnMinY, mnMaxY)."""
        kps = np.ascontiguousarray(kps)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        b = np.ascontiguousarray(bounds, np.float32)
        k = np.zeros(4, np.float32) if K4 is None else np.ascontiguousarray(K4, np.float32)
        d = np.zeros(0, np.float32) if dist is None else np.ascontiguousarray(dist, np.float32)
        check(lib().orbx_frame_set(self._h, ptr(kps), ptr(desc), len(kps), 0, ptr(k), ptr(d), len(d), ptr(b)))
        self.N = len(kps)
        self._dtype = kps.dtype

    def set_device(self, d_kps_ptr, d_desc_ptr, n, bounds, K4=None, dist=None):
        """The same from device pointers (e.g. one frame's rows of orbx_extract_batch_device's outputs)."""
        b = np.ascontiguousarray(bounds, np.float32)
        k = np.zeros(4, np.float32) if K4 is None else np.ascontiguousarray(K4, np.float32)
        d = np.zeros(0, np.float32) if dist is None else np.ascontiguousarray(dist, np.float32)
        check(lib().orbx_frame_set(self._h, d_kps_ptr, d_desc_ptr, n, 1, ptr(k), ptr(d), len(d), ptr(b)))
        self.N = n

    def keys_un(self):
        from ._capi import KP_DTYPE
        out = np.empty(max(self.N, 1), KP_DTYPE)
        check(lib().orbx_frame_get_keys_un(self._h, ptr(out)))
        return out[:self.N]

    def match_window(self, q_xy, q_r, q_min_level, q_max_level, q_desc):
        """ORBmatcher.match_window against the resident frame: (cand_off, cand_idx, cand_dist)."""
        q_xy = np.ascontiguousarray(q_xy, np.float32).reshape(-1, 2)
        nq = q_xy.shape[0]
        q_r = np.ascontiguousarray(q_r, np.float32)
        q_min_level = np.ascontiguousarray(q_min_level, np.int32)
        q_max_level = np.ascontiguousarray(q_max_level, np.int32)
        q_desc = np.ascontiguousarray(q_desc, np.uint8).reshape(-1, 32)
        if not (len(q_r) == len(q_min_level) == len(q_max_level) == len(q_desc) == nq):
            raise ValueError("query arrays disagree in length")
        off = np.zeros(nq + 1, np.int32)
        cap = max(64, 32 * nq)
        while True:
            idx, dist = np.empty(cap, np.int32), np.empty(cap, np.int32)
            n = C.c_int32(0)
            rc = lib().orbx_frame_match_window(self._h, ptr(q_xy), ptr(q_r), ptr(q_min_level), ptr(q_max_level), ptr(q_desc), nq,
                                               ptr(off), ptr(idx), ptr(dist), cap, C.byref(n))
            if rc == -3 and n.value > cap:
                cap = n.value
                continue
            check(rc)
            return off, idx[:n.value].copy(), dist[:n.value].copy()


class Vocabulary:
    """The tree of a DBoW2 vocabulary resident on the device (orbx_vocabulary_*): parent[i] < i per node (node 0 =
    root), node descriptors (nodes, 32) uint8, depth L."""

    def __init__(self, matcher, parent, desc, depth):
        self.matcher = matcher
        parent = np.ascontiguousarray(parent, np.int32)
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        if len(parent) != len(desc):
            raise ValueError("parent and desc disagree in length")
        h = C.c_void_p()
        check(lib().orbx_vocabulary_create(matcher._h, len(parent), ptr(parent), ptr(desc), depth, C.byref(h)))
        self._h = h

    def close(self):
        if getattr(self, "_h", None):
            lib().orbx_vocabulary_destroy(self._h)
            self._h = None

    __del__ = close

    def descend(self, desc, levelsup):
        """Per descriptor: (leaf node, node at level L - levelsup) of TemplatedVocabulary::transform's descent."""
        desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
        n = len(desc)
        leaf, level = np.empty(max(n, 1), np.int32), np.empty(max(n, 1), np.int32)
        check(lib().orbx_vocabulary_descend(self._h, ptr(desc), n, levelsup, ptr(leaf), ptr(level)))
        return leaf[:n], level[:n]
