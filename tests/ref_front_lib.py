"""ctypes binding of oracle/_ref/libref_front.so -- TEST INFRASTRUCTURE ONLY.

The library is the reference's own Frame.cpp / ORBmatcher.cpp / MapPoint.cpp / Map.cpp / ORBextractor.cpp compiled
where they lie against the cv shim (oracle/Makefile, oracle/ref_shim/ref_front_capi.cpp). It exists only where it was
built in the authoring container (git-ignored; it travels to the GPU box with the snapshot). Tests skip without it.
"""
import ctypes as C
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
LIB_PATH = ROOT / "oracle" / "_ref" / "libref_front.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
P = C.c_void_p
_lib = None


def available():
    return LIB_PATH.exists()


def lib():
    global _lib
    if _lib is None:
        L = C.CDLL(str(LIB_PATH))
        L.ref_set_allocator_mode.argtypes = [C.c_int]
        L.ref_front_create.restype = P
        L.ref_front_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int] + [C.c_float] * 6
        L.ref_front_destroy.argtypes = [P]
        L.ref_frame_mono.restype = P
        L.ref_frame_mono.argtypes = [P, P, C.c_int, C.c_int, C.c_int]
        L.ref_frame_stereo.restype = P
        L.ref_frame_stereo.argtypes = [P, P, P, C.c_int, C.c_int, C.c_int]
        L.ref_frame_destroy.argtypes = [P]
        L.ref_frame_n.argtypes = [P]
        L.ref_frame_n_right.argtypes = [P]
        L.ref_frame_get.argtypes = [P] * 8
        L.ref_frame_get_right.argtypes = [P] * 3
        L.ref_front_level.argtypes = [P, C.c_int, C.c_int, P, P, P, C.c_int]
        L.ref_frame_set_pose.argtypes = [P, P]
        L.ref_search_for_initialization.argtypes = [P, P, P, P, C.c_int, C.c_float, C.c_int]
        L.ref_map_point_create.argtypes = [P, P, C.c_int, P]
        L.ref_map_point_set_track.argtypes = [P, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int, C.c_float]
        L.ref_map_point_set_nobs.argtypes = [P, C.c_int, C.c_int]
        L.ref_frame_set_map_point.argtypes = [P, P, C.c_int, C.c_int]
        L.ref_frame_set_outlier.argtypes = [P, C.c_int, C.c_int]
        L.ref_frame_get_map_points.argtypes = [P, P, P]
        L.ref_search_by_projection_mp.argtypes = [P, P, P, C.c_int, C.c_float, C.c_float, C.c_int]
        L.ref_search_by_projection_frame.argtypes = [P, P, C.c_float, C.c_int, C.c_float, C.c_int]
        L.ref_project_last_frame.argtypes = [P, P, C.c_int] + [P] * 6
        L.ref_keyframe_create.restype = P
        L.ref_keyframe_create.argtypes = [P, P]
        L.ref_frame_set_featvec.argtypes = [P, P, C.c_int]
        L.ref_keyframe_set_featvec.argtypes = [P, P, C.c_int]
        L.ref_keyframe_add_map_point.argtypes = [P, P, C.c_int, C.c_int]
        L.ref_keyframe_address_rank.argtypes = [P, P]
        L.ref_search_by_bow_kf_frame.argtypes = [P, P, P, C.c_float, C.c_int, P]
        L.ref_search_by_bow_kf_kf.argtypes = [P, P, P, C.c_float, C.c_int, P]
        L.ref_epipole.argtypes = [P, P, P, P]
        L.ref_search_for_triangulation.argtypes = [P, P, P, C.c_int, C.c_float, C.c_int, P, C.c_int]
        L.ref_keyframe_tables.argtypes = [P, P, P]
        L.ref_map_point_observed_rows.argtypes = [P, C.c_int, P, C.c_int]
        L.ref_map_point_distinctive.argtypes = [P, C.c_int, P]
        L.ref_fuse_queries.argtypes = [P, P, P, C.c_int, C.c_float] + [P] * 6
        L.ref_fuse.argtypes = [P, P, P, C.c_int, C.c_float]
        L.ref_keyframe_get_map_points.argtypes = [P, P, P]
        L.ref_keyframe_inv_level_sigma2.argtypes = [P, P]
        L.ref_map_point_observations.argtypes = [P, C.c_int]
        L.ref_map_point_is_bad.argtypes = [P, C.c_int]
        L.ref_scw_queries.argtypes = [P, P, P, P, C.c_int, P, C.c_int] + [P] * 5
        L.ref_search_by_projection_kf_scw.argtypes = [P, P, P, P, C.c_int, P, C.c_int]
        L.ref_search_by_projection_frame_kf.argtypes = [P, P, P, P, C.c_int, C.c_float, C.c_int, C.c_float, C.c_int]
        L.ref_fuse_scw.argtypes = [P, P, P, P, C.c_int, C.c_float, P]
        L.ref_search_by_sim3.argtypes = [P, P, P, P, C.c_float, P, P, C.c_float]
        L.ref_descriptor_distance.argtypes = [P, P]
        L.ref_set_allocator_mode(1)   # pointer order == creation order: the canonical DistributeOctTree tie-break
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(P)


# the reference's objects with the product's ORBmatcher.cc (matcher) / ORBmatcher.cc and ORBextractor.cc (front)
DROPIN_MATCHER_PATH = ROOT / "oracle" / "_ref" / "libdropin_matcher.so"
DROPIN_PATH = ROOT / "oracle" / "_ref" / "libdropin_front.so"


def variant(path, name=None):
    """A second instance of this module bound to another build of the same harness: oracle/_ref/libdropin_front.so =
    the reference's Frame / KeyFrame / MapPoint objects with R/src/ORBmatcher.cpp replaced by the product's
    orb-slam2-_b200/host/ORBmatcher.cc over liborbx.so (needs a GPU)."""
    import importlib.util
    import sys
    name = name or "ref_front_lib_" + Path(path).stem
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, __file__)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    mod.LIB_PATH = Path(path)
    return mod


class Frame:
    def __init__(self, front, handle):
        self.front, self._h = front, handle
        self.N = lib().ref_frame_n(handle)
        n = self.N
        self.keys, self.keys_un = np.empty(n, KP_DTYPE), np.empty(n, KP_DTYPE)
        self.desc = np.empty((n, 32), np.uint8)
        self.u_right, self.depth = np.empty(n, np.float32), np.empty(n, np.float32)
        self.bounds = np.empty(4, np.float32)
        mb = np.empty(1, np.float32)
        lib().ref_frame_get(handle, _p(self.keys), _p(self.keys_un), _p(self.desc), _p(self.u_right), _p(self.depth),
                            _p(self.bounds), _p(mb))
        self.mb = mb[0]
        nr = lib().ref_frame_n_right(handle)
        self.keys_right, self.desc_right = np.empty(nr, KP_DTYPE), np.empty((nr, 32), np.uint8)
        if nr:
            lib().ref_frame_get_right(handle, _p(self.keys_right), _p(self.desc_right))

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_frame_destroy(self._h)
            self._h = None

    def set_pose(self, Tcw):
        T = np.ascontiguousarray(Tcw, np.float32).reshape(4, 4)
        lib().ref_frame_set_pose(self._h, _p(T))

    def set_map_point(self, idx, mp):
        lib().ref_frame_set_map_point(self.front._h, self._h, idx, -1 if mp is None else mp)

    def set_outlier(self, idx, flag):
        lib().ref_frame_set_outlier(self._h, idx, int(flag))

    def map_points(self):
        out = np.empty(self.N, np.int32)
        lib().ref_frame_get_map_points(self.front._h, self._h, _p(out))
        return out

    def set_featvec(self, node):
        node = np.ascontiguousarray(node, np.int32)
        lib().ref_frame_set_featvec(self._h, _p(node), len(node))


class KeyFrame:
    """A reference KeyFrame built from a reference Frame (owned by the Front)."""

    def __init__(self, front, frame):
        self.front, self.frame = front, frame
        self._h = lib().ref_keyframe_create(front._h, frame._h)
        self.N = frame.N

    def set_featvec(self, node):
        node = np.ascontiguousarray(node, np.int32)
        lib().ref_keyframe_set_featvec(self._h, _p(node), len(node))

    def add_map_point(self, idx, mp):
        lib().ref_keyframe_add_map_point(self.front._h, self._h, idx, mp)

    def tables(self):
        sf, ls = np.empty(self.front.nlevels, np.float32), np.empty(self.front.nlevels, np.float32)
        lib().ref_keyframe_tables(self._h, _p(sf), _p(ls))
        return sf, ls

    def inv_level_sigma2(self):
        out = np.empty(self.front.nlevels, np.float32)
        lib().ref_keyframe_inv_level_sigma2(self._h, _p(out))
        return out

    def map_points(self):
        out = np.empty(max(self.N, 1), np.int32)
        lib().ref_keyframe_get_map_points(self.front._h, self._h, _p(out))
        return out[:self.N]


class Front:
    """Two reference ORBextractors + camera + Map, and factories for reference Frame / MapPoint objects."""

    def __init__(self, nfeatures=1000, scale=1.2, nlevels=8, ini_th=20, min_th=7, fx=500.0, fy=500.0, cx=320.0, cy=240.0,
                 bf=40.0, th_depth=35.0):
        self.fx, self.fy, self.cx, self.cy, self.bf = (np.float32(v) for v in (fx, fy, cx, cy, bf))
        self.nlevels = nlevels
        self._h = lib().ref_front_create(nfeatures, scale, nlevels, ini_th, min_th, fx, fy, cx, cy, bf, th_depth)

    def __del__(self):
        if getattr(self, "_h", None):
            lib().ref_front_destroy(self._h)
            self._h = None

    def frame_mono(self, img):
        img = np.ascontiguousarray(img, np.uint8)
        return Frame(self, lib().ref_frame_mono(self._h, _p(img), img.shape[1], img.shape[0], img.strides[0]))

    def frame_stereo(self, left, right):
        left, right = np.ascontiguousarray(left, np.uint8), np.ascontiguousarray(right, np.uint8)
        assert left.shape == right.shape
        return Frame(self, lib().ref_frame_stereo(self._h, _p(left), _p(right), left.shape[1], left.shape[0], left.strides[0]))

    def level(self, side, level):
        w, h = C.c_int(), C.c_int()
        lib().ref_front_level(self._h, side, level, C.byref(w), C.byref(h), None, 0)
        out = np.empty((h.value, w.value), np.uint8)
        lib().ref_front_level(self._h, side, level, None, None, _p(out), out.strides[0])
        return out

    def map_point(self, frame, idx, pos, n_obs=1):
        pos = np.ascontiguousarray(pos, np.float32)
        mp = lib().ref_map_point_create(self._h, frame._h, idx, _p(pos))
        lib().ref_map_point_set_nobs(self._h, mp, n_obs)
        return mp

    def set_track(self, mp, in_view, u, v, u_r, level, view_cos):
        lib().ref_map_point_set_track(self._h, mp, int(in_view), u, v, u_r, level, view_cos)

    def search_for_initialization(self, f1, f2, prev_matched, window, nnratio, check_ori=True):
        pm = np.ascontiguousarray(prev_matched, np.float32).copy()
        m12 = np.empty(f1.N, np.int32)
        n = lib().ref_search_for_initialization(f1._h, f2._h, _p(pm), _p(m12), window, nnratio, int(check_ori))
        return n, m12, pm

    def search_by_projection_mp(self, frame, mps, th, nnratio, check_ori=True):
        mps = np.ascontiguousarray(mps, np.int32)
        return lib().ref_search_by_projection_mp(self._h, frame._h, _p(mps), len(mps), th, nnratio, int(check_ori))

    def search_by_projection_frame(self, cur, last, th, mono, nnratio=0.9, check_ori=True):
        return lib().ref_search_by_projection_frame(cur._h, last._h, th, int(mono), nnratio, int(check_ori))

    def project_last_frame(self, cur, last, mono):
        n = last.N
        valid = np.empty(n, np.uint8)
        u, v, invz = (np.empty(n, np.float32) for _ in range(3))
        fwd, bwd = C.c_int(), C.c_int()
        lib().ref_project_last_frame(cur._h, last._h, int(mono), _p(valid), _p(u), _p(v), _p(invz), C.byref(fwd), C.byref(bwd))
        return valid.astype(bool), u, v, invz, bool(fwd.value), bool(bwd.value)

    # ---- key frames, bag-of-words restricted searches, distinctive descriptors ----
    def keyframe(self, frame):
        return KeyFrame(self, frame)

    def search_by_bow_kf_frame(self, kf, frame, nnratio, check_ori=True):
        out = np.empty(max(frame.N, 1), np.int32)
        n = lib().ref_search_by_bow_kf_frame(self._h, kf._h, frame._h, nnratio, int(check_ori), _p(out))
        return n, out[:frame.N]

    def search_by_bow_kf_kf(self, kf1, kf2, nnratio, check_ori=True):
        out = np.empty(max(kf1.N, 1), np.int32)
        n = lib().ref_search_by_bow_kf_kf(self._h, kf1._h, kf2._h, nnratio, int(check_ori), _p(out))
        return n, out[:kf1.N]

    def epipole(self, kf1, kf2):
        ex, ey = C.c_float(), C.c_float()
        lib().ref_epipole(kf1._h, kf2._h, C.byref(ex), C.byref(ey))
        return np.float32(ex.value), np.float32(ey.value)

    def search_for_triangulation(self, kf1, kf2, F12, only_stereo, nnratio=0.6, check_ori=False):
        F = np.ascontiguousarray(F12, np.float32).reshape(9)
        pairs = np.empty((max(kf1.N, 1), 2), np.int32)
        n = lib().ref_search_for_triangulation(kf1._h, kf2._h, _p(F), int(only_stereo), nnratio, int(check_ori), _p(pairs),
                                               len(pairs))
        assert n >= 0
        return n, pairs[:n]

    def observed_rows(self, mp):
        n = lib().ref_map_point_observed_rows(self._h, mp, None, 0)
        rows = np.empty((max(n, 1), 32), np.uint8)
        lib().ref_map_point_observed_rows(self._h, mp, _p(rows), n)
        return rows[:n]

    def distinctive(self, mp):
        out = np.empty(32, np.uint8)
        lib().ref_map_point_distinctive(self._h, mp, _p(out))
        return out

    # ---- Fuse / SearchByProjection(KeyFrame*, Scw, ...) ----
    def fuse_queries(self, kf, mps, th):
        mps = np.ascontiguousarray(mps, np.int32)
        n = len(mps)
        valid, uv, ur, radius = np.empty(n, np.uint8), np.empty((n, 2), np.float32), np.empty(n, np.float32), np.empty(n, np.float32)
        level, desc = np.empty(n, np.int32), np.empty((n, 32), np.uint8)
        lib().ref_fuse_queries(self._h, kf._h, _p(mps), n, th, _p(valid), _p(uv), _p(ur), _p(radius), _p(level), _p(desc))
        return valid.astype(bool), uv, ur, radius, level, desc

    def fuse(self, kf, mps, th):
        mps = np.ascontiguousarray(mps, np.int32)
        return lib().ref_fuse(self._h, kf._h, _p(mps), len(mps), th)

    def observations(self, mp):
        return lib().ref_map_point_observations(self._h, int(mp))

    def is_bad(self, mp):
        return bool(lib().ref_map_point_is_bad(self._h, int(mp)))

    def scw_queries(self, kf, Scw, mps, matched, th):
        mps, matched = np.ascontiguousarray(mps, np.int32), np.ascontiguousarray(matched, np.int32)
        S = np.ascontiguousarray(Scw, np.float32).reshape(16)
        n = len(mps)
        valid, uv, radius = np.empty(n, np.uint8), np.empty((n, 2), np.float32), np.empty(n, np.float32)
        level, desc = np.empty(n, np.int32), np.empty((n, 32), np.uint8)
        lib().ref_scw_queries(self._h, kf._h, _p(S), _p(mps), n, _p(matched), th, _p(valid), _p(uv), _p(radius), _p(level), _p(desc))
        return valid.astype(bool), uv, radius, level, desc

    def search_by_projection_kf_scw(self, kf, Scw, mps, matched, th):
        mps = np.ascontiguousarray(mps, np.int32)
        out = np.ascontiguousarray(matched, np.int32).copy()
        S = np.ascontiguousarray(Scw, np.float32).reshape(16)
        n = lib().ref_search_by_projection_kf_scw(self._h, kf._h, _p(S), _p(mps), len(mps), _p(out), th)
        return n, out

    # ---- the remaining ORBmatcher members (relocalization, loop closing) ----
    def search_by_projection_frame_kf(self, cur, kf, found, th, orb_dist, nnratio=0.9, check_ori=True):
        found = np.ascontiguousarray(found, np.int32)
        return lib().ref_search_by_projection_frame_kf(self._h, cur._h, kf._h, _p(found), len(found), th, orb_dist, nnratio,
                                                       int(check_ori))

    def fuse_scw(self, kf, Scw, mps, th):
        mps = np.ascontiguousarray(mps, np.int32)
        S = np.ascontiguousarray(Scw, np.float32).reshape(16)
        rep = np.empty(max(len(mps), 1), np.int32)
        n = lib().ref_fuse_scw(self._h, kf._h, _p(S), _p(mps), len(mps), th, _p(rep))
        return n, rep[:len(mps)]

    def search_by_sim3(self, kf1, kf2, matches12, s12, R12, t12, th):
        out = np.ascontiguousarray(matches12, np.int32).copy()
        R, t = np.ascontiguousarray(R12, np.float32).reshape(9), np.ascontiguousarray(t12, np.float32).reshape(3)
        n = lib().ref_search_by_sim3(self._h, kf1._h, kf2._h, _p(out), s12, _p(R), _p(t), th)
        return n, out

    def descriptor_distance(self, a, b):
        a, b = np.ascontiguousarray(a, np.uint8), np.ascontiguousarray(b, np.uint8)
        return lib().ref_descriptor_distance(_p(a), _p(b))
