"""ctypes binding of oracle/liborb_oracle.so -- TEST INFRASTRUCTURE ONLY.

May be imported from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs. The product package (orb-slam2-_b200/) never imports this module.
"""
import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
ORACLE_DIR = ROOT / "oracle"
LIB_PATH = ORACLE_DIR / "liborb_oracle.so"

KP_DTYPE = np.dtype([("x", "<f4"), ("y", "<f4"), ("size", "<f4"), ("angle", "<f4"),
                     ("response", "<f4"), ("octave", "<i4"), ("class_id", "<i4")])
assert KP_DTYPE.itemsize == 28

_u8p = C.POINTER(C.c_uint8)
_i32p = C.POINTER(C.c_int32)
_f32p = C.POINTER(C.c_float)


def build(force=False):
    """Compile the oracle with its Makefile (gcc only; seconds)."""
    srcs = list(ORACLE_DIR.glob("*.cpp")) + list(ORACLE_DIR.glob("*.h"))
    stale = (not LIB_PATH.exists()) or any(s.stat().st_mtime > LIB_PATH.stat().st_mtime for s in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", str(ORACLE_DIR), "-j8"], stdout=subprocess.DEVNULL)
    return LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(str(LIB_PATH))
        L = _lib
        L.orb_oracle_fast_atan2.restype = C.c_float
        L.orb_oracle_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.orb_oracle_sincosf_port.argtypes = [C.c_float, _f32p, _f32p]
        L.orb_oracle_sincosf_libm.argtypes = [C.c_float, _f32p, _f32p]
        L.orb_oracle_cv_round.argtypes = [C.c_double]
        L.orb_oracle_sincosf_sweep.restype = C.c_int64
        L.orb_oracle_sincosf_sweep.argtypes = [C.c_uint32, C.c_uint32, C.c_int, C.POINTER(C.c_uint32)]
        L.orb_oracle_extractor_create.restype = C.c_void_p
        L.orb_oracle_extractor_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        L.orb_oracle_extractor_destroy.argtypes = [C.c_void_p]
        L.orb_oracle_extractor_set_tie_mode.argtypes = [C.c_void_p, C.c_int]
        L.orb_oracle_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                         C.c_void_p, C.c_int]
        L.orb_oracle_extractor_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 6
        L.orb_oracle_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.orb_oracle_get_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_get_blurred_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_level_candidates.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_level_kept.argtypes = [C.c_void_p, C.c_int]
        L.orb_oracle_level_tie_sensitive.argtypes = [C.c_void_p, C.c_int]
        L.orb_oracle_distribute.argtypes = [C.c_void_p, C.c_int] + [C.c_int] * 6 + [C.c_void_p, C.c_int]
        L.orb_oracle_resize.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.orb_oracle_blur7.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_fast.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_descriptor_distance.argtypes = [C.c_void_p, C.c_void_p]
        L.orb_oracle_match_bf2.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int64, C.c_float, C.c_int,
                                           C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
        L.orb_oracle_grid_create.restype = C.c_void_p
        L.orb_oracle_grid_create.argtypes = [C.c_void_p, C.c_int] + [C.c_float] * 4
        L.orb_oracle_grid_destroy.argtypes = [C.c_void_p]
        L.orb_oracle_grid_query.argtypes = [C.c_void_p, C.c_float, C.c_float, C.c_float, C.c_int, C.c_int, C.c_void_p, C.c_int]
        L.orb_oracle_search_for_initialization.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + \
            [C.c_float] * 4 + [C.c_void_p, C.c_void_p, C.c_int, C.c_float, C.c_int]
        L.orb_oracle_search_by_projection_mp.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p] + \
            [C.c_float] * 4 + [C.c_void_p] * 5 + [C.c_int, C.c_void_p, C.c_float, C.c_void_p, C.c_void_p]
        L.orb_oracle_search_by_projection_frame.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p] + \
            [C.c_float] * 4 + [C.c_void_p] * 7 + [C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
        L.orb_oracle_stereo_matches.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int] + \
            [C.c_void_p] * 5 + [C.c_int, C.c_void_p, C.c_void_p, C.c_float, C.c_float, C.c_void_p, C.c_void_p]
        V, I, F = C.c_void_p, C.c_int, C.c_float
        L.orb_oracle_block_distances.argtypes = [V] * 6 + [I, V, V]
        L.orb_oracle_block_distances.restype = None
        L.orb_oracle_search_by_bow_kf_frame.argtypes = [V, V, V, V, V, V, I, V, V, I, V, V, V, I, F, I, V]
        L.orb_oracle_search_by_bow_kf_kf.argtypes = [V, V, V, I, V, V, V, I, V, V, V, I, V, V, V, I, F, I, V]
        L.orb_oracle_search_for_triangulation.argtypes = [V, V, V, V, I, V, V, V, I, V, V, V, V, I, V, V, V, I, V, F, F, V, V,
                                                          I, I, V]
        L.orb_oracle_distinctive.argtypes = [V, V, I, V, V]
        L.orb_oracle_distinctive.restype = None
        L.orb_oracle_search_best_in_window.argtypes = [V, V, I, V, F, F, F, F, V, V, V, V, V, V, I, V, I, V, V]
        L.orb_oracle_search_best_in_window.restype = None
        L.orb_oracle_undistort_points.argtypes = [V, C.c_longlong, V, V, I, V]
        L.orb_oracle_undistort_points.restype = None
        L.orb_oracle_voc_load.restype = V
        L.orb_oracle_voc_load.argtypes = [C.c_char_p]
        L.orb_oracle_voc_destroy.argtypes = [V]
        L.orb_oracle_voc_nodes.argtypes = [V]
        L.orb_oracle_voc_words.argtypes = [V]
        L.orb_oracle_voc_export.argtypes = [V] * 6
        L.orb_oracle_voc_export.restype = None
        L.orb_oracle_voc_transform.argtypes = [V, V, I, I] + [V] * 7
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


def _img(a):
    a = np.asarray(a)
    assert a.dtype == np.uint8 and a.ndim == 2 and a.strides[1] == 1
    return a


# ---- primitives ------------------------------------------------------------------------------
def resize(src, dw, dh):
    src = _img(src)
    dst = np.empty((dh, dw), np.uint8)
    lib().orb_oracle_resize(_ptr(src), src.shape[1], src.shape[0], src.strides[0], _ptr(dst), dw, dh, dw)
    return dst


def blur7(src):
    src = _img(src)
    dst = np.empty(src.shape, np.uint8)
    lib().orb_oracle_blur7(_ptr(src), src.shape[1], src.shape[0], src.strides[0], _ptr(dst), src.shape[1])
    return dst


def fast(img, threshold, nms=True):
    img = _img(img)
    cap = max(16, img.size)
    out = np.empty((cap, 3), np.int32)
    n = lib().orb_oracle_fast(_ptr(img), img.shape[1], img.shape[0], img.strides[0], threshold, int(nms), _ptr(out), cap)
    return out[:n].copy()


def fast_atan2(y, x):
    return lib().orb_oracle_fast_atan2(float(y), float(x))


def sincosf_port(a):
    s, c = C.c_float(), C.c_float()
    lib().orb_oracle_sincosf_port(float(a), C.byref(s), C.byref(c))
    return s.value, c.value


def sincosf_libm(a):
    s, c = C.c_float(), C.c_float()
    lib().orb_oracle_sincosf_libm(float(a), C.byref(s), C.byref(c))
    return s.value, c.value


def sincosf_sweep(lo_bits, hi_bits, nthreads=None):
    first = C.c_uint32(0)
    n = lib().orb_oracle_sincosf_sweep(lo_bits, hi_bits, nthreads or os.cpu_count() or 1, C.byref(first))
    return n, first.value


def descriptor_distance(a, b):
    a = np.ascontiguousarray(a, np.uint8)
    b = np.ascontiguousarray(b, np.uint8)
    return lib().orb_oracle_descriptor_distance(_ptr(a), _ptr(b))


def match_bf2(q, db, ratio, th, nthreads=None):
    q = np.ascontiguousarray(q, np.uint8)
    db = np.ascontiguousarray(db, np.uint8)
    nq, ndb = q.shape[0], db.shape[0]
    idx = np.empty(nq, np.int32)
    d1 = np.empty(nq, np.int32)
    d2 = np.empty(nq, np.int32)
    acc = np.empty(nq, np.uint8)
    lib().orb_oracle_match_bf2(_ptr(q), nq, _ptr(db), ndb, ratio, th, _ptr(idx), _ptr(d1), _ptr(d2), _ptr(acc),
                               nthreads or os.cpu_count() or 1)
    return idx, d1, d2, acc


def distribute(xys, minX, maxX, minY, maxY, N, tie_mode=0):
    xys = np.ascontiguousarray(xys, np.int32).reshape(-1, 3)
    out = np.empty(max(1, len(xys)), np.int32)
    n = lib().orb_oracle_distribute(_ptr(xys), len(xys), minX, maxX, minY, maxY, N, tie_mode, _ptr(out), len(out))
    return out[:n].copy()


# ---- extractor -------------------------------------------------------------------------------
class Extractor:
    """Mirror of ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST)."""

    def __init__(self, nfeatures=1000, scale_factor=1.2, nlevels=8, ini_th=20, min_th=7):
        self.nlevels = nlevels
        self.nfeatures = nfeatures
        self._h = lib().orb_oracle_extractor_create(nfeatures, scale_factor, nlevels, ini_th, min_th)
        if not self._h:
            raise ValueError("bad extractor parameters")

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_oracle_extractor_destroy(self._h)
            self._h = None

    def set_tie_mode(self, m):
        lib().orb_oracle_extractor_set_tie_mode(self._h, m)

    def tables(self):
        n = self.nlevels
        sc, isc, s2, is2 = (np.empty(n, np.float32) for _ in range(4))
        feats = np.empty(n, np.int32)
        umax = np.empty(16, np.int32)
        lib().orb_oracle_extractor_tables(self._h, _ptr(sc), _ptr(isc), _ptr(s2), _ptr(is2), _ptr(feats), _ptr(umax))
        return dict(scale=sc, inv_scale=isc, sigma2=s2, inv_sigma2=is2, feats=feats, umax=umax)

    def __call__(self, img, cap=None):
        img = _img(img)
        cap = cap or (self.nfeatures * 2 + 4096)
        kps = np.empty(cap, KP_DTYPE)
        desc = np.empty((cap, 32), np.uint8)
        n = lib().orb_oracle_extract(self._h, _ptr(img), img.shape[1], img.shape[0], img.strides[0],
                                     _ptr(kps), _ptr(desc), cap)
        if n < 0:
            return self(img, cap=-n)
        return kps[:n].copy(), desc[:n].copy()

    def level(self, l, blurred=False):
        w, h = C.c_int(), C.c_int()
        lib().orb_oracle_level_size(self._h, l, C.byref(w), C.byref(h))
        out = np.empty((h.value, w.value), np.uint8)
        if blurred:
            lib().orb_oracle_get_blurred_level(self._h, l, _ptr(out), w.value)
        else:
            lib().orb_oracle_get_level(self._h, l, _ptr(out), w.value)
        return out

    def candidates(self, l):
        n = lib().orb_oracle_level_candidates(self._h, l, None, 0)
        out = np.empty((max(n, 1), 3), np.int32)
        lib().orb_oracle_level_candidates(self._h, l, _ptr(out), n)
        return out[:n]

    def kept(self, l):
        return lib().orb_oracle_level_kept(self._h, l)

    def tie_sensitive(self, l):
        return lib().orb_oracle_level_tie_sensitive(self._h, l)


# ---- windowed matchers / stereo ------------------------------------------------------------------
def _opt(a):
    return None if a is None else _ptr(a)


class Grid:
    """Frame grid + GetFeaturesInArea (R/src/Frame.cpp:244-260, 387-452)."""

    def __init__(self, kps_un, bounds):
        self._k = np.ascontiguousarray(kps_un)
        self._h = lib().orb_oracle_grid_create(_ptr(self._k), len(self._k), *[float(b) for b in bounds])

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_oracle_grid_destroy(self._h)
            self._h = None

    def query(self, x, y, r, min_level, max_level):
        out = np.empty(max(len(self._k), 1), np.int32)
        n = lib().orb_oracle_grid_query(self._h, x, y, r, min_level, max_level, _ptr(out), len(out))
        return out[:n].copy()


def search_for_initialization(kps1, desc1, kps2, desc2, bounds, prev_matched, window, nnratio, check_ori=True):
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    desc1, desc2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    pm = np.ascontiguousarray(prev_matched, np.float32).copy()
    m12 = np.empty(len(kps1), np.int32)
    n = lib().orb_oracle_search_for_initialization(_ptr(kps1), _ptr(desc1), len(kps1), _ptr(kps2), _ptr(desc2), len(kps2),
                                                   *[float(b) for b in bounds], _ptr(pm), _ptr(m12), window, nnratio,
                                                   int(check_ori))
    return n, m12, pm


def search_by_projection_mp(kps_un, desc, uright, has_mp, bounds, mp_uv, mp_r, mp_level, mp_ur, mp_desc, scale_factors,
                            nnratio, mp_has_obs=None):
    kps_un = np.ascontiguousarray(kps_un)
    desc = np.ascontiguousarray(desc, np.uint8)
    args = [np.ascontiguousarray(mp_uv, np.float32), np.ascontiguousarray(mp_r, np.float32),
            np.ascontiguousarray(mp_level, np.int32), np.ascontiguousarray(mp_ur, np.float32),
            np.ascontiguousarray(mp_desc, np.uint8)]
    has_mp = np.ascontiguousarray(has_mp, np.uint8)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    match = np.empty(len(args[1]), np.int32)
    n = lib().orb_oracle_search_by_projection_mp(_ptr(kps_un), _ptr(desc), len(kps_un), _opt(ur), _ptr(has_mp),
                                                 *[float(b) for b in bounds], *[_ptr(a) for a in args], len(args[1]),
                                                 _ptr(sf), nnratio, _ptr(match),
                                                 _opt(None if mp_has_obs is None else np.ascontiguousarray(mp_has_obs, np.uint8)))
    return n, match


def search_by_projection_frame(kps_un, desc, uright, has_mp, bounds, q_uv, q_r, q_min_level, q_max_level, q_ur, q_angle,
                               q_desc, th_dist, check_ori=True, q_has_obs=None):
    kps_un = np.ascontiguousarray(kps_un)
    desc = np.ascontiguousarray(desc, np.uint8)
    args = [np.ascontiguousarray(q_uv, np.float32), np.ascontiguousarray(q_r, np.float32),
            np.ascontiguousarray(q_min_level, np.int32), np.ascontiguousarray(q_max_level, np.int32),
            np.ascontiguousarray(q_ur, np.float32), np.ascontiguousarray(q_angle, np.float32),
            np.ascontiguousarray(q_desc, np.uint8)]
    has_mp = np.ascontiguousarray(has_mp, np.uint8)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    match = np.empty(len(kps_un), np.int32)
    n = lib().orb_oracle_search_by_projection_frame(_ptr(kps_un), _ptr(desc), len(kps_un), _opt(ur), _ptr(has_mp),
                                                    *[float(b) for b in bounds], *[_ptr(a) for a in args], len(args[1]),
                                                    th_dist, int(check_ori), _ptr(match),
                                                    _opt(None if q_has_obs is None else np.ascontiguousarray(q_has_obs, np.uint8)))
    return n, match


def stereo_matches(kps_l, desc_l, kps_r, desc_r, pyr_l, pyr_r, scale_factors, inv_scale_factors, mbf, mb):
    kps_l, kps_r = np.ascontiguousarray(kps_l), np.ascontiguousarray(kps_r)
    desc_l, desc_r = np.ascontiguousarray(desc_l, np.uint8), np.ascontiguousarray(desc_r, np.uint8)
    pl = [np.ascontiguousarray(p, np.uint8) for p in pyr_l]
    pr = [np.ascontiguousarray(p, np.uint8) for p in pyr_r]
    nlev = len(pl)
    PL = (C.c_void_p * nlev)(*[p.ctypes.data for p in pl])
    PR = (C.c_void_p * nlev)(*[p.ctypes.data for p in pr])
    w = np.array([p.shape[1] for p in pl], np.int32)
    h = np.array([p.shape[0] for p in pl], np.int32)
    st = np.array([p.strides[0] for p in pl], np.int32)
    sf = np.ascontiguousarray(scale_factors, np.float32)
    isf = np.ascontiguousarray(inv_scale_factors, np.float32)
    u = np.empty(max(len(kps_l), 1), np.float32)
    z = np.empty(max(len(kps_l), 1), np.float32)
    n = lib().orb_oracle_stereo_matches(_ptr(kps_l), _ptr(desc_l), len(kps_l), _ptr(kps_r), _ptr(desc_r), len(kps_r),
                                        PL, PR, _ptr(w), _ptr(h), _ptr(st), nlev, _ptr(sf), _ptr(isf), mbf, mb,
                                        _ptr(u), _ptr(z))
    return u[:len(kps_l)], z[:len(kps_l)], n


# ---- bag-of-words restricted searches, distinctive descriptors -----------------------
def featvec(node):
    """Flatten what DBoW2::FeatureVector::addFeature builds when features are added in index order: node[i] = vocabulary
    node of feature i (< 0: none). Returns (node ids ascending, CSR offsets, feature indices) as int32 arrays."""
    node = np.asarray(node, np.int64)
    keep = np.nonzero(node >= 0)[0]
    order = keep[np.argsort(node[keep], kind="stable")]
    ids, counts = np.unique(node[keep], return_counts=True)
    off = np.zeros(len(ids) + 1, np.int32)
    off[1:] = np.cumsum(counts)
    return ids.astype(np.int32), off, order.astype(np.int32)


def common_blocks(fv_a, fv_b):
    """The nodes both feature vectors hold, as the (a_off, a_idx, b_off, b_idx) CSR orbx_match_blocks takes."""
    ids_a, off_a, idx_a = fv_a
    ids_b, off_b, idx_b = fv_b
    common, ia, ib = np.intersect1d(ids_a, ids_b, return_indices=True)
    a_off, b_off, a_idx, b_idx = [0], [0], [], []
    for x, y in zip(ia, ib):
        a_idx.append(idx_a[off_a[x]:off_a[x + 1]])
        b_idx.append(idx_b[off_b[y]:off_b[y + 1]])
        a_off.append(a_off[-1] + len(a_idx[-1]))
        b_off.append(b_off[-1] + len(b_idx[-1]))
    cat = lambda v: np.concatenate(v).astype(np.int32) if v else np.zeros(0, np.int32)
    return np.array(a_off, np.int32), cat(a_idx), np.array(b_off, np.int32), cat(b_idx)


def block_distances(desc_a, desc_b, a_off, a_idx, b_off, b_idx):
    desc_a, desc_b = np.ascontiguousarray(desc_a, np.uint8), np.ascontiguousarray(desc_b, np.uint8)
    arrs = [np.ascontiguousarray(a, np.int32) for a in (a_off, a_idx, b_off, b_idx)]
    nblocks = len(arrs[0]) - 1
    total = int(np.sum(np.diff(arrs[0]).astype(np.int64) * np.diff(arrs[2]).astype(np.int64)))
    off = np.zeros(nblocks + 1, np.int64)
    dist = np.empty(max(total, 1), np.uint16)
    lib().orb_oracle_block_distances(_ptr(desc_a), _ptr(desc_b), *[_ptr(a) for a in arrs], nblocks, _ptr(off), _ptr(dist))
    return off, dist[:total]


def search_by_bow_kf_frame(desc_kf, angle_kf, kf_usable, fv_kf, desc_f, angle_f, fv_f, nnratio, check_ori=True):
    desc_kf, desc_f = np.ascontiguousarray(desc_kf, np.uint8), np.ascontiguousarray(desc_f, np.uint8)
    angle_kf, angle_f = np.ascontiguousarray(angle_kf, np.float32), np.ascontiguousarray(angle_f, np.float32)
    usable = np.ascontiguousarray(kf_usable, np.uint8)
    match = np.empty(max(len(desc_f), 1), np.int32)
    n = lib().orb_oracle_search_by_bow_kf_frame(_ptr(desc_kf), _ptr(angle_kf), _ptr(usable), *[_ptr(a) for a in fv_kf],
                                                len(fv_kf[0]), _ptr(desc_f), _ptr(angle_f), len(desc_f),
                                                *[_ptr(a) for a in fv_f], len(fv_f[0]), nnratio, int(check_ori), _ptr(match))
    return n, match[:len(desc_f)]


def search_by_bow_kf_kf(desc1, angle1, usable1, fv1, desc2, angle2, usable2, fv2, nnratio, check_ori=True):
    desc1, desc2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    angle1, angle2 = np.ascontiguousarray(angle1, np.float32), np.ascontiguousarray(angle2, np.float32)
    u1, u2 = np.ascontiguousarray(usable1, np.uint8), np.ascontiguousarray(usable2, np.uint8)
    match = np.empty(max(len(desc1), 1), np.int32)
    n = lib().orb_oracle_search_by_bow_kf_kf(_ptr(desc1), _ptr(angle1), _ptr(u1), len(desc1), *[_ptr(a) for a in fv1],
                                             len(fv1[0]), _ptr(desc2), _ptr(angle2), _ptr(u2), len(desc2),
                                             *[_ptr(a) for a in fv2], len(fv2[0]), nnratio, int(check_ori), _ptr(match))
    return n, match[:len(desc1)]


def search_for_triangulation(kps1, desc1, has_mp1, uright1, fv1, kps2, desc2, has_mp2, uright2, fv2, F12, ex, ey,
                             scale_factors2, level_sigma2_2, only_stereo, check_ori=True):
    kps1, kps2 = np.ascontiguousarray(kps1), np.ascontiguousarray(kps2)
    desc1, desc2 = np.ascontiguousarray(desc1, np.uint8), np.ascontiguousarray(desc2, np.uint8)
    h1, h2 = np.ascontiguousarray(has_mp1, np.uint8), np.ascontiguousarray(has_mp2, np.uint8)
    u1, u2 = np.ascontiguousarray(uright1, np.float32), np.ascontiguousarray(uright2, np.float32)
    F = np.ascontiguousarray(F12, np.float32).reshape(9)
    sf, ls = np.ascontiguousarray(scale_factors2, np.float32), np.ascontiguousarray(level_sigma2_2, np.float32)
    match = np.empty(max(len(kps1), 1), np.int32)
    n = lib().orb_oracle_search_for_triangulation(_ptr(kps1), _ptr(desc1), _ptr(h1), _ptr(u1), len(kps1),
                                                  *[_ptr(a) for a in fv1], len(fv1[0]), _ptr(kps2), _ptr(desc2), _ptr(h2),
                                                  _ptr(u2), len(kps2), *[_ptr(a) for a in fv2], len(fv2[0]), _ptr(F),
                                                  float(ex), float(ey), _ptr(sf), _ptr(ls), int(only_stereo), int(check_ori),
                                                  _ptr(match))
    return n, match[:len(kps1)]


def distinctive(desc, obs_off):
    desc = np.ascontiguousarray(desc, np.uint8).reshape(-1, 32)
    obs_off = np.ascontiguousarray(obs_off, np.int32)
    npoints = len(obs_off) - 1
    best, med = np.empty(max(npoints, 1), np.int32), np.empty(max(npoints, 1), np.int32)
    lib().orb_oracle_distinctive(_ptr(desc), _ptr(obs_off), npoints, _ptr(best), _ptr(med))
    return best[:npoints], med[:npoints]


def search_best_in_window(kps_un, desc, uright, bounds, inv_level_sigma2, q_uv, q_ur, q_radius, q_level, q_desc, taken, th):
    """Windowed best-match core of Fuse x2 / SearchByProjection(KeyFrame*, Scw) / SearchBySim3. `taken` (uint8 array or
    None) is updated in place. Returns (best_idx, best_dist)."""
    kps_un = np.ascontiguousarray(kps_un)
    desc = np.ascontiguousarray(desc, np.uint8)
    ur = None if uright is None else np.ascontiguousarray(uright, np.float32)
    ils = None if inv_level_sigma2 is None else np.ascontiguousarray(inv_level_sigma2, np.float32)
    q_uv = np.ascontiguousarray(q_uv, np.float32).reshape(-1, 2)
    nq = len(q_uv)
    q_ur = np.zeros(nq, np.float32) if q_ur is None else np.ascontiguousarray(q_ur, np.float32)
    q_radius, q_level = np.ascontiguousarray(q_radius, np.float32), np.ascontiguousarray(q_level, np.int32)
    q_desc = np.ascontiguousarray(q_desc, np.uint8)
    assert taken is None or (taken.dtype == np.uint8 and taken.flags.c_contiguous and len(taken) == len(kps_un))
    bi, bd = np.empty(max(nq, 1), np.int32), np.empty(max(nq, 1), np.int32)
    lib().orb_oracle_search_best_in_window(_ptr(kps_un), _ptr(desc), len(kps_un), _opt(ur), *[float(b) for b in bounds], _opt(ils),
                                           _ptr(q_uv), _ptr(q_ur), _ptr(q_radius), _ptr(q_level), _ptr(q_desc), nq,
                                           _opt(taken), th, _ptr(bi), _ptr(bd))
    return bi[:nq], bd[:nq]


def undistort_points(xy, K4, dist):
    """cv::undistortPoints(xy, K, dist, None, K): (n, 2) float32 in and out."""
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    K4 = np.ascontiguousarray(K4, np.float32)
    dist = np.ascontiguousarray(dist, np.float32)
    out = np.empty_like(xy)
    lib().orb_oracle_undistort_points(_ptr(xy), len(xy), _ptr(K4), _ptr(dist), len(dist), _ptr(out))
    return out


class Vocabulary:
    """Restatement of the vendored DBoW2 vocabulary: loadFromTextFile + transform (oracle/bow_oracle.cpp)."""

    def __init__(self, path):
        self._h = lib().orb_oracle_voc_load(str(path).encode())
        assert self._h, "not a vocabulary text file"

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orb_oracle_voc_destroy(self._h)
            self._h = None

    def words(self):
        return lib().orb_oracle_voc_words(self._h)

    def export(self):
        n = lib().orb_oracle_voc_nodes(self._h)
        parent, desc = np.empty(n, np.int32), np.empty((n, 32), np.uint8)
        weight, word, klsw = np.empty(n, np.float64), np.empty(n, np.int32), np.empty(4, np.int32)
        lib().orb_oracle_voc_export(self._h, _ptr(parent), _ptr(desc), _ptr(weight), _ptr(word), _ptr(klsw))
        return parent, desc, weight, word, klsw

    def transform(self, desc, levelsup):
        import ref_voc_lib
        return ref_voc_lib.flat_transform(lib().orb_oracle_voc_transform, self._h, desc, levelsup)
