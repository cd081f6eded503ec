"""SURVEY.md 8f rows 1-2 on the device: node-restricted distance blocks (orbx_match_blocks) with the C++ host resolves of
ORBmatcher::SearchByBoW x2 / SearchForTriangulation (orb-slam2-_b200/host/orbx_search.cc), and
MapPoint::ComputeDistinctiveDescriptors for many points (orbx_distinctive) -- against the oracle and, where
oracle/_ref/libref_front.so travelled to this box, against the reference's own ORBmatcher.cpp / MapPoint.cpp /
KeyFrame.cpp run on real KeyFrame and MapPoint objects."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

import bow_case
import ref_front_lib as rf

pytestmark = pytest.mark.gpu

LIB = Path(__file__).resolve().parents[1] / "orb-slam2-_b200" / "host" / "libmirror_test.so"
P, I, F = C.c_void_p, C.c_int, C.c_float


@pytest.fixture(scope="module")
def mirror():
    L = C.CDLL(str(LIB))
    L.mirror_search_by_bow_kf_frame.argtypes = [P, P, P, I, P, P, P, I, P, P, P, I, P, P, P, I, F, I, P]
    L.mirror_search_by_bow_kf_kf.argtypes = [P, P, P, I, P, P, P, I, P, P, P, I, P, P, P, I, P, F, I, P]
    L.mirror_search_for_triangulation.argtypes = [P, P, P, I, P, P, P, I, P, P, P, P, I, P, P, P, I, P, P, P, F, F, P, P, I, I, P]
    L.mirror_search_best_in_window.argtypes = [P, P, P, I, P, P, P, P, P, P, P, P, I, P, I, P, P]
    L.mirror_search_best_in_window.restype = None
    return L


def _p(a):
    return None if a is None else np.ascontiguousarray(a).ctypes.data_as(P)


def _u8(mask):
    return np.ascontiguousarray(mask, np.uint8)


@pytest.fixture(scope="module", params=["mono", "stereo"])
def case(request, pkg):
    stereo = request.param == "stereo"
    return bow_case.build(rf, stereo) if rf.available() else bow_case.build_arrays(pkg, stereo)


def test_match_blocks_vs_oracle(pkg, oracle):
    rng = np.random.default_rng(3)
    m = pkg.ORBmatcher(0.6)
    da, db = pkg.synth.descriptors(700, 11), pkg.synth.descriptors(900, 12)
    sizes_a = rng.integers(0, 40, 60)
    sizes_b = rng.integers(0, 70, 60)
    sizes_a[5] = 0
    sizes_b[9] = 0
    sizes_b[20] = 333   # more than one lane pass per row
    a_off = np.concatenate([[0], np.cumsum(sizes_a)]).astype(np.int32)
    b_off = np.concatenate([[0], np.cumsum(sizes_b)]).astype(np.int32)
    a_idx = rng.integers(0, len(da), a_off[-1]).astype(np.int32)
    b_idx = rng.integers(0, len(db), b_off[-1]).astype(np.int32)
    goff, gd = m.match_blocks(da, db, a_off, a_idx, b_off, b_idx)
    eoff, ed = oracle.block_distances(da, db, a_off, a_idx, b_off, b_idx)
    assert np.array_equal(goff, eoff) and np.array_equal(gd, ed) and len(gd) > 30000
    # spot check against the bit count
    k = 20
    blk = gd[goff[k]:goff[k + 1]].reshape(sizes_a[k], sizes_b[k])
    ref = np.unpackbits(da[a_idx[a_off[k]:a_off[k + 1]]][:, None, :] ^ db[b_idx[b_off[k]:b_off[k + 1]]][None, :, :], axis=2).sum(2)
    assert np.array_equal(blk, ref)
    # no blocks / only empty blocks
    z = np.zeros(1, np.int32)
    off0, d0 = m.match_blocks(da, db, z, np.zeros(0, np.int32), z, np.zeros(0, np.int32))
    assert off0.tolist() == [0] and len(d0) == 0
    off1, d1 = m.match_blocks(da, db, np.array([0, 3], np.int32), np.array([1, 2, 3], np.int32), np.array([0, 0], np.int32),
                              np.zeros(0, np.int32))
    assert off1.tolist() == [0, 0] and len(d1) == 0
    with pytest.raises(pkg.OrbxError):   # index outside the descriptor array
        m.match_blocks(da, db, np.array([0, 1], np.int32), np.array([len(da)], np.int32), np.array([0, 1], np.int32),
                       np.array([0], np.int32))


def test_distinctive_vs_oracle(pkg, oracle):
    rng = np.random.default_rng(5)
    m = pkg.ORBmatcher(0.6)
    sizes = np.array([0, 1, 2, 3, 4, 5, 8, 31, 32, 33, 64, 100, 300, 0, 7, 2], np.int32)
    sizes = np.concatenate([sizes, rng.integers(1, 40, 500).astype(np.int32)])
    off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    base = pkg.synth.descriptors(len(sizes), 21)
    desc = np.repeat(base, sizes, axis=0)
    # observations of one point = its descriptor with up to 40 flipped bits (so medians tie often); some exact duplicates
    flips = rng.integers(0, 41, len(desc))
    for r in range(len(desc)):
        bits = rng.choice(256, flips[r], replace=False)
        np.bitwise_xor.at(desc[r], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    desc[off[12] + 5] = desc[off[12] + 200]
    gb, gm = m.distinctive(desc, off)
    eb, em = oracle.distinctive(desc, off)
    assert np.array_equal(gb, eb) and np.array_equal(gm, em)
    assert gb[0] == -1 and gb[13] == -1 and gb[1] == 0 and gm[1] == 0
    assert m.distinctive(np.zeros((0, 32), np.uint8), np.zeros(1, np.int32))[0].size == 0


def test_distinctive_vs_reference(pkg):
    if not rf.available():
        pytest.skip("oracle/_ref/libref_front.so not built")
    c = bow_case.build_observations(rf)
    rows = [c.front.observed_rows(mp) for mp in c.points]
    off = np.concatenate([[0], np.cumsum([len(r) for r in rows])]).astype(np.int32)
    best, _ = pkg.ORBmatcher(0.6).distinctive(np.concatenate(rows), off)
    for k, mp in enumerate(c.points):
        assert np.array_equal(c.front.distinctive(mp), rows[k][best[k]])


@pytest.mark.parametrize("check_ori", [1, 0])
@pytest.mark.parametrize("ratio", [0.7, 0.9])
def test_search_by_bow_keyframe_frame(pkg, oracle, mirror, case, check_ori, ratio):
    c = case
    m = pkg.ORBmatcher(ratio)
    ratio = np.float32(ratio)
    usable = _u8(c.mp_of1 >= 0)
    gm = np.empty(c.f2.N, np.int32)
    gn = mirror.mirror_search_by_bow_kf_frame(m._h, _p(c.f1.keys_un), _p(c.f1.desc), c.f1.N, *[_p(a) for a in c.fv1], len(c.fv1[0]),
                                              _p(usable), _p(c.f2.keys_un), _p(c.f2.desc), c.f2.N, *[_p(a) for a in c.fv2],
                                              len(c.fv2[0]), ratio, check_ori, _p(gm))
    en, em = oracle.search_by_bow_kf_frame(c.f1.desc, c.f1.keys_un["angle"], usable, c.fv1, c.f2.desc, c.f2.keys_un["angle"],
                                           c.fv2, ratio, bool(check_ori))
    assert gn == en and np.array_equal(gm, em) and gn > 40
    if c.front is not None:
        rn, rmp = c.front.search_by_bow_kf_frame(c.kf1, c.f2, ratio, bool(check_ori))
        assert rn == gn and np.array_equal(rmp, np.where(gm >= 0, c.mp_of1[np.maximum(gm, 0)], -1))


@pytest.mark.parametrize("check_ori", [1, 0])
def test_search_by_bow_keyframes(pkg, oracle, mirror, case, check_ori):
    c = case
    ratio = np.float32(0.75)
    m = pkg.ORBmatcher(ratio)
    u1, u2 = _u8(c.mp_of1 >= 0), _u8(c.mp_of2 >= 0)
    gm = np.empty(c.f1.N, np.int32)
    gn = mirror.mirror_search_by_bow_kf_kf(m._h, _p(c.f1.keys_un), _p(c.f1.desc), c.f1.N, *[_p(a) for a in c.fv1], len(c.fv1[0]),
                                           _p(u1), _p(c.f2.keys_un), _p(c.f2.desc), c.f2.N, *[_p(a) for a in c.fv2], len(c.fv2[0]),
                                           _p(u2), ratio, check_ori, _p(gm))
    en, em = oracle.search_by_bow_kf_kf(c.f1.desc, c.f1.keys_un["angle"], u1, c.fv1, c.f2.desc, c.f2.keys_un["angle"], u2, c.fv2,
                                        ratio, bool(check_ori))
    assert gn == en and np.array_equal(gm, em) and gn > 20
    if c.front is not None:
        rn, rmp = c.front.search_by_bow_kf_kf(c.kf1, c.kf2, ratio, bool(check_ori))
        assert rn == gn and np.array_equal(rmp, np.where(gm >= 0, c.mp_of2[np.maximum(gm, 0)], -1))


@pytest.mark.parametrize("only_stereo", [0, 1])
@pytest.mark.parametrize("check_ori", [0, 1])
@pytest.mark.parametrize("near_epipole", [False, True])
def test_search_for_triangulation(pkg, oracle, mirror, case, only_stereo, check_ori, near_epipole):
    c = case
    m = pkg.ORBmatcher(0.6)
    if c.front is not None:
        kf2 = c.kf2_center if near_epipole else c.kf2
        ex, ey = c.front.epipole(c.kf1, kf2)
        sf, ls = kf2.tables()
    else:
        ex, ey = bow_case.epipole_of(bow_case.T_CENTER if near_epipole else bow_case.T_FAR)
        sf, ls = c.scale_factors, c.level_sigma2
    h1, h2 = _u8(c.mp_of1 >= 0), _u8(c.mp_of2 >= 0)
    pairs = np.empty((max(c.f1.N, 1), 2), np.int32)
    gn = mirror.mirror_search_for_triangulation(m._h, _p(c.f1.keys_un), _p(c.f1.desc), c.f1.N, *[_p(a) for a in c.fv1], len(c.fv1[0]),
                                                _p(h1), _p(c.f1.u_right), _p(c.f2.keys_un), _p(c.f2.desc), c.f2.N,
                                                *[_p(a) for a in c.fv2], len(c.fv2[0]), _p(h2), _p(c.f2.u_right), _p(c.F12),
                                                np.float32(ex), np.float32(ey), _p(sf), _p(ls), only_stereo, check_ori, _p(pairs))
    en, em = oracle.search_for_triangulation(c.f1.keys_un, c.f1.desc, h1, c.f1.u_right, c.fv1, c.f2.keys_un, c.f2.desc, h2,
                                             c.f2.u_right, c.fv2, c.F12, ex, ey, sf, ls, bool(only_stereo), bool(check_ori))
    i1 = np.nonzero(em >= 0)[0]
    assert gn == en == len(i1) and np.array_equal(pairs[:gn], np.stack([i1, em[i1]], 1))
    if c.front is not None:
        rn, rpairs = c.front.search_for_triangulation(c.kf1, kf2, c.F12, bool(only_stereo), 0.6, bool(check_ori))
        assert rn == gn and np.array_equal(rpairs, pairs[:gn])


def _best_in_window(mirror, m, f, inv_sigma2, uv, ur, radius, level, desc, taken, th):
    nq = len(radius)
    bi, bd = np.empty(max(nq, 1), np.int32), np.empty(max(nq, 1), np.int32)
    uv, radius, level, desc = (np.ascontiguousarray(a) for a in (uv, radius, level, desc))
    ur = np.zeros(nq, np.float32) if ur is None else np.ascontiguousarray(ur)
    mirror.mirror_search_best_in_window(m._h, _p(f.keys_un), _p(f.desc), f.N, _p(f.u_right), _p(f.bounds), _p(inv_sigma2), _p(uv),
                                        _p(ur), _p(radius), _p(level), _p(desc), nq, None if taken is None else taken.ctypes.data_as(P),
                                        th, _p(bi), _p(bd))
    return bi[:nq], bd[:nq]


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [3.0, 8.0])
def test_fuse(pkg, oracle, mirror, stereo, th):
    """Fuse(pKF, vpMapPoints, th) (R/src/ORBmatcher.cpp:995-1160): device candidates + distances, host gates and best
    match; against the oracle and against the reference's Fuse run on the real key frame."""
    if not rf.available():
        pytest.skip("oracle/_ref/libref_front.so not built (the projection prologue comes from the reference objects)")
    c = bow_case.build_fuse(rf, stereo)
    m = pkg.ORBmatcher(0.6)
    valid, uv, ur, radius, level, desc = c.front.fuse_queries(c.kf2, c.src, th)
    ils = c.kf2.inv_level_sigma2()
    gi, gd = _best_in_window(mirror, m, c.f2, ils, uv[valid], ur[valid], radius[valid], level[valid], desc[valid], None, 50)
    ei, ed = oracle.search_best_in_window(c.f2.keys_un, c.f2.desc, c.f2.u_right, c.f2.bounds, ils, uv[valid], ur[valid],
                                          radius[valid], level[valid], desc[valid], None, 50)
    assert np.array_equal(gi, ei) and np.array_equal(gd, ed)
    best_idx, best_dist = np.full(len(c.src), -1, np.int32), np.full(len(c.src), 256, np.int32)
    best_idx[valid], best_dist[valid] = gi, gd
    n_points = int(max(c.src.max(), c.mp_of2.max())) + 1
    obs = np.array([c.front.observations(k) for k in range(n_points)])
    state = c.mp_of2.copy()
    gn = bow_case.simulate_fuse(state, obs, c.src, valid, best_idx, best_dist, c.f2.u_right >= 0)
    rn = c.front.fuse(c.kf2, c.src, th)
    assert rn == gn and rn > (20 if stereo else 300)
    assert np.array_equal(c.kf2.map_points(), state)


@pytest.mark.parametrize("stereo", [False, True])
def test_search_by_projection_keyframe_scw(pkg, oracle, mirror, stereo):
    """SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (R/src/ORBmatcher.cpp:370-497): no gates, vpMatched as the
    `taken` flags; the same core serves Fuse(pKF, Scw, ...) (:1164-1296) and SearchBySim3 (:1305-1547)."""
    if not rf.available():
        pytest.skip("oracle/_ref/libref_front.so not built")
    c = bow_case.build_fuse(rf, stereo)
    m = pkg.ORBmatcher(0.75)
    th = 10
    Scw = bow_case.pose(bow_case.T_FUSE)
    Scw[:3, :] *= np.float32(1.25)
    matched = np.full(c.f2.N, -1, np.int32)
    matched[5::9] = c.src[np.arange(len(matched[5::9])) % len(c.src)]
    valid, uv, radius, level, desc = c.front.scw_queries(c.kf2, Scw, c.src, matched, th)
    tg, te = ((matched >= 0).astype(np.uint8) for _ in range(2))
    gi, gd = _best_in_window(mirror, m, c.f2, None, uv[valid], None, radius[valid], level[valid], desc[valid], tg, 50)
    ei, ed = oracle.search_best_in_window(c.f2.keys_un, c.f2.desc, None, c.f2.bounds, None, uv[valid], None, radius[valid],
                                          level[valid], desc[valid], te, 50)
    assert np.array_equal(gi, ei) and np.array_equal(gd, ed) and np.array_equal(tg, te)
    expect = matched.copy()
    hit = gd <= 50
    expect[gi[hit]] = c.src[valid][hit]
    rn, rmatched = c.front.search_by_projection_kf_scw(c.kf2, Scw, c.src, matched, th)
    assert rn == int(hit.sum()) and rn > 200 and np.array_equal(rmatched, expect)
