"""ORBmatcher's windowed searches end to end: device candidates + distances (orbx_match_window) and the
order-dependent host resolve (orb-slam2-_b200/host/orbx_search.cc, C++), against the oracle's restatement of
R/src/ORBmatcher.cpp SearchForInitialization :499-617 and SearchByProjection :63-163 / :1564-1718."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LIB = Path(__file__).resolve().parents[1] / "orb-slam2-_b200" / "host" / "libmirror_test.so"
P = C.c_void_p


@pytest.fixture(scope="module")
def mirror():
    L = C.CDLL(str(LIB))
    L.mirror_search_for_initialization.argtypes = [P, P, P, C.c_int, P, P, C.c_int, P, P, P, C.c_int, C.c_float, C.c_int]
    L.mirror_search_by_projection_mp.argtypes = [P, P, P, C.c_int, P, P, P, P, P, P, P, P, C.c_int, P, C.c_float, P, P]
    L.mirror_search_by_projection_frame.argtypes = [P, P, P, C.c_int, P, P, P, P, P, P, P, P, P, P, C.c_int, C.c_int,
                                                   C.c_int, P, P]
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(P)


@pytest.fixture(scope="module")
def two_frames(pkg):
    w, h = 640, 480
    img1 = pkg.synth.frame(w, h, 70)
    img2 = np.roll(img1, (3, 6), (0, 1))
    ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k1, d1 = ex(img1)
    k2, d2 = ex(img2)
    return k1, d1, k2, d2, np.array([0, w, 0, h], np.float32), ex.GetScaleFactors()


@pytest.mark.parametrize("check_ori", [1, 0])
def test_search_for_initialization(pkg, oracle, mirror, two_frames, check_ori):
    k1, d1, k2, d2, b, _ = two_frames
    m = pkg.ORBmatcher(0.9)
    pm = np.ascontiguousarray(np.stack([k1["x"], k1["y"]], 1), np.float32)
    en, em12, epm = oracle.search_for_initialization(k1, d1, k2, d2, b, pm, 100, np.float32(0.9), bool(check_ori))
    gpm = pm.copy()
    gm12 = np.empty(len(k1), np.int32)
    gn = mirror.mirror_search_for_initialization(m._h, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(b), _p(gpm),
                                                 _p(gm12), 100, np.float32(0.9), check_ori)
    assert en > 100
    assert gn == en and np.array_equal(gm12, em12) and np.array_equal(gpm.view(np.uint32), epm.view(np.uint32))
    # second iteration with the updated vbPrevMatched (Tracking calls it once per frame while initialising)
    en2, em2, _ = oracle.search_for_initialization(k1, d1, k2, d2, b, epm, 100, np.float32(0.9), bool(check_ori))
    gn2 = mirror.mirror_search_for_initialization(m._h, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(b), _p(gpm),
                                                  _p(gm12), 100, np.float32(0.9), check_ori)
    assert gn2 == en2 and np.array_equal(gm12, em2)


@pytest.mark.parametrize("stereo,some_without_obs", [(False, False), (True, False), (True, True)])
def test_search_by_projection_map_points(pkg, oracle, mirror, two_frames, stereo, some_without_obs):
    k1, d1, k2, d2, b, sf = two_frames
    rng = np.random.default_rng(5)
    nmp = len(k1)
    uv = np.stack([k1["x"] + 6, k1["y"] + 3], 1).astype(np.float32) + rng.normal(0, 1.5, (nmp, 2)).astype(np.float32)
    r = np.where(rng.random(nmp) < 0.5, 2.5, 4.0).astype(np.float32) * np.float32(3.0)   # RadiusByViewingCos * th
    level = np.clip(k1["octave"] + rng.integers(-1, 2, nmp), 0, 7).astype(np.int32)
    ur = (uv[:, 0] - rng.uniform(1, 40, nmp)).astype(np.float32)
    has_mp = (rng.random(len(k2)) < 0.1).astype(np.uint8)
    uright = None
    if stereo:
        uright = np.where(rng.random(len(k2)) < 0.6, k2["x"] - rng.uniform(1, 40, len(k2)), -1).astype(np.float32)
    obs = (rng.random(nmp) < 0.7).astype(np.uint8) if some_without_obs else None   # Observations()>0 per map point
    en, ematch = oracle.search_by_projection_mp(k2, d2, uright, has_mp, b, uv, r, level, ur, d1, sf, np.float32(0.8), obs)
    m = pkg.ORBmatcher(0.8)
    gmatch = np.empty(nmp, np.int32)
    gh = has_mp.copy()
    gn = mirror.mirror_search_by_projection_mp(m._h, _p(k2), _p(d2), len(k2), _p(uright), _p(gh), _p(b), _p(uv), _p(r),
                                               _p(level), _p(ur), _p(np.ascontiguousarray(d1)), nmp,
                                               _p(np.ascontiguousarray(sf, np.float32)), np.float32(0.8), _p(gmatch), _p(obs))
    assert en > 100
    assert gn == en and np.array_equal(gmatch, ematch)


@pytest.mark.parametrize("stereo,check_ori,some_without_obs", [(False, 1, False), (True, 1, False), (False, 0, False),
                                                               (True, 1, True)])
def test_search_by_projection_frame(pkg, oracle, mirror, two_frames, stereo, check_ori, some_without_obs):
    k1, d1, k2, d2, b, sf = two_frames
    rng = np.random.default_rng(6)
    nq = len(k1)
    uv = np.stack([k1["x"] + 6, k1["y"] + 3], 1).astype(np.float32)
    th = np.float32(15.0 if not stereo else 7.0)
    r = (th * sf[k1["octave"]]).astype(np.float32)
    mode = rng.integers(0, 3, nq)      # forward / backward / neither (:1629-1634)
    lo = np.where(mode == 0, k1["octave"], np.where(mode == 1, 0, k1["octave"] - 1)).astype(np.int32)
    hi = np.where(mode == 0, -1, np.where(mode == 1, k1["octave"], k1["octave"] + 1)).astype(np.int32)
    ur = (uv[:, 0] - rng.uniform(1, 40, nq)).astype(np.float32)
    has_mp = (rng.random(len(k2)) < 0.05).astype(np.uint8)
    uright = None
    if stereo:
        uright = np.where(rng.random(len(k2)) < 0.6, k2["x"] - rng.uniform(1, 40, len(k2)), -1).astype(np.float32)
    angle = np.ascontiguousarray(k1["angle"], np.float32)
    obs = (rng.random(nq) < 0.6).astype(np.uint8) if some_without_obs else None   # temporal points have none
    en, ematch = oracle.search_by_projection_frame(k2, d2, uright, has_mp, b, uv, r, lo, hi, ur, angle, d1, 100,
                                                   bool(check_ori), obs)
    m = pkg.ORBmatcher(0.9)
    gmatch = np.empty(len(k2), np.int32)
    gh = has_mp.copy()
    gn = mirror.mirror_search_by_projection_frame(m._h, _p(k2), _p(d2), len(k2), _p(uright), _p(gh), _p(b), _p(uv), _p(r),
                                                  _p(lo), _p(hi), _p(ur), _p(angle), _p(np.ascontiguousarray(d1)), nq,
                                                  100, check_ori, _p(gmatch), _p(obs))
    assert en > 100
    assert gn == en and np.array_equal(gmatch, ematch)
