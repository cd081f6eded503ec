"""The CUDA path against the REFERENCE'S OWN Frame / ORBmatcher code (oracle/_ref/libref_front.so: R/src/Frame.cpp,
ORBmatcher.cpp, MapPoint.cpp compiled where they lie against the cv shim), without the oracle in between:
extraction inside Frame::Frame, Frame::ComputeStereoMatches, ORBmatcher::SearchForInitialization and
ORBmatcher::SearchByProjection(CurrentFrame, LastFrame)."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

import ref_front_lib as rf
from test_oracle_vs_ref_front import BASELINE, FX, _frame_pair_with_points, u32

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not rf.available(), reason="oracle/_ref/libref_front.so not built")]

LIB = Path(__file__).resolve().parents[1] / "orb-slam2-_b200" / "host" / "libmirror_test.so"
P = C.c_void_p


def _p(a):
    return None if a is None else a.ctypes.data_as(P)


@pytest.fixture(scope="module")
def mirror():
    L = C.CDLL(str(LIB))
    L.mirror_search_for_initialization.argtypes = [P, P, P, C.c_int, P, P, C.c_int, P, P, P, C.c_int, C.c_float, C.c_int]
    L.mirror_search_by_projection_frame.argtypes = [P, P, P, C.c_int, P, P, P, P, P, P, P, P, P, P, C.c_int, C.c_int,
                                                   C.c_int, P, P]
    return L


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 5), (1241, 376, 2000, 6)])
def test_stereo_frame_vs_reference(pkg, w, h, nf, seed):
    """Frame::Frame(imLeft, imRight, ...): both extractions and ComputeStereoMatches (which runs with mb == 0, see
    tests/test_oracle_vs_ref_front.py) against orbx_extract + orbx_stereo_match."""
    left, right = pkg.synth.stereo_pair(w, h, seed)
    front = rf.Front(nf, fx=FX, fy=FX, cx=w / 2, cy=h / 2, bf=FX * BASELINE)
    F = front.frame_stereo(left, right)
    exl = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl(left)
    kr, dr = exr(right)
    assert np.array_equal(kl.view(np.uint8), F.keys.view(np.uint8)) and np.array_equal(dl, F.desc)
    assert np.array_equal(kr.view(np.uint8), F.keys_right.view(np.uint8)) and np.array_equal(dr, F.desc_right)
    u, z, n = pkg.ORBmatcher().stereo_match(exl, exr, kl, dl, kr, dr, np.float32(FX * BASELINE), np.float32(0.0))
    assert n == int((F.u_right >= 0).sum()) > 100
    assert np.array_equal(u32(u), u32(F.u_right)) and np.array_equal(u32(z), u32(F.depth))


@pytest.mark.parametrize("check_ori", [1, 0])
def test_search_for_initialization_vs_reference(pkg, mirror, check_ori):
    w, h = 640, 480
    img1 = pkg.synth.frame(w, h, 70)
    img2 = np.roll(img1, (3, 6), (0, 1))
    front = rf.Front(2000, fx=500.0, fy=500.0, cx=320.0, cy=240.0, bf=40.0)
    f1, f2 = front.frame_mono(img1), front.frame_mono(img2)
    ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k1, d1 = ex(img1)
    k2, d2 = ex(img2)
    assert np.array_equal(k1.view(np.uint8), f1.keys_un.view(np.uint8)) and np.array_equal(k2.view(np.uint8), f2.keys_un.view(np.uint8))
    pm = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    rn, rm, rpm = front.search_for_initialization(f1, f2, pm, 100, 0.9, bool(check_ori))
    m = pkg.ORBmatcher(0.9)
    gpm, gm12 = pm.copy(), np.empty(len(k1), np.int32)
    b = np.array([0, w, 0, h], np.float32)
    gn = mirror.mirror_search_for_initialization(m._h, _p(k1), _p(d1), len(k1), _p(k2), _p(d2), len(k2), _p(b), _p(gpm),
                                                 _p(gm12), 100, np.float32(0.9), check_ori)
    assert rn > 100 and gn == rn and np.array_equal(gm12, rm) and np.array_equal(u32(gpm), u32(rpm))


@pytest.mark.parametrize("kind", ["mono", "forward", "backward"])
def test_search_by_projection_last_frame_vs_reference(pkg, mirror, kind):
    rng = np.random.default_rng(11)
    front, last, cur, mp_of, has_obs = _frame_pair_with_points(kind, rng)
    mono = kind == "mono"
    th = 15.0 if mono else 7.0
    valid, u, v, invz, fwd, bwd = front.project_last_frame(cur, last, mono)
    rn = front.search_by_projection_frame(cur, last, th, mono, 0.9, True)
    got = cur.map_points()
    got_last_idx = np.array([mp_of[m] if m >= 0 else -1 for m in got], np.int32)
    ex = pkg.ORBextractor(1500, 1.2, 8, 20, 7, max_width=800, max_height=480)
    sf = ex.GetScaleFactors()
    q = np.nonzero(valid)[0]
    octv = last.keys["octave"][q].astype(np.int32)
    q_r = (np.float32(th) * sf[octv]).astype(np.float32)
    if fwd:
        lo, hi = octv, np.full(len(q), -1, np.int32)
    elif bwd:
        lo, hi = np.zeros(len(q), np.int32), octv
    else:
        lo, hi = octv - 1, octv + 1
    lo, hi = np.ascontiguousarray(lo, np.int32), np.ascontiguousarray(hi, np.int32)
    q_uv = np.ascontiguousarray(np.stack([u[q], v[q]], 1), np.float32)
    q_ur = (u[q] - front.bf * invz[q]).astype(np.float32)
    q_angle = np.ascontiguousarray(last.keys_un["angle"][q], np.float32)
    q_desc = np.ascontiguousarray(last.desc[q])
    q_obs = np.ascontiguousarray(has_obs[q])
    m = pkg.ORBmatcher(0.9)
    gmatch = np.empty(cur.N, np.int32)
    gh = np.zeros(cur.N, np.uint8)
    gn = mirror.mirror_search_by_projection_frame(m._h, _p(cur.keys_un), _p(cur.desc), cur.N, _p(cur.u_right), _p(gh),
                                                  _p(cur.bounds), _p(q_uv), _p(q_r), _p(lo), _p(hi), _p(q_ur), _p(q_angle),
                                                  _p(q_desc), len(q), 100, 1, _p(gmatch), _p(q_obs))
    expect = np.where(gmatch >= 0, q[np.maximum(gmatch, 0)], -1).astype(np.int32)
    assert rn > 100 and gn == rn and np.array_equal(expect, got_last_idx)
