"""Pins the matcher / stereo rows of the oracle against the REFERENCE'S OWN CODE:
R/src/Frame.cpp, R/src/ORBmatcher.cpp, R/src/MapPoint.cpp compiled where they lie against the cv shim
(oracle/_ref/libref_front.so, tests/ref_front_lib.py). Real Frame and MapPoint objects are built through the
reference's constructors; the reference's member functions run; the oracle gets the same plain arrays.

Skipped where the library was not built (it needs /root/reference at build time; the built .so travels)."""
import importlib

import numpy as np
import pytest

import oracle_lib as ol
import ref_front_lib as rf

pytestmark = pytest.mark.skipif(not rf.available(), reason="oracle/_ref/libref_front.so not built")
synth = importlib.import_module("orb-slam2-_b200.synth")

FX = 435.2
BASELINE = 0.11


def u32(a):
    return np.ascontiguousarray(a).view(np.uint32)


@pytest.fixture(scope="module")
def mono_pair():
    w, h = 640, 480
    img1 = synth.frame(w, h, 70)
    img2 = np.roll(img1, (3, 6), (0, 1))
    front = rf.Front(2000, fx=500.0, fy=500.0, cx=320.0, cy=240.0, bf=40.0)
    f1, f2 = front.frame_mono(img1), front.frame_mono(img2)
    return front, f1, f2, img1, img2


def test_reference_frame_equals_oracle_extraction(mono_pair):
    """Frame::Frame (mono) = ExtractORB + identity undistortion: same keypoints as the oracle, bounds = image."""
    front, f1, f2, img1, img2 = mono_pair
    ox = ol.Extractor(2000, 1.2, 8, 20, 7)
    for f, img in ((f1, img1), (f2, img2)):
        k, d = ox(img)
        assert np.array_equal(k.view(np.uint8), f.keys.view(np.uint8)) and np.array_equal(d, f.desc)
        assert np.array_equal(f.keys.view(np.uint8), f.keys_un.view(np.uint8))
    assert f1.bounds.tolist() == [0.0, 640.0, 0.0, 480.0]


@pytest.mark.parametrize("check_ori", [True, False])
@pytest.mark.parametrize("window,ratio", [(100, 0.9), (30, 0.7)])
def test_search_for_initialization_vs_reference(mono_pair, check_ori, window, ratio):
    front, f1, f2, _, _ = mono_pair
    pm = np.stack([f1.keys_un["x"], f1.keys_un["y"]], 1).astype(np.float32)
    rn, rm, rpm = front.search_for_initialization(f1, f2, pm, window, ratio, check_ori)
    en, em, epm = ol.search_for_initialization(f1.keys_un, f1.desc, f2.keys_un, f2.desc, f1.bounds, pm, window,
                                               np.float32(ratio), check_ori)
    assert rn > 50
    assert rn == en and np.array_equal(rm, em) and np.array_equal(u32(rpm), u32(epm))
    # second call with the updated vbPrevMatched, as Tracking does while initialising
    rn2, rm2, _ = front.search_for_initialization(f1, f2, rpm, window, ratio, check_ori)
    en2, em2, _ = ol.search_for_initialization(f1.keys_un, f1.desc, f2.keys_un, f2.desc, f1.bounds, epm, window,
                                               np.float32(ratio), check_ori)
    assert rn2 == en2 and np.array_equal(rm2, em2)


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 5), (1241, 376, 2000, 6)])
def test_compute_stereo_matches_vs_reference(w, h, nf, seed):
    """Frame::Frame (stereo) runs ComputeStereoMatches with the member mb still 0 (R/src/Frame.cpp:69 initialises
    mb(0), :101 calls ComputeStereoMatches, :126 sets mb = mbf/fx), so the reference's disparity limit maxD = mbf/mb
    is +inf: the oracle is called with mb = 0 and must then agree bit for bit."""
    left, right = synth.stereo_pair(w, h, seed)
    front = rf.Front(nf, fx=FX, fy=FX, cx=w / 2, cy=h / 2, bf=FX * BASELINE)
    F = front.frame_stereo(left, right)
    ox = ol.Extractor(nf, 1.2, 8, 20, 7)
    kl, dl = ox(left)
    pl = [ox.level(l) for l in range(8)]
    kr, dr = ox(right)
    pr = [ox.level(l) for l in range(8)]
    t = ox.tables()
    assert np.array_equal(kl.view(np.uint8), F.keys.view(np.uint8)) and np.array_equal(kr.view(np.uint8), F.keys_right.view(np.uint8))
    for l in range(8):
        assert np.array_equal(front.level(0, l), pl[l]) and np.array_equal(front.level(1, l), pr[l])
    assert F.mb == np.float32(FX * BASELINE) / np.float32(FX)
    eu, ez, en = ol.stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], np.float32(FX * BASELINE), np.float32(0.0))
    assert en > 100 and en == int((F.u_right >= 0).sum())
    assert np.array_equal(u32(F.u_right), u32(eu)) and np.array_equal(u32(F.depth), u32(ez))
    # with the baseline the caller might expect (mb = mbf/fx) the limit bites and the result differs
    _, _, en2 = ol.stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], np.float32(FX * BASELINE), F.mb)
    assert en2 < en


def _track_fields(rng, k1, sf):
    n = len(k1)
    uv = np.stack([k1["x"] + 6, k1["y"] + 3], 1).astype(np.float32) + rng.normal(0, 1.5, (n, 2)).astype(np.float32)
    view_cos = np.where(rng.random(n) < 0.5, 0.999, 0.9).astype(np.float32)
    level = np.clip(k1["octave"] + rng.integers(-1, 2, n), 0, len(sf) - 1).astype(np.int32)
    ur = (uv[:, 0] - rng.uniform(1, 40, n)).astype(np.float32)
    in_view = rng.random(n) < 0.9
    return uv, view_cos, level, ur, in_view


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [1.0, 3.0])
def test_search_by_projection_map_points_vs_reference(stereo, th):
    """ORBmatcher::SearchByProjection(Frame&, vector<MapPoint*>&, th) on real MapPoint objects carrying the fields
    Frame::isInFrustum leaves on them."""
    rng = np.random.default_rng(5)
    if stereo:
        w, h = 752, 480
        left, right = synth.stereo_pair(w, h, 8)
        front = rf.Front(1500, fx=FX, fy=FX, cx=w / 2, cy=h / 2, bf=FX * BASELINE)
        src = front.frame_mono(np.roll(left, (-3, -6), (0, 1)))
        F = front.frame_stereo(left, right)
        assert (F.u_right > 0).sum() > 50
    else:
        w, h = 640, 480
        img1 = synth.frame(w, h, 70)
        front = rf.Front(1500, fx=500.0, fy=500.0, cx=320.0, cy=240.0, bf=40.0)
        src = front.frame_mono(img1)
        F = front.frame_mono(np.roll(img1, (3, 6), (0, 1)))
    src.set_pose(np.eye(4))
    sf = ol.Extractor(1500, 1.2, 8, 20, 7).tables()["scale"]
    k1 = src.keys_un
    uv, view_cos, level, ur, in_view = _track_fields(rng, k1, sf)
    mps = []
    obs = (rng.random(src.N) < 0.7).astype(np.uint8)      # Observations() > 0; the others leave their keypoint available
    for i in range(src.N):
        mp = front.map_point(src, i, [0.0, 0.0, 5.0], n_obs=int(obs[i]))
        front.set_track(mp, in_view[i], uv[i, 0], uv[i, 1], ur[i], int(level[i]), view_cos[i])
        mps.append(mp)
    # a tenth of the frame's keypoints already carry a map point with observations
    has_mp = (rng.random(F.N) < 0.1)
    owner = front.map_point(src, 0, [0.0, 0.0, 5.0])
    for i in np.nonzero(has_mp)[0]:
        F.set_map_point(int(i), owner)
    rn = front.search_by_projection_mp(F, mps, th, 0.8)
    got = F.map_points()
    sel = np.nonzero(in_view)[0]
    r = np.where(view_cos[sel] > 0.998, 2.5, 4.0).astype(np.float32)
    if th != 1.0:
        r = r * np.float32(th)
    en, ematch = ol.search_by_projection_mp(F.keys_un, F.desc, F.u_right if stereo else None, has_mp.astype(np.uint8),
                                            F.bounds, uv[sel], r, level[sel], ur[sel], src.desc[sel], sf, np.float32(0.8),
                                            obs[sel])
    assert rn > 100 and rn == en
    expect = np.full(F.N, -1, np.int32)
    expect[has_mp] = owner
    for j, idx in enumerate(ematch):     # ascending map-point order: a later point overwrites an unobserved earlier one
        if idx >= 0:
            expect[idx] = mps[sel[j]]
    assert np.array_equal(got, expect)
    assert len(set(ematch[ematch >= 0].tolist())) < int((ematch >= 0).sum())   # the overwrite case really occurs


def _frame_pair_with_points(kind, rng):
    """Last frame with map points on 70 % of its keypoints (5 % of them flagged outliers), current frame displaced."""
    if kind == "mono":
        w, h = 640, 480
        img1 = synth.frame(w, h, 70)
        front = rf.Front(1500, fx=500.0, fy=500.0, cx=320.0, cy=240.0, bf=40.0)
        last, cur = front.frame_mono(img1), front.frame_mono(np.roll(img1, (3, 6), (0, 1)))
        tz = 0.0
    else:
        w, h = 752, 480
        left, right = synth.stereo_pair(w, h, 8)
        front = rf.Front(1500, fx=FX, fy=FX, cx=w / 2, cy=h / 2, bf=FX * BASELINE)
        last = front.frame_stereo(left, right)
        cur = front.frame_stereo(np.roll(left, (3, 6), (0, 1)), np.roll(right, (3, 6), (0, 1)))
        tz = {"forward": -0.3, "backward": 0.3, "still": 0.0}[kind]   # camera centre twc = -t for R = I
    last.set_pose(np.eye(4))
    z = 5.0
    T = np.eye(4, dtype=np.float32)
    T[0, 3], T[1, 3], T[2, 3] = 6 * z / float(front.fx), 3 * z / float(front.fy), tz
    cur.set_pose(T)
    mp_of = {}
    has_obs = np.zeros(last.N, np.uint8)
    for i in range(last.N):
        if rng.random() < 0.7:
            zi = z * (1 + rng.uniform(-0.05, 0.05))
            x, y = float(last.keys_un["x"][i]), float(last.keys_un["y"][i])
            pos = [(x - float(front.cx)) / float(front.fx) * zi, (y - float(front.cy)) / float(front.fy) * zi, zi]
            has_obs[i] = rng.random() < 0.6     # the temporal points of Tracking::UpdateLastFrame have no observations
            mp = front.map_point(last, i, pos, n_obs=int(has_obs[i]))
            last.set_map_point(i, mp)
            mp_of[mp] = i
            if rng.random() < 0.05:
                last.set_outlier(i, True)
    return front, last, cur, mp_of, has_obs


@pytest.mark.parametrize("kind", ["mono", "still", "forward", "backward"])
@pytest.mark.parametrize("th", [7.0, 15.0])
def test_search_by_projection_last_frame_vs_reference(kind, th):
    """ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, th, bMono): the harness exports the projections the
    reference computes in its prologue; the oracle gets them as queries."""
    rng = np.random.default_rng(11)
    front, last, cur, mp_of, has_obs = _frame_pair_with_points(kind, rng)
    mono = kind == "mono"
    valid, u, v, invz, fwd, bwd = front.project_last_frame(cur, last, mono)
    assert (fwd, bwd) == (kind == "forward", kind == "backward")
    rn = front.search_by_projection_frame(cur, last, th, mono, 0.9, True)
    got = cur.map_points()
    sf = ol.Extractor(1500, 1.2, 8, 20, 7).tables()["scale"]
    q = np.nonzero(valid)[0]
    octv = last.keys["octave"][q].astype(np.int32)
    q_r = (np.float32(th) * sf[octv]).astype(np.float32)
    if fwd:
        lo, hi = octv, np.full(len(q), -1, np.int32)
    elif bwd:
        lo, hi = np.zeros(len(q), np.int32), octv
    else:
        lo, hi = octv - 1, octv + 1
    q_ur = (u[q] - front.bf * invz[q]).astype(np.float32)
    en, ematch = ol.search_by_projection_frame(cur.keys_un, cur.desc, cur.u_right, np.zeros(cur.N, np.uint8), cur.bounds,
                                               np.stack([u[q], v[q]], 1), q_r, lo, hi, q_ur, last.keys_un["angle"][q],
                                               last.desc[q], 100, True, has_obs[q])
    assert rn > 100 and rn == en
    expect = np.full(cur.N, -1, np.int32)
    for i2, j in enumerate(ematch):
        if j >= 0:
            expect[i2] = q[j]
    got_last_idx = np.array([mp_of[m] if m >= 0 else -1 for m in got], np.int32)
    assert np.array_equal(got_last_idx, expect)
