"""The drop-in ORBmatcher (orb-slam2-_b200/host/ORBmatcher.cc over liborbx.so) against the reference's own
R/src/ORBmatcher.cpp, member by member, on REAL reference objects.

oracle/_ref/libref_front.so and oracle/_ref/libdropin_front.so hold the same harness and the same reference Frame.cpp /
KeyFrame.cpp / MapPoint.cpp / Map.cpp / ORBextractor.cpp objects; they differ in one translation unit: the first links
R/src/ORBmatcher.cpp, the second the product's ORBmatcher.cc (compiled against the reference's unchanged ORBmatcher.h).
Every scenario is built twice, once in each library, the same member function is called, and what it leaves behind
(return value, output vectors, the pointer graph: Frame::mvpMapPoints, KeyFrame map points, MapPoint observations and
bad flags) must be identical. Map points are named by creation index, which is the same in both libraries."""
import importlib

import numpy as np
import pytest

import bow_case
import ref_front_lib as rf

synth = importlib.import_module("orb-slam2-_b200.synth")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (rf.available() and rf.DROPIN_PATH.exists() and rf.DROPIN_MATCHER_PATH.exists()),
                                 reason="oracle/_ref libraries not built")]

FX, BASELINE = 435.2, 0.11
# "matcher": only R/src/ORBmatcher.cpp replaced; "front": R/src/ORBextractor.cpp and Frame::ComputeStereoMatches replaced
# as well (host/ORBextractor.cc, host/FrameStereo.cc), so the frames the searches run on were extracted and
# stereo-matched on the device
PATHS = {"matcher": rf.DROPIN_MATCHER_PATH, "front": rf.DROPIN_PATH}
_current = {"path": rf.DROPIN_MATCHER_PATH}


@pytest.fixture(autouse=True, scope="module", params=["matcher", "front"])
def dropin_variant(request):
    _current["path"] = PATHS[request.param]
    return request.param


def libs():
    return rf, rf.variant(_current["path"])


def same(a, b):
    """Results of the two libraries: equal return values and bit-equal arrays."""
    if isinstance(a, tuple):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            same(x, y)
    elif isinstance(a, np.ndarray):
        assert a.shape == b.shape and np.array_equal(a.view(np.uint8), b.view(np.uint8))
    else:
        assert a == b


def graph(front, n_points, *holders):
    """The part of the pointer graph the matcher may touch."""
    obs = np.array([front.observations(k) for k in range(n_points)], np.int32)
    bad = np.array([front.is_bad(k) for k in range(n_points)], np.uint8)
    return (obs, bad) + tuple(h.map_points() for h in holders)


def test_the_two_libraries_differ_in_the_matcher_only():
    import subprocess
    def syms(path, flag):
        out = subprocess.run(["nm", "-D", flag, str(path)], capture_output=True, text=True).stdout
        return {l.split()[-1] for l in out.splitlines() if l.strip()}
    assert not any(s.startswith("orbx_") for s in syms(rf.LIB_PATH, "-u"))
    need = syms(rf.DROPIN_MATCHER_PATH, "-u")
    assert {"orbx_matcher_create", "orbx_match_window", "orbx_match_blocks"} <= need and "orbx_extract" not in need
    need = syms(rf.DROPIN_PATH, "-u")
    assert {"orbx_matcher_create", "orbx_match_window", "orbx_match_blocks", "orbx_extract", "orbx_stereo_match"} <= need
    assert not any("ExtractorNode" in s for s in syms(rf.DROPIN_PATH, "--defined-only"))   # no reference extractor inside


def test_descriptor_distance_and_stereo_frames():
    """ORBmatcher::DescriptorDistance / TH_HIGH / TH_LOW are also what Frame::ComputeStereoMatches (R/src/Frame.cpp:551-770)
    and MapPoint::ComputeDistinctiveDescriptors call: frames built by the two libraries must agree."""
    r, d = libs()
    rng = np.random.default_rng(3)
    a, b = rng.integers(0, 256, (64, 32), dtype=np.uint8), rng.integers(0, 256, (64, 32), dtype=np.uint8)
    b[:8] = a[:8]
    fr, fd = r.Front(500), d.Front(500)
    for i in range(64):
        e = int(np.unpackbits(a[i] ^ b[i]).sum())
        assert fr.descriptor_distance(a[i], b[i]) == fd.descriptor_distance(a[i], b[i]) == e
    left, right = synth.stereo_pair(752, 480, 5)
    out = []
    for m in (r, d):
        F = m.Front(1200, fx=FX, fy=FX, cx=376, cy=240, bf=FX * BASELINE).frame_stereo(left, right)
        out.append((F.keys, F.desc, F.u_right, F.depth))
    assert (out[0][2] >= 0).sum() > 100
    same(out[0], out[1])
    c = [bow_case.build_observations(m) for m in (r, d)]
    for k in range(len(c[0].points)):
        same(c[0].front.distinctive(c[0].points[k]), c[1].front.distinctive(c[1].points[k]))


@pytest.mark.parametrize("check_ori", [True, False])
@pytest.mark.parametrize("window,ratio", [(100, 0.9), (30, 0.7)])
def test_search_for_initialization(check_ori, window, ratio):
    img = synth.frame(640, 480, 70)
    out = []
    for m in libs():
        front = m.Front(2000)
        f1, f2 = front.frame_mono(img), front.frame_mono(np.roll(img, (3, 6), (0, 1)))
        prev = np.stack([f1.keys_un["x"], f1.keys_un["y"]], 1)
        n1, m1, p1 = front.search_for_initialization(f1, f2, prev, window, ratio, check_ori)
        n2, m2, p2 = front.search_for_initialization(f1, f2, p1, window, ratio, check_ori)   # second call: updated windows
        out.append((n1, m1, p1, n2, m2, p2))
    assert out[0][0] > 100
    same(out[0], out[1])


def _local_points(m, stereo, th):
    rng = np.random.default_rng(5)
    if stereo:
        left, right = synth.stereo_pair(752, 480, 8)
        front = m.Front(1500, fx=FX, fy=FX, cx=376, cy=240, bf=FX * BASELINE)
        src, F = front.frame_mono(np.roll(left, (-3, -6), (0, 1))), front.frame_stereo(left, right)
    else:
        img = synth.frame(640, 480, 70)
        front = m.Front(1500)
        src, F = front.frame_mono(img), front.frame_mono(np.roll(img, (3, 6), (0, 1)))
    src.set_pose(np.eye(4))
    k1 = src.keys_un
    n = src.N
    uv = np.stack([k1["x"] + 6 + rng.uniform(-2, 2, n), k1["y"] + 3 + rng.uniform(-2, 2, n)], 1).astype(np.float32)
    view_cos = rng.choice(np.array([0.9985, 0.99], np.float32), n)
    level = k1["octave"].astype(np.int32)
    ur = (uv[:, 0] - np.float32(8.0) + rng.uniform(-3, 3, n)).astype(np.float32)
    in_view = rng.random(n) < 0.85
    obs = rng.random(n) < 0.7
    mps = []
    for i in range(n):
        mp = front.map_point(src, i, [0.0, 0.0, 5.0], n_obs=int(obs[i]))
        front.set_track(mp, bool(in_view[i]), float(uv[i, 0]), float(uv[i, 1]), float(ur[i]), int(level[i]), float(view_cos[i]))
        mps.append(mp)
    owner = front.map_point(src, 0, [0.0, 0.0, 5.0])
    for i in np.nonzero(rng.random(F.N) < 0.1)[0]:
        F.set_map_point(int(i), owner)
    n_found = front.search_by_projection_mp(F, mps, th, 0.8)
    return n_found, F.map_points()


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [1.0, 3.0])
def test_search_by_projection_local_map_points(stereo, th):
    a, b = (_local_points(m, stereo, th) for m in libs())
    assert a[0] > 100
    same(a, b)


def _frame_pair(m, kind):
    rng = np.random.default_rng(11)
    if kind == "mono":
        img = synth.frame(640, 480, 70)
        front = m.Front(1500)
        last, cur = front.frame_mono(img), front.frame_mono(np.roll(img, (3, 6), (0, 1)))
        tz = 0.0
    else:
        left, right = synth.stereo_pair(752, 480, 8)
        front = m.Front(1500, fx=FX, fy=FX, cx=376, cy=240, bf=FX * BASELINE)
        last = front.frame_stereo(left, right)
        cur = front.frame_stereo(np.roll(left, (3, 6), (0, 1)), np.roll(right, (3, 6), (0, 1)))
        tz = {"forward": -0.3, "backward": 0.3, "still": 0.0}[kind]
    last.set_pose(np.eye(4))
    z = 5.0
    T = np.eye(4, dtype=np.float32)
    T[0, 3], T[1, 3], T[2, 3] = 6 * z / float(front.fx), 3 * z / float(front.fy), tz
    cur.set_pose(T)
    for i in range(last.N):
        if rng.random() < 0.7:
            zi = z * (1 + rng.uniform(-0.05, 0.05))
            x, y = float(last.keys_un["x"][i]), float(last.keys_un["y"][i])
            pos = [(x - float(front.cx)) / float(front.fx) * zi, (y - float(front.cy)) / float(front.fy) * zi, zi]
            mp = front.map_point(last, i, pos, n_obs=int(rng.random() < 0.6))
            last.set_map_point(i, mp)
            if rng.random() < 0.05:
                last.set_outlier(i, True)
    # some keypoints of the current frame already carry points, with and without observations
    for i in range(0, cur.N, 13):
        cur.set_map_point(i, front.map_point(last, 0, [0.0, 0.0, 5.0], n_obs=(i // 13) % 2))
    return front, last, cur


@pytest.mark.parametrize("kind", ["mono", "still", "forward", "backward"])
@pytest.mark.parametrize("th", [7.0, 15.0])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_projection_last_frame(kind, th, check_ori):
    out = []
    for m in libs():
        front, last, cur = _frame_pair(m, kind)
        n = front.search_by_projection_frame(cur, last, th, kind == "mono", 0.9, check_ori)
        out.append((n, cur.map_points()))
    assert out[0][0] > 100
    same(out[0], out[1])


@pytest.fixture(scope="module", params=["mono", "stereo"])
def bow_pair(request, dropin_variant):
    return [bow_case.build(m, stereo=request.param == "stereo") for m in libs()], request.param


@pytest.mark.parametrize("check_ori", [True, False])
@pytest.mark.parametrize("ratio", [0.7, 0.9])
def test_search_by_bow(bow_pair, check_ori, ratio):
    (a, b), _ = bow_pair
    ra = a.front.search_by_bow_kf_frame(a.kf1, a.f2, ratio, check_ori)
    rb = b.front.search_by_bow_kf_frame(b.kf1, b.f2, ratio, check_ori)
    assert ra[0] > 40
    same(ra, rb)
    ra = a.front.search_by_bow_kf_kf(a.kf1, a.kf2, ratio, check_ori)
    rb = b.front.search_by_bow_kf_kf(b.kf1, b.kf2, ratio, check_ori)
    assert ra[0] > 20
    same(ra, rb)


@pytest.mark.parametrize("only_stereo", [False, True])
@pytest.mark.parametrize("check_ori", [False, True])
@pytest.mark.parametrize("epipole_inside", [False, True])
def test_search_for_triangulation(bow_pair, only_stereo, check_ori, epipole_inside):
    (a, b), kind = bow_pair
    out = []
    for c in (a, b):
        kf2 = c.kf2_center if epipole_inside else c.kf2
        out.append(c.front.search_for_triangulation(c.kf1, kf2, c.F12, only_stereo, 0.6, check_ori))
    if kind == "mono":
        assert out[0][0] == 0 if only_stereo else out[0][0] > 20
    else:
        assert out[0][0] > 5
    same(out[0], out[1])


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [3.0, 8.0])
def test_fuse(stereo, th):
    out = []
    for m in libs():
        c = bow_case.build_fuse(m, stereo)
        src = np.concatenate([c.src, c.src[:40], [c.mp_of2[5]]]).astype(np.int32)   # duplicates and a point of the key frame
        n_points = int(max(c.src.max(), c.mp_of2.max())) + 1
        n = c.front.fuse(c.kf2, src, th)
        out.append((n,) + graph(c.front, n_points, c.kf1, c.kf2))
    assert out[0][0] > (20 if stereo else 300)
    same(out[0], out[1])


def _scw(scale):
    S = bow_case.pose(bow_case.T_FUSE) * np.float32(1.0)
    S[:3, :] *= np.float32(scale)
    return S


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th,scale", [(4, 1.25), (10, 1.0)])
def test_search_by_projection_keyframe_scw(stereo, th, scale):
    out = []
    for m in libs():
        c = bow_case.build_fuse(m, stereo)
        matched = np.full(c.f2.N, -1, np.int32)
        matched[5::9] = c.src[np.arange(len(matched[5::9])) % len(c.src)]
        out.append(c.front.search_by_projection_kf_scw(c.kf2, _scw(scale), c.src, matched, th))
    assert out[0][0] > 200
    same(out[0], out[1])


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th,scale", [(4.0, 1.25), (10.0, 1.0)])
def test_fuse_scw(stereo, th, scale):
    out = []
    for m in libs():
        c = bow_case.build_fuse(m, stereo)
        src = np.concatenate([c.src, c.src[:25], c.mp_of2[c.mp_of2 >= 0][:30]]).astype(np.int32)
        n_points = int(max(c.src.max(), c.mp_of2.max())) + 1
        n, rep = c.front.fuse_scw(c.kf2, _scw(scale), src, th)
        out.append((n, rep) + graph(c.front, n_points, c.kf1, c.kf2))
    n, rep = out[0][0], out[0][1]
    assert n > 300 and (rep >= 0).sum() > 100 and n > (rep >= 0).sum()   # both branches: replace and add
    same(out[0], out[1])


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th,s12", [(7.5, 1.0), (10.0, 1.02)])
def test_search_by_sim3(stereo, th, s12):
    out = []
    for m in libs():
        c = bow_case.build_fuse(m, stereo)
        # camera 1 at [I|0], camera 2 at [I|T_FUSE]: p_c1 = s12 * p_c2 + t12
        t12 = -np.float32(s12) * bow_case.T_FUSE
        matches12 = np.full(c.f1.N, -1, np.int32)
        owned2 = np.nonzero(c.mp_of2 >= 0)[0]
        for k, i1 in enumerate(range(0, c.f1.N, 14)):      # matches the caller already holds
            matches12[i1] = c.mp_of2[owned2[(7 * k) % len(owned2)]]
        out.append(c.front.search_by_sim3(c.kf1, c.kf2, matches12, np.float32(s12), np.eye(3, dtype=np.float32), t12, th))
    assert out[0][0] > (100 if s12 == 1.0 else 20)
    same(out[0], out[1])


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th,orb_dist", [(10.0, 100), (3.0, 64)])
@pytest.mark.parametrize("check_ori", [True, False])
def test_search_by_projection_relocalization(stereo, th, orb_dist, check_ori):
    out = []
    for m in libs():
        c = bow_case.build_fuse(m, stereo)
        cur = c.f2                                   # pose T_FUSE; key frame 1 holds the points `src`
        for i in range(3, cur.N, 11):
            cur.set_map_point(i, int(c.mp_of2[i]) if c.mp_of2[i] >= 0 else int(c.src[0]))
        found = c.src[::6]
        n = c.front.search_by_projection_frame_kf(cur, c.kf1, found, th, orb_dist, 0.9, check_ori)
        out.append((n, cur.map_points()))
    assert out[0][0] > 150
    same(out[0], out[1])


def test_three_threads_use_the_matcher_at_once():
    """Tracking, LocalMapping and LoopClosing construct ORBmatcher objects on their own threads and search concurrently
    (SURVEY.md 8b). The drop-in keeps one device handle per host thread; results must not depend on what the other
    threads are doing."""
    import threading
    ref, drop = libs()
    rounds = 3

    def jobs(m):
        a = [T for T in (_frame_pair(m, "mono") for _ in range(rounds))]
        b = [bow_case.build_fuse(m, False) for _ in range(rounds)]
        c = [bow_case.build(m, stereo=False) for _ in range(rounds)]
        return [
            [lambda s=s: (s[0].search_by_projection_frame(s[2], s[1], 15.0, True, 0.9, True), s[2].map_points()) for s in a],
            [lambda s=s: (s.front.fuse(s.kf2, s.src, 3.0), s.kf2.map_points()) for s in b],
            [lambda s=s: (s.front.search_by_bow_kf_kf(s.kf1, s.kf2, 0.75, True),
                          s.front.search_for_triangulation(s.kf1, s.kf2, s.F12, False, 0.6, True)) for s in c],
        ]

    expect = [[fn() for fn in lane] for lane in jobs(ref)]
    lanes = jobs(drop)
    got = [[None] * rounds for _ in lanes]
    errors = []
    start = threading.Barrier(len(lanes))

    def run(k):
        try:
            start.wait()
            for r, fn in enumerate(lanes[k]):
                got[k][r] = fn()
        except BaseException as e:   # noqa: BLE001 - reported by the main thread
            errors.append(e)

    threads = [threading.Thread(target=run, args=(k,)) for k in range(len(lanes))]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    assert not errors, errors
    for k in range(len(lanes)):
        for r in range(rounds):
            same(expect[k][r], got[k][r])


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 5), (1241, 376, 2000, 6), (640, 480, 1000, 7)])
def test_stereo_frame_constructor(w, h, nf, seed, dropin_variant):
    """Frame::Frame(imLeft, imRight, ...) (R/src/Frame.cpp:67-128): in the "front" library both extractions and
    Frame::ComputeStereoMatches (host/FrameStereo.cc -> orbx_stereo_match on the device pyramids) are the product's."""
    left, right = synth.stereo_pair(w, h, seed)
    out = []
    for m in libs():
        F = m.Front(nf, fx=FX, fy=FX, cx=w / 2, cy=h / 2, bf=FX * BASELINE).frame_stereo(left, right)
        out.append((F.keys, F.keys_un, F.desc, F.keys_right, F.desc_right, F.u_right, F.depth))
    assert (out[0][5] >= 0).sum() > 100
    same(out[0], out[1])
