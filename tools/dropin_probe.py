#!/usr/bin/env python3
"""Wall time of every ORBmatcher member through the reference-facing C++ interface: the reference's own
R/src/ORBmatcher.cpp (oracle/_ref/libref_front.so, one host thread -- these members are single-threaded in the
reference) next to the drop-in orb-slam2-_b200/host/ORBmatcher.cc over liborbx.so (oracle/_ref/libdropin_matcher.so),
on the same reference Frame / KeyFrame / MapPoint objects (640x480, 2000 features, the scenarios of
tests/test_gpu_dropin_matcher.py). The searches change the objects they run on, so every repetition builds a fresh
scenario and times the member call alone. Prints one JSON line."""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import bow_case  # noqa: E402
import ref_front_lib as rf  # noqa: E402
import test_gpu_dropin_matcher as T  # noqa: E402

REPS = int(sys.argv[1]) if len(sys.argv) > 1 else 7


def timed(fn):
    t0 = time.perf_counter()
    r = fn()
    return (time.perf_counter() - t0) * 1e3, r


def case_init(m):
    img = T.synth.frame(640, 480, 70)
    front = m.Front(2000)
    f1, f2 = front.frame_mono(img), front.frame_mono(np.roll(img, (3, 6), (0, 1)))
    prev = np.stack([f1.keys_un["x"], f1.keys_un["y"]], 1)
    return lambda: front.search_for_initialization(f1, f2, prev, 100, 0.9, True)[0]


def case_last_frame(m):
    front, last, cur = T._frame_pair(m, "mono")
    return lambda: front.search_by_projection_frame(cur, last, 15.0, True, 0.9, True)


def case_bow(m, which):
    c = bow_case.build(m, stereo=False)
    if which == "kf_frame":
        return lambda: c.front.search_by_bow_kf_frame(c.kf1, c.f2, 0.7, True)[0]
    if which == "kf_kf":
        return lambda: c.front.search_by_bow_kf_kf(c.kf1, c.kf2, 0.75, True)[0]
    return lambda: c.front.search_for_triangulation(c.kf1, c.kf2, c.F12, False, 0.6, False)[0]


def case_fuse(m, which):
    c = bow_case.build_fuse(m, False)
    S = T._scw(1.0)
    if which == "fuse":
        return lambda: c.front.fuse(c.kf2, c.src, 3.0)
    if which == "fuse_scw":
        return lambda: c.front.fuse_scw(c.kf2, S, c.src, 4.0)[0]
    if which == "scw":
        matched = np.full(c.f2.N, -1, np.int32)
        return lambda: c.front.search_by_projection_kf_scw(c.kf2, S, c.src, matched, 10)[0]
    if which == "sim3":
        m12 = np.full(c.f1.N, -1, np.int32)
        return lambda: c.front.search_by_sim3(c.kf1, c.kf2, m12, np.float32(1.0), np.eye(3, dtype=np.float32), -bow_case.T_FUSE, 7.5)[0]
    return lambda: c.front.search_by_projection_frame_kf(c.f2, c.kf1, c.src[::6], 10.0, 100, 0.9, True)


CASES = [
    ("SearchForInitialization", case_init),
    ("SearchByProjection(Cur,Last)", case_last_frame),
    ("SearchByProjection(F,MapPoints)", lambda m: (lambda: T._local_points(m, False, 1.0)[0])),   # includes its scenario
    ("SearchByBoW(KF,F)", lambda m: case_bow(m, "kf_frame")),
    ("SearchByBoW(KF,KF)", lambda m: case_bow(m, "kf_kf")),
    ("SearchForTriangulation", lambda m: case_bow(m, "tri")),
    ("Fuse(KF,MapPoints)", lambda m: case_fuse(m, "fuse")),
    ("Fuse(KF,Scw)", lambda m: case_fuse(m, "fuse_scw")),
    ("SearchByProjection(KF,Scw)", lambda m: case_fuse(m, "scw")),
    ("SearchBySim3", lambda m: case_fuse(m, "sim3")),
    ("SearchByProjection(Cur,KF,found)", lambda m: case_fuse(m, "reloc")),
]


def frame_rows():
    """Frame::Frame itself (R/src/Frame.cpp:67-128, :187-250) in libref_front.so and in libdropin_front.so, where
    ORBextractor.cpp and Frame::ComputeStereoMatches are the product's: extraction of the image(s) (the reference runs
    the two eyes on two threads), stereo matching, undistortion, grid assignment."""
    full = rf.variant(rf.DROPIN_PATH)
    out = {}
    for name, (w, h, nf, stereo) in {"Frame::Frame mono 640x480/1000": (640, 480, 1000, False),
                                     "Frame::Frame stereo 752x480/1200": (752, 480, 1200, True),
                                     "Frame::Frame stereo 1241x376/2000": (1241, 376, 2000, True)}.items():
        if stereo:
            left, right = T.synth.stereo_pair(w, h, 5)
        else:
            left = right = T.synth.frame(w, h, 5)
        ms = {"reference": [], "dropin": []}
        n = {}
        for label, m in (("reference", rf), ("dropin", full)):
            front = m.Front(nf, fx=435.2, fy=435.2, cx=w / 2, cy=h / 2, bf=47.9)
            for rep in range(REPS + 1):
                t0 = time.perf_counter()
                hnd = (m.lib().ref_frame_stereo(front._h, rf._p(left), rf._p(right), w, h, left.strides[0]) if stereo
                       else m.lib().ref_frame_mono(front._h, rf._p(left), w, h, left.strides[0]))
                t = (time.perf_counter() - t0) * 1e3
                F = m.Frame(front, hnd)
                if rep:
                    ms[label].append(t)
                n[label] = (F.N, int((F.u_right >= 0).sum()))
                del F
        assert n["reference"] == n["dropin"], (name, n)
        out[name] = {"keypoints": n["reference"][0], "stereo_matches": n["reference"][1],
                     "reference_ms": round(float(np.median(ms["reference"])), 3), "dropin_ms": round(float(np.median(ms["dropin"])), 3)}
    return out


def main():
    drop = rf.variant(rf.DROPIN_MATCHER_PATH)
    rows = frame_rows()
    for name, make in CASES:
        if name == "SearchByProjection(F,MapPoints)":
            continue   # its scenario builder runs the search itself; covered by the tests, not timed here
        ms = {"reference": [], "dropin": []}
        result = {}
        for rep in range(REPS + 1):
            for label, m in (("reference", rf), ("dropin", drop)):
                t, r = timed(make(m))
                if rep:      # repetition 0 warms the device handle, the pinned staging and the page cache
                    ms[label].append(t)
                result[label] = r
        assert result["reference"] == result["dropin"], (name, result)
        rows[name] = {"matches": int(result["reference"]), "reference_ms": round(float(np.median(ms["reference"])), 3),
                      "dropin_ms": round(float(np.median(ms["dropin"])), 3)}
    print(json.dumps({"probe": "dropin_matcher", "reps": REPS, "frame": "640x480, 2000 features", "rows": rows}))


if __name__ == "__main__":
    main()
