"""Matcher / stereo probe for profiling runs: one pass over each device path of the matcher with realistic
sizes (product code only, no oracle). Prints host-side wall times per call; ncu captures the kernels.

  python tools/match_probe.py
"""
import importlib
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
pkg = importlib.import_module("orb-slam2-_b200")


def timed(name, fn, reps=3):
    fn()
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    dt = (time.perf_counter() - t0) / reps
    print(f"{name}: {dt * 1e3:.3f} ms per call (host wall, host buffers)")
    return out


def main():
    rng = np.random.default_rng(3)
    m = pkg.ORBmatcher(0.6)
    # M1: brute force top-2, 2048 queries x 1M rows
    q = rng.integers(0, 256, (2048, 32), dtype=np.uint8)
    db = rng.integers(0, 256, (1 << 20, 32), dtype=np.uint8)
    timed("match_bf2 2048 x 1M", lambda: m.match_bf2(q, db), reps=2)

    # M2-M5: windowed search, 2000-keypoint frames
    w, h = 640, 480
    img1 = pkg.synth.frame(w, h, 70)
    img2 = np.roll(img1, (3, 6), (0, 1))
    ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k1, d1 = ex(img1)
    k2, d2 = ex(img2)
    bounds = (0.0, float(w), 0.0, float(h))
    qxy = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
    nq = len(k1)
    r = np.full(nq, 100.0, np.float32)
    lo = np.zeros(nq, np.int32)
    off, idx, dist = timed("match_window r=100 level 0", lambda: m.match_window(k2, d2, bounds, qxy, r, lo, lo, d1))
    print("  candidates:", int(off[-1]))
    r = np.full(nq, 15.0, np.float32)
    lv = k1["octave"].astype(np.int32)
    off, idx, dist = timed("match_window r=15 level+-1",
                           lambda: m.match_window(k2, d2, bounds, qxy, r, np.maximum(lv - 1, 0), lv + 1, d1))
    print("  candidates:", int(off[-1]))

    # S1: stereo
    w, h = 1241, 376
    left, right = pkg.synth.stereo_pair(w, h, 6)
    exl = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl(left)
    kr, dr = exr(right)
    mb = np.float32(0.11)
    mbf = np.float32(435.2) * mb
    u, z, n = timed("stereo_match 1241x376 2000 kps", lambda: m.stereo_match(exl, exr, kl, dl, kr, dr, mbf, mb))
    print("  stereo matches:", n, "of", len(kl))


if __name__ == "__main__":
    main()
