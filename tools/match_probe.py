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

    # node-restricted blocks, distinctive descriptors, frame epilogue, vocabulary descent
    node = lambda d: ((d[:, 0] >> 4).astype(np.int64) << 3) | (d[:, 5] >> 5).astype(np.int64)

    def featvec(n):
        order = np.argsort(n, kind="stable")
        ids, counts = np.unique(n, return_counts=True)
        return ids, np.concatenate([[0], np.cumsum(counts)]).astype(np.int32), order.astype(np.int32)
    (ia, oa, xa), (ib, ob, xb) = featvec(node(d1)), featvec(node(d2))
    common, pa, pb = np.intersect1d(ia, ib, return_indices=True)
    a_idx = np.concatenate([xa[oa[i]:oa[i + 1]] for i in pa]).astype(np.int32)
    b_idx = np.concatenate([xb[ob[i]:ob[i + 1]] for i in pb]).astype(np.int32)
    a_off = np.concatenate([[0], np.cumsum([oa[i + 1] - oa[i] for i in pa])]).astype(np.int32)
    b_off = np.concatenate([[0], np.cumsum([ob[i + 1] - ob[i] for i in pb])]).astype(np.int32)
    off, dist = timed("match_blocks 2000 x 2000 features, 128 nodes", lambda: m.match_blocks(d1, d2, a_off, a_idx, b_off, b_idx), reps=20)
    print("  pairs:", len(dist))
    sizes = rng.integers(2, 13, 2000).astype(np.int32)
    ooff = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    obs = pkg.synth.descriptors(int(ooff[-1]), 4)
    timed("distinctive 2000 map points x 2-12 observations", lambda: m.distinctive(obs, ooff), reps=20)
    K4 = np.array([517.3, 516.5, 318.6, 255.3], np.float32)
    D = np.array([0.2624, -0.9531, -0.0054, 0.0026, 1.1633], np.float32)
    fr = pkg.Frame(m)
    b4 = np.array(bounds, np.float32)
    timed("frame_set 2000 keypoints, undistort + grid", lambda: fr.set(k2, d2, b4, K4, D), reps=20)
    fr.set(k2, d2, b4)
    timed("frame_match_window r=15 level+-1 (resident frame)",
          lambda: fr.match_window(qxy, r, np.maximum(lv - 1, 0), lv + 1, d1), reps=20)
    kk, LL = 10, 4
    counts = [kk ** l for l in range(LL + 1)]
    parent = np.concatenate([[0]] + [np.repeat(np.arange(sum(counts[:l - 1]), sum(counts[:l])), kk) for l in range(1, LL + 1)]).astype(np.int32)
    voc = pkg.Vocabulary(m, parent, pkg.synth.descriptors(len(parent), 9), LL)
    timed(f"vocabulary_descend 2000 descriptors, {len(parent)} nodes", lambda: voc.descend(d1, 2), reps=20)


if __name__ == "__main__":
    main()
