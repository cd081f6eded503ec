#!/usr/bin/env python3
"""Host wall time of a windowed search against a device-resident frame (orbx_frame_match_window) next to the
upload-per-call orbx_match_window: 2000 keypoints, 1000 queries (SearchByProjection's size), radius 15."""
import sys
import time
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parents[1]))
import orbslam2_b200 as pkg  # noqa: E402

w, h = 640, 480
ex = pkg.ORBextractor(2000, 1.2, 8, 20, 7, max_width=w, max_height=h)
m = pkg.ORBmatcher(0.9)
k, d = ex(pkg.synth.frame(w, h, 3))
rng = np.random.default_rng(1)
nq = 1000
i = rng.integers(0, len(k), nq)
q = (np.stack([k["x"][i] + 2, k["y"][i] - 1], 1).astype(np.float32), np.full(nq, 15, np.float32),
     np.maximum(k["octave"][i] - 1, 0).astype(np.int32), (k["octave"][i] + 1).astype(np.int32), d[i].copy())
b = np.array([0, w, 0, h], np.float32)
fr = pkg.Frame(m)
K4 = np.array([517.3, 516.5, 318.6, 255.3], np.float32)
D = np.array([0.2624, -0.9531, -0.0054, 0.0026, 1.1633], np.float32)


def t(fn, n=200):
    for _ in range(20):
        fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n * 1e3


print(f"keypoints {len(k)}, queries {nq}")
print(f"orbx_match_window (uploads frame, builds grid):      {t(lambda: m.match_window(k, d, b, *q)):.3f} ms per call")
print(f"orbx_frame_set (host arrays, no distortion):          {t(lambda: fr.set(k, d, b)):.3f} ms per call")
print(f"orbx_frame_set (host arrays, undistort 5 coeffs):     {t(lambda: fr.set(k, d, b, K4, D)):.3f} ms per call")
fr.set(k, d, b)
print(f"orbx_frame_match_window (resident frame):             {t(lambda: fr.match_window(*q)):.3f} ms per call")
print(f"candidates per call: {len(fr.match_window(*q)[1])}")
