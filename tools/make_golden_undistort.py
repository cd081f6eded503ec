#!/usr/bin/env python3
"""tests/golden/undistort_cv2.npz: inputs and the outputs of the REAL cv2.undistortPoints (4.13.0) for the camera
models of the reference's example settings (TUM1 / TUM2 / EuRoC-like with 4 and 5 coefficients, one strongly distorted
lens that drives the fixed-point iteration's denominator negative for far points). Pins the oracle's restatement and
the CUDA kernel without cv2 at test time. Run in the authoring container."""
import sys
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parents[1]
CAMERAS = {   # fx fy cx cy | k1 k2 p1 p2 [k3]   (R/Examples/Monocular/TUM1.yaml, TUM2.yaml, EuRoC.yaml values)
    "tum1": ([517.306408, 516.469215, 318.643040, 255.313989], [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]),
    "tum2": ([520.908620, 521.007327, 325.141442, 249.701764], [0.231222, -0.784899, -0.003257, -0.000105, 0.917205]),
    "euroc": ([458.654, 457.296, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]),
    "strong": ([300.0, 300.0, 320.0, 240.0], [-0.9, 0.1, 0.01, -0.02, 0.0]),
}


def main():
    assert cv2.__version__.startswith("4.13"), cv2.__version__
    rng = np.random.Generator(np.random.PCG64(2024))
    d = {"cv2_version": np.array(cv2.__version__)}
    for name, (K4, dist) in CAMERAS.items():
        K4, dist = np.array(K4, np.float32), np.array(dist, np.float32)
        pts = np.stack([rng.uniform(-40, 800, 4000), rng.uniform(-40, 560, 4000)], 1).astype(np.float32)
        pts[:4] = [[0, 0], [752, 0], [0, 480], [752, 480]]       # ComputeImageBounds' corners
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        out = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, dist.reshape(-1, 1), None, K).reshape(-1, 2)
        d[name + "_K4"], d[name + "_dist"], d[name + "_in"], d[name + "_out"] = K4, dist, pts, out.astype(np.float32)
    np.savez_compressed(ROOT / "tests" / "golden" / "undistort_cv2.npz", **d)
    print("wrote", ROOT / "tests" / "golden" / "undistort_cv2.npz")


if __name__ == "__main__":
    sys.exit(main())
