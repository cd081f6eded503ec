"""Oracle primitives pinned live against the installed OpenCV (cv2 4.13) when it imports, and the
sincosf restatement against the host libm."""
import struct

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")


@pytest.mark.parametrize("w,h", [(640, 480), (752, 480), (1241, 376), (533, 400), (101, 77)])
def test_resize_vs_cv2(oracle, pkg, w, h):
    im = pkg.synth.frame(w, h, 1)
    for s in (1.2, 1.44, 1.5, 3.0):
        dw, dh = int(round(w / s)), int(round(h / s))
        assert np.array_equal(oracle.resize(im, dw, dh), cv2.resize(im, (dw, dh), interpolation=cv2.INTER_LINEAR))


@pytest.mark.parametrize("w,h", [(640, 480), (1241, 376), (37, 29), (9, 9), (7, 20)])
def test_blur_vs_cv2(oracle, pkg, w, h):
    for kind in ("shapes", "noise"):
        im = pkg.synth.frame(w, h, 2, kind)
        assert np.array_equal(oracle.blur7(im), cv2.GaussianBlur(im, (7, 7), 2, 2, borderType=cv2.BORDER_REFLECT_101))


def test_fast_vs_cv2_random_cells(oracle, pkg):
    rng = np.random.default_rng(0)
    total = 0
    for it in range(60):
        w, h = int(rng.integers(7, 60)), int(rng.integers(7, 60))
        big = pkg.synth.frame(200, 200, it, "shapes" if it % 3 else "noise")
        x0, y0 = int(rng.integers(0, 200 - w)), int(rng.integers(0, 200 - h))
        cell = big[y0:y0 + h, x0:x0 + w]  # strided view, like the reference's rowRange/colRange ROI
        for th in (20, 7, 0, 100):
            det = cv2.FastFeatureDetector_create(threshold=th, nonmaxSuppression=True,
                                                 type=cv2.FAST_FEATURE_DETECTOR_TYPE_9_16)
            kps = det.detect(np.ascontiguousarray(cell))
            exp = np.array([[int(k.pt[0]), int(k.pt[1]), int(k.response)] for k in kps], np.int32).reshape(-1, 3)
            assert np.array_equal(oracle.fast(cell, th, True), exp)
            total += len(exp)
    assert total > 1000


def test_fast_atan2_vs_cv2(oracle):
    rng = np.random.default_rng(3)
    for _ in range(20000):
        y, x = float(rng.integers(-70000, 70000)), float(rng.integers(-70000, 70000))
        assert oracle.fast_atan2(y, x) == np.float32(cv2.fastAtan2(y, x))


def test_sincosf_port_equals_libm_on_orb_domain(oracle):
    """Every float in [3.0, 3.0+2^24 ulps) and a strided sweep of [0, 2*pi]: 0 mismatches. The
    exhaustive [0, 2*pi] sweep (1.09e9 values, 0 mismatches on this host) is tools/sincosf_exhaustive.py."""
    lo = struct.unpack("<I", struct.pack("<f", 3.0))[0]
    bad, first = oracle.sincosf_sweep(lo, lo + (1 << 24))
    assert bad == 0, hex(first)
    hi = struct.unpack("<I", struct.pack("<f", 6.2831855))[0]
    for start in range(0, hi, hi // 16):
        bad, first = oracle.sincosf_sweep(start, min(hi, start + (1 << 21)))
        assert bad == 0, hex(first)


def test_undistort_points_vs_cv2(oracle):
    """cv::undistortPoints(src, dst, K, dist, noArray(), K) as Frame::UndistortKeyPoints / ComputeImageBounds call it
    (R/src/Frame.cpp:491, :523): every coordinate bit-identical, 4 and 5 coefficients, weak and strong lenses."""
    rng = np.random.default_rng(11)
    cams = [([517.306408, 516.469215, 318.643040, 255.313989], [0.262383, -0.953104, -0.005358, 0.002628, 1.163314]),
            ([458.654, 457.296, 367.215, 248.375], [-0.28340811, 0.07395907, 0.00019359, 1.76187114e-05]),
            ([300.0, 300.0, 320.0, 240.0], [-0.9, 0.1, 0.01, -0.02, 0.0])]
    for K4, dist in cams:
        K4, dist = np.array(K4, np.float32), np.array(dist, np.float32)
        pts = np.stack([rng.uniform(-50, 810, 50000), rng.uniform(-50, 570, 50000)], 1).astype(np.float32)
        K = np.array([[K4[0], 0, K4[2]], [0, K4[1], K4[3]], [0, 0, 1]], np.float32)
        exp = cv2.undistortPoints(pts.reshape(-1, 1, 2), K, dist.reshape(-1, 1), None, K).reshape(-1, 2)
        got = oracle.undistort_points(pts, K4, dist)
        assert np.array_equal(got.view(np.uint32), exp.view(np.uint32))
