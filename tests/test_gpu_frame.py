"""SURVEY.md 8f row 3 on the device: cv::undistortPoints (Frame::UndistortKeyPoints, R/src/Frame.cpp:470-504) and the
device-resident frame handle whose keypoints, descriptors and grid stay in HBM between extraction and the windowed
searches (orbx_frame_*), against the oracle, the cv2 golden vectors and the upload-per-call path."""
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
G = Path(__file__).resolve().parent / "golden"
TUM1_K = np.array([517.306408, 516.469215, 318.643040, 255.313989], np.float32)
TUM1_D = np.array([0.262383, -0.953104, -0.005358, 0.002628, 1.163314], np.float32)


def test_undistort_points_vs_cv2_golden_and_oracle(pkg, oracle):
    d = np.load(G / "undistort_cv2.npz")
    for cam in ("tum1", "tum2", "euroc", "strong"):
        got = pkg.undistort_points(d[cam + "_in"], d[cam + "_K4"], d[cam + "_dist"])
        assert np.array_equal(got.view(np.uint32), d[cam + "_out"].view(np.uint32)), cam
    rng = np.random.default_rng(7)
    pts = np.stack([rng.uniform(-100, 2000, 300000), rng.uniform(-100, 1200, 300000)], 1).astype(np.float32)
    for K4, dist in ((TUM1_K, TUM1_D), (np.array([718.856, 718.856, 607.1928, 185.2157], np.float32), np.zeros(4, np.float32)),
                     (np.array([300, 300, 320, 240], np.float32), np.array([-0.9, 0.1, 0.01, -0.02], np.float32))):
        got = pkg.undistort_points(pts, K4, dist)
        exp = oracle.undistort_points(pts, K4, dist)
        assert np.array_equal(got.view(np.uint32), exp.view(np.uint32))
    assert pkg.undistort_points(np.zeros((0, 2), np.float32), TUM1_K, TUM1_D).shape == (0, 2)


def _queries(pkg, k2, d2, rng, nq=1500):
    i = rng.integers(0, len(k2), nq)
    xy = np.stack([k2["x"][i] + rng.uniform(-8, 8, nq), k2["y"][i] + rng.uniform(-8, 8, nq)], 1).astype(np.float32)
    r = rng.uniform(3, 40, nq).astype(np.float32)
    lo = rng.integers(-1, 6, nq).astype(np.int32)
    hi = np.where(rng.random(nq) < 0.3, -1, lo + rng.integers(0, 3, nq)).astype(np.int32)
    return xy, r, lo, hi, d2[i].copy()


@pytest.mark.parametrize("distorted", [False, True])
def test_resident_frame_equals_upload_per_call(pkg, oracle, distorted):
    """orbx_frame_set + orbx_frame_match_window against orbx_match_window on the (oracle-)undistorted keypoints: same
    candidates in the same order with the same distances; mvKeysUn read back equals the oracle's undistortion."""
    w, h = 640, 480
    ex = pkg.ORBextractor(1500, 1.2, 8, 20, 7, max_width=w, max_height=h)
    m = pkg.ORBmatcher(0.9)
    rng = np.random.default_rng(5)
    fr = pkg.Frame(m)
    for seed in (31, 32):   # the handle is refilled for the second frame
        k, d = ex(pkg.synth.frame(w, h, seed))
        if distorted:
            K4, dist = TUM1_K, TUM1_D
            corners = oracle.undistort_points(np.array([[0, 0], [w, 0], [0, h], [w, h]], np.float32), K4, dist)
            bounds = np.array([min(corners[0, 0], corners[2, 0]), max(corners[1, 0], corners[3, 0]),
                               min(corners[0, 1], corners[1, 1]), max(corners[2, 1], corners[3, 1])], np.float32)   # :527-530
            ku = k.copy()
            un = oracle.undistort_points(np.stack([k["x"], k["y"]], 1), K4, dist)
            ku["x"], ku["y"] = un[:, 0], un[:, 1]
        else:
            K4, dist, bounds, ku = None, None, np.array([0, w, 0, h], np.float32), k
        fr.set(k, d, bounds, K4, dist)
        assert np.array_equal(fr.keys_un().view(np.uint8), ku.view(np.uint8))
        q = _queries(pkg, ku, d, rng)
        eo, ei, ed = m.match_window(ku, d, bounds, *q)
        go, gi, gd = fr.match_window(*q)
        assert len(ei) > 5000
        assert np.array_equal(go, eo) and np.array_equal(gi, ei) and np.array_equal(gd, ed)
        # and against the oracle's grid for a sample of the queries
        grid = oracle.Grid(ku, bounds)
        for j in range(0, len(q[1]), 97):
            exp = grid.query(q[0][j, 0], q[0][j, 1], q[1][j], int(q[2][j]), int(q[3][j]))
            assert np.array_equal(gi[go[j]:go[j + 1]], exp)
    # an empty frame and an empty query list
    fr.set(np.zeros(0, pkg.KP_DTYPE), np.zeros((0, 32), np.uint8), np.array([0, w, 0, h], np.float32))
    o, i_, d_ = fr.match_window(q[0][:5], q[1][:5], q[2][:5], q[3][:5], q[4][:5])
    assert o.tolist() == [0] * 6 and len(i_) == 0
    o, i_, d_ = fr.match_window(np.zeros((0, 2), np.float32), np.zeros(0, np.float32), np.zeros(0, np.int32),
                                np.zeros(0, np.int32), np.zeros((0, 32), np.uint8))
    assert o.tolist() == [0]


def test_resident_frame_from_device_batch(pkg):
    """Keypoints and descriptors go from orbx_extract_batch_device's outputs into the frame handle without crossing PCIe."""
    torch = pytest.importorskip("torch")
    w, h, B, cap = 320, 240, 3, 700
    imgs = np.stack([pkg.synth.frame(w, h, 800 + i) for i in range(B)])
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    m = pkg.ORBmatcher(0.9)
    d_img = torch.from_numpy(imgs).cuda()
    d_k = torch.empty((B, cap, 7), dtype=torch.float32, device="cuda")
    d_d = torch.empty((B, cap, 32), dtype=torch.uint8, device="cuda")
    d_n = torch.zeros(B, dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    ex.extract_batch_device(d_img.data_ptr(), B, w, h, w, w * h, d_k.data_ptr(), d_d.data_ptr(), cap, d_n.data_ptr(), s)
    torch.cuda.synchronize()
    n = d_n.cpu().numpy()
    bounds = np.array([0, w, 0, h], np.float32)
    fr = pkg.Frame(m)
    rng = np.random.default_rng(9)
    for f in range(B):
        hk, hd = ex(imgs[f])
        assert n[f] == len(hk)
        fr.set_device(d_k[f].data_ptr(), d_d[f].data_ptr(), int(n[f]), bounds)
        assert np.array_equal(fr.keys_un().view(np.uint8), hk.view(np.uint8))
        q = _queries(pkg, hk, hd, rng, 300)
        assert all(np.array_equal(a, b) for a, b in zip(fr.match_window(*q), m.match_window(hk, hd, bounds, *q)))
