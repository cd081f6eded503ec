"""The C++ drop-in class ORB_SLAM2::ORBextractor (orb-slam2-_b200/host/ORBextractor.{h,cc}), driven the way
Frame.cpp drives the reference's class, against the oracle. The test build compiles the class against the cv
shim (no OpenCV C++ on either box); the compute goes through liborbx.so."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

LIB = Path(__file__).resolve().parents[1] / "orb-slam2-_b200" / "host" / "libmirror_test.so"


@pytest.fixture(scope="module")
def mirror():
    L = C.CDLL(str(LIB))
    L.mirror_create.restype = C.c_void_p
    L.mirror_create.argtypes = [C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
    L.mirror_destroy.argtypes = [C.c_void_p]
    L.mirror_extract.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    L.mirror_tables.argtypes = [C.c_void_p] + [C.c_void_p] * 4
    L.mirror_level_size.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.mirror_get_level.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    return L


def _extract(L, h, img, kp_dtype):
    cap = 20000
    k = np.empty(cap, kp_dtype)
    d = np.empty((cap, 32), np.uint8)
    n = L.mirror_extract(h, img.ctypes.data, img.shape[1], img.shape[0], img.strides[0], k.ctypes.data, d.ctypes.data, cap)
    assert n >= 0, n
    return k[:n].copy(), d[:n].copy()


def test_cpp_mirror_equals_oracle(pkg, oracle, mirror):
    h = mirror.mirror_create(1000, 1.2, 8, 20, 7)
    ox = oracle.Extractor(1000, 1.2, 8, 20, 7)
    try:
        # one instance, several shapes in a row (Tracking reuses its extractors for the whole sequence)
        for (w, hh, seed, kind) in [(640, 480, 0, "shapes"), (640, 480, 1, "noise"), (752, 480, 2, "shapes"),
                                    (640, 480, 3, "constant"), (640, 480, 4, "shapes")]:
            img = pkg.synth.frame(w, hh, seed, kind)
            gk, gd = _extract(mirror, h, img, pkg.KP_DTYPE)
            ok, od = ox(img)
            assert len(gk) == len(ok)
            assert np.array_equal(gk.view(np.uint8), ok.view(np.uint8))
            assert np.array_equal(gd, od)
            if kind != "constant":
                for l in range(8):  # public mvImagePyramid, read by Frame::ComputeStereoMatches
                    ww, h2 = C.c_int(), C.c_int()
                    mirror.mirror_level_size(h, l, C.byref(ww), C.byref(h2))
                    buf = np.empty((h2.value, ww.value), np.uint8)
                    mirror.mirror_get_level(h, l, buf.ctypes.data, ww.value)
                    assert np.array_equal(buf, ox.level(l)), l
        t = [np.empty(8, np.float32) for _ in range(4)]
        mirror.mirror_tables(h, *[a.ctypes.data for a in t])
        o = ox.tables()
        for got, key in zip(t, ("scale", "inv_scale", "sigma2", "inv_sigma2")):
            assert np.array_equal(got.view(np.uint32), o[key].view(np.uint32)), key
    finally:
        mirror.mirror_destroy(h)


def test_cpp_mirror_strided_roi_input(pkg, oracle, mirror):
    """cv::Mat ROIs have a row step larger than their width (and possibly unaligned rows)."""
    big = pkg.synth.frame(700, 500, 9)
    roi = big[7:487, 13:653]  # 640x480 view, step 700, start offset 13
    h = mirror.mirror_create(1000, 1.2, 8, 20, 7)
    try:
        gk, gd = _extract(mirror, h, roi, pkg.KP_DTYPE)
        ok, od = oracle.Extractor(1000, 1.2, 8, 20, 7)(np.ascontiguousarray(roi))
        assert np.array_equal(gk.view(np.uint8), ok.view(np.uint8)) and np.array_equal(gd, od)
    finally:
        mirror.mirror_destroy(h)
