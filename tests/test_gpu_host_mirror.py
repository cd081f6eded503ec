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


@pytest.mark.parametrize("ndev", [1, 2])
def test_cpp_shard_group_wrapper(pkg, mirror, ndev):
    """orbx_host::ShardGroup (host/orbx_shard.h, the C++ face of orbx_shard_*): sharded top-2 and a sequence extracted
    by the group equal the single-GPU entry points."""
    if pkg.device_count() < ndev:
        pytest.skip(f"needs {ndev} GPUs")
    ndb, nq = 60_000, 300
    db = pkg.synth.descriptors(ndb, 51)
    q = pkg.synth.planted_queries(db, nq, 52)
    idx, d1, d2 = (np.empty(nq, np.int32) for _ in range(3))
    acc = np.empty(nq, np.uint8)
    w, h, nf, count = 320, 240, 300, 11
    frames = np.stack([pkg.synth.frame(w, h, 500 + i) for i in range(count)])
    cap = nf + 64
    kps = pkg.pinned_empty((count, cap), pkg.KP_DTYPE)
    desc = pkg.pinned_empty((count, cap, 32), np.uint8)
    n = pkg.pinned_empty(count, np.int32)
    err = C.create_string_buffer(512)
    fn = mirror.mirror_shard_run
    fn.argtypes = [C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_float, C.c_int] + [C.c_void_p] * 5 + \
                  [C.c_int] * 4 + [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_char_p, C.c_int]
    rc = fn(ndev, db.ctypes.data, ndb, q.ctypes.data, nq, 0.6, 50, idx.ctypes.data, d1.ctypes.data, d2.ctypes.data,
            acc.ctypes.data, frames.ctypes.data, count, w, h, nf, kps.ctypes.data, desc.ctypes.data, cap, n.ctypes.data,
            err, 512)
    assert rc == 0, err.value
    exp = pkg.ORBmatcher(0.6).match_bf2(q, db)
    for a, b in zip((idx, d1, d2, acc), exp):
        assert np.array_equal(a, b)
    one = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    for f in range(count):
        k1, dd = one(frames[f])
        assert n[f] == len(k1) and np.array_equal(kps[f, :n[f]].view(np.uint8), k1.view(np.uint8)) and np.array_equal(desc[f, :n[f]], dd)
