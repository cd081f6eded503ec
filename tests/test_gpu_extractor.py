"""GPU parity: every CUDA stage of the extractor against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SHAPES = [(640, 480), (752, 480), (1241, 376), (533, 400), (200, 150)]


@pytest.mark.parametrize("w,h", SHAPES)
def test_resize_bit_exact(pkg, oracle, w, h):
    img = pkg.synth.frame(w, h, 11)
    for s in (1.2, 1.44, 1.7):
        dw, dh = int(round(w / s)), int(round(h / s))
        assert np.array_equal(pkg.resize_u8(img, dw, dh), oracle.resize(img, dw, dh))


@pytest.mark.parametrize("w,h", SHAPES + [(37, 29), (64, 32), (65, 33), (4, 4), (38, 30), (39, 31), (12, 12), (13, 9),
                                          (11, 11), (33, 70), (130, 5)])
def test_blur_bit_exact(pkg, oracle, w, h):
    for kind in ("shapes", "noise"):
        img = pkg.synth.frame(w, h, 12, kind)
        assert np.array_equal(pkg.blur7_u8(img), oracle.blur7(img))


def test_fast_atan2_bit_exact(pkg, oracle):
    rng = np.random.default_rng(0)
    y = rng.integers(-70000, 70000, 200000).astype(np.float32)
    x = rng.integers(-70000, 70000, 200000).astype(np.float32)
    y[:8] = [0, 0, 5, -5, 1, -1, 0, 3]
    x[:8] = [0, 5, 0, 0, 1, -1, -7, 3]
    got = pkg.fast_atan2(y, x)
    exp = np.array([oracle.fast_atan2(a, b) for a, b in zip(y[:20000], x[:20000])], np.float32)
    assert np.array_equal(got[:20000].view(np.uint32), exp.view(np.uint32))


def test_sincosf_matches_host_libm_sample(pkg, oracle):
    rng = np.random.default_rng(1)
    a = (rng.random(100000) * 360.0).astype(np.float32) * np.float32(np.pi / 180.0)
    s, c = pkg.sincosf(a)
    for i in range(0, 100000, 7):
        es, ec = oracle.sincosf_libm(a[i])
        assert np.float32(es).view(np.uint32) == s[i].view(np.uint32)
        assert np.float32(ec).view(np.uint32) == c[i].view(np.uint32)


def test_sincosf_exhaustive_block_vs_port(pkg, oracle):
    """2^24 consecutive float bit patterns around pi (the port is itself pinned to libm exhaustively
    on the CPU side); the full [0, 2pi] sweep is tools/sincosf_exhaustive.py."""
    import struct
    lo = struct.unpack("<I", struct.pack("<f", 3.0))[0]
    hi = lo + (1 << 22)
    s, c = pkg.sincosf_sweep(lo, hi)
    bits = np.arange(lo, hi, dtype=np.uint32)
    a = bits.view(np.float32)
    idx = np.random.default_rng(2).integers(0, hi - lo, 3000)
    for i in idx:
        es, ec = oracle.sincosf_libm(a[i])
        assert np.float32(es).view(np.uint32) == s[i].view(np.uint32)
        assert np.float32(ec).view(np.uint32) == c[i].view(np.uint32)


def _cands_set(a):
    return set(map(tuple, np.asarray(a).tolist()))


@pytest.mark.parametrize("w,h,nf,kind", [(640, 480, 1000, "shapes"), (640, 480, 1000, "noise"),
                                         (640, 480, 1000, "lowcontrast"), (752, 480, 1200, "shapes"),
                                         (1241, 376, 2000, "shapes")])
def test_pyramid_blur_and_fast_candidates(pkg, oracle, w, h, nf, kind):
    img = pkg.synth.frame(w, h, 21, kind)
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ox = oracle.Extractor(nf, 1.2, 8, 20, 7)
    ex(img)
    ox(img)
    for l in range(8):
        assert np.array_equal(ex.image_pyramid_level(l), ox.level(l)), f"pyramid level {l}"
        gc, oc = ex.level_candidates(l), ox.candidates(l)
        assert len(gc) == len(oc), f"level {l}: {len(gc)} vs {len(oc)} candidates"
        assert _cands_set(gc) == _cands_set(oc), f"level {l} candidate set"
        if ox.kept(l) > 0:
            assert np.array_equal(ex.image_pyramid_level(l, blurred=True), ox.level(l, blurred=True)), f"blur {l}"


def test_fast_candidates_weak_corner_in_cell_with_strong_bounds_only(pkg, oracle):
    """A cell whose only corners are weak (minThFAST < margin <= iniThFAST) while every pixel that passes the kernel's
    cheap first test has a bound ABOVE iniThFAST: the fallback suppression at minThFAST must still run although the
    list of weakly bounded tasks is empty (frame 3 of the synthetic set has such a cell on level 0)."""
    img = pkg.synth.frame(640, 480, 3)
    ex = pkg.ORBextractor(1000, 1.2, 1, 20, 7, max_width=640, max_height=480)
    ox = oracle.Extractor(1000, 1.2, 1, 20, 7)
    ex(img)
    ox(img)
    assert _cands_set(ex.level_candidates(0)) == _cands_set(ox.candidates(0))


def test_octree_matches_oracle_in_order(pkg, oracle):
    ox = oracle.Extractor(1000, 1.2, 8, 20, 7)
    for seed, kind in [(31, "shapes"), (32, "shapes"), (33, "noise")]:
        img = pkg.synth.frame(640, 480, seed, kind)
        ox(img)
        feats = ox.tables()["feats"]
        for l in range(8):
            lv = ox.level(l)
            hh, ww = lv.shape
            minB, maxBX, maxBY = 16, ww - 16, hh - 16
            width, height = maxBX - minB, maxBY - minB
            wCell = int(np.ceil(width / (width // 30)))
            hCell = int(np.ceil(height / (height // 30)))
            c = ox.candidates(l)
            exp_idx = oracle.distribute(c, minB, maxBX, minB, maxBY, int(feats[l]))
            got = pkg.distribute(c, minB, maxBX, minB, maxBY, int(feats[l]), wCell, hCell)
            assert np.array_equal(got, c[exp_idx]), f"seed {seed} level {l}"


@pytest.mark.parametrize("w,h,nf,kind,seed", [
    (640, 480, 1000, "shapes", 0), (640, 480, 1000, "shapes", 1), (640, 480, 1000, "noise", 2),
    (640, 480, 1000, "lowcontrast", 3), (752, 480, 1200, "shapes", 4), (1241, 376, 2000, "shapes", 5),
    (1920, 1080, 4000, "shapes", 6)])
def test_extract_bit_exact(pkg, oracle, w, h, nf, kind, seed):
    img = pkg.synth.frame(w, h, seed, kind)
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ox = oracle.Extractor(nf, 1.2, 8, 20, 7)
    gk, gd = ex(img)
    ok, od = ox(img)
    assert len(gk) == len(ok)
    for f in ("x", "y", "size", "response", "octave", "class_id"):
        assert np.array_equal(gk[f], ok[f]), f
    # north_star allows 1e-3 deg on the angle; we require bit equality
    assert np.array_equal(gk["angle"].view(np.uint32), ok["angle"].view(np.uint32))
    assert np.array_equal(gd, od)


def test_constant_frame_gives_no_keypoints(pkg):
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    k, d = ex(pkg.synth.frame(640, 480, 0, "constant"))
    assert len(k) == 0 and d is None
    k, d = ex(np.empty((0, 0), np.uint8))
    assert len(k) == 0 and d is None


def test_batch_equals_single(pkg, oracle):
    imgs = [pkg.synth.frame(640, 480, 40 + i) for i in range(5)]
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=8)
    ox = oracle.Extractor(1000, 1.2, 8, 20, 7)
    res = ex.extract_batch(imgs)
    for im, (gk, gd) in zip(imgs, res):
        ok, od = ox(im)
        assert np.array_equal(gk.view(np.uint8), ok.view(np.uint8))
        assert np.array_equal(gd, od)


def test_errors_are_loud(pkg):
    with pytest.raises(pkg.OrbxError):
        pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=64, max_height=48)  # top level too small for the grid
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    with pytest.raises(pkg.OrbxError):
        ex(np.zeros((1080, 1920), np.uint8))  # larger than the handle was sized for
    with pytest.raises(TypeError):
        ex(np.zeros((480, 640), np.float32))


def test_extract_matches_reference_golden(pkg):
    """CUDA path vs golden vectors produced by the REFERENCE's own ORBextractor.cpp (tools/make_golden_ref.py;
    /root/reference does not exist on the GPU box, the vectors do)."""
    import hashlib
    from pathlib import Path
    e = np.load(Path(__file__).resolve().parent / "golden" / "ref_extract.npz")
    keys = sorted(k[:-4] for k in e.files if k.endswith("_sha"))
    assert len(keys) >= 6
    for key in keys:
        wh, nf, seed, kind = key.split("_")
        w, h = map(int, wh.split("x"))
        img = pkg.synth.frame(w, h, int(seed), kind)
        assert hashlib.sha256(img.tobytes()).hexdigest() == str(e[key + "_imgsha"])
        ex = pkg.ORBextractor(int(nf), 1.2, 8, 20, 7, max_width=w, max_height=h)
        k, d = ex(img)
        assert len(k) == int(e[key + "_n"]), key
        assert np.array_equal(k[:8].view(np.uint8), e[key + "_k8"].view(np.uint8)), key
        assert hashlib.sha256(k.tobytes() + d.tobytes()).hexdigest() == str(e[key + "_sha"]), key


def test_chunked_host_batch_equals_single(pkg, oracle):
    """More frames than one chunk: the two-lane pipeline of orbx_extract_batch must give every frame the
    result of a single-frame call (and of the oracle for a sample)."""
    B = 70
    frames = np.stack([pkg.synth.frame(320, 240, 200 + i) for i in range(B)])
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240, max_batch=128)
    ex.set_pipeline(chunk_frames=8, lanes=3)   # 9 chunks over 3 lanes: every lane is reused twice
    cap = 700
    kps = np.empty((B, cap), pkg.KP_DTYPE)
    desc = np.empty((B, cap, 32), np.uint8)
    n = np.zeros(B, np.int32)
    ex.extract_batch_into(frames, kps, desc, n)
    ox = oracle.Extractor(500, 1.2, 8, 20, 7)
    single = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240)
    for f in range(B):
        sk, sd = single(frames[f])
        assert n[f] == len(sk)
        assert np.array_equal(kps[f, :n[f]].view(np.uint8), sk.view(np.uint8)), f
        assert np.array_equal(desc[f, :n[f]], sd), f
        if f % 16 == 0:
            ok, od = ox(frames[f])
            assert np.array_equal(sk.view(np.uint8), ok.view(np.uint8)) and np.array_equal(sd, od)
    # a second call on the same handle (lanes are reused) and a non-contiguous list of frames
    res = ex.extract_batch([frames[i] for i in (5, 3, 69)])
    for (gk, gd), i in zip(res, (5, 3, 69)):
        assert np.array_equal(gk.view(np.uint8), kps[i, :n[i]].view(np.uint8)) and np.array_equal(gd, desc[i, :n[i]])


@pytest.mark.parametrize("cap", [700, 400])
def test_pinned_outputs_take_rows_directly(pkg, cap):
    """Pinned result arrays receive the device rows with no staging copy (orbx_host_alloc + the direct D2H path of
    orbx_extract_batch): same rows as the pageable path for every frame; cap below the per-frame row count
    truncates the copy but never the reported n."""
    B = 70
    frames = pkg.pinned_empty((B, 240, 320), np.uint8)
    for i in range(B):
        frames[i] = pkg.synth.frame(320, 240, 300 + i)
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240, max_batch=128)
    ex.set_pipeline(chunk_frames=16, lanes=4)  # 5 chunks over 4 lanes
    big = 700
    rk, rd, rn = np.empty((B, big), pkg.KP_DTYPE), np.empty((B, big, 32), np.uint8), np.zeros(B, np.int32)
    ex.extract_batch_into(frames, rk, rd, rn)
    pk, pd, pn = pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), np.zeros(B, np.int32)
    if cap >= rn.max():
        ex.extract_batch_into(frames, pk, pd, pn)
    else:
        with pytest.raises(pkg.OrbxError):
            ex.extract_batch_into(frames, pk, pd, pn)
    assert np.array_equal(pn, rn)
    for f in range(B):
        m = min(int(rn[f]), cap)
        if rn[f] <= cap:
            assert np.array_equal(pk[f, :m].view(np.uint8), rk[f, :m].view(np.uint8)), f
            assert np.array_equal(pd[f, :m], rd[f, :m]), f


def test_streaming_submissions_equal_synchronous_batch(pkg):
    """orbx_extract_batch_submit / _wait: three submissions in flight on the same handle (lanes keep rotating across
    submissions), waited for out of submission order, must give what the synchronous call gives; pageable result
    arrays, a stale ticket and a too-small capacity are refused."""
    B, S = 40, 3
    frames = [pkg.pinned_empty((B, 240, 320), np.uint8) for _ in range(S)]
    for s in range(S):
        for i in range(B):
            frames[s][i] = pkg.synth.frame(320, 240, 500 + 100 * s + i)
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240, max_batch=64)
    ex.set_pipeline(chunk_frames=8, lanes=3)   # 5 chunks per submission over 3 lanes
    cap = 700
    ref = []
    for s in range(S):
        rk, rd, rn = np.empty((B, cap), pkg.KP_DTYPE), np.empty((B, cap, 32), np.uint8), np.zeros(B, np.int32)
        ex.extract_batch_into(frames[s], rk, rd, rn)
        ref.append((rk, rd, rn))
    outs = [(pkg.pinned_empty((B, cap), pkg.KP_DTYPE), pkg.pinned_empty((B, cap, 32), np.uint8), pkg.pinned_empty(B, np.int32))
            for _ in range(S)]
    for rounds in range(2):   # the second round reuses tickets' ring slots and the lanes' arenas
        for o in outs:
            o[2][:] = -1
        tickets = [ex.extract_batch_submit(frames[s], *outs[s]) for s in range(S)]
        for s in (2, 0, 1):
            ex.extract_batch_wait(tickets[s])
        for s in range(S):
            rk, rd, rn = ref[s]
            gk, gd, gn = outs[s]
            assert np.array_equal(gn, rn)
            for f in range(B):
                assert np.array_equal(gk[f, :rn[f]].view(np.uint8), rk[f, :rn[f]].view(np.uint8)), (s, f)
                assert np.array_equal(gd[f, :rn[f]], rd[f, :rn[f]]), (s, f)
    with pytest.raises(pkg.OrbxError):
        ex.extract_batch_wait(tickets[0])          # already collected
    with pytest.raises(pkg.OrbxError):               # pageable results cannot take asynchronous copies
        ex.extract_batch_submit(frames[0], np.empty((B, cap), pkg.KP_DTYPE), np.empty((B, cap, 32), np.uint8), np.zeros(B, np.int32))
    small = (pkg.pinned_empty((B, 300), pkg.KP_DTYPE), pkg.pinned_empty((B, 300, 32), np.uint8), pkg.pinned_empty(B, np.int32))
    t = ex.extract_batch_submit(frames[0], *small)
    with pytest.raises(pkg.OrbxError):
        ex.extract_batch_wait(t)
    assert np.array_equal(small[2], ref[0][2])       # n is still reported


@pytest.mark.parametrize("w,h", [(321, 240), (1241, 376)])
def test_odd_width_batch_equals_single(pkg, w, h):
    """Tightly packed frames whose width is not a multiple of the arena pitch cross PCIe flat and are re-pitched on the
    device (host batch), or re-pitched with one 2-D copy (device batch): same rows as single-frame calls."""
    import torch
    B = 6
    frames = np.stack([pkg.synth.frame(w, h, 400 + i) for i in range(B)])
    ex = pkg.ORBextractor(800, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    ex.set_pipeline(chunk_frames=4, lanes=2)
    cap = 1000
    kps, desc, n = np.empty((B, cap), pkg.KP_DTYPE), np.empty((B, cap, 32), np.uint8), np.zeros(B, np.int32)
    ex.extract_batch_into(frames, kps, desc, n)
    d_frames = torch.from_numpy(frames).cuda()
    dk = torch.zeros((B, cap, 7), dtype=torch.float32, device="cuda")
    dd = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    dn = torch.zeros(B, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ex.extract_batch_device(d_frames.data_ptr(), B, w, h, w, w * h, dk.data_ptr(), dd.data_ptr(), cap, dn.data_ptr())
    torch.cuda.synchronize()
    single = pkg.ORBextractor(800, 1.2, 8, 20, 7, max_width=w, max_height=h)
    for f in range(B):
        sk, sd = single(frames[f])
        assert n[f] == len(sk) == int(dn[f])
        assert np.array_equal(kps[f, :n[f]].view(np.uint8), sk.view(np.uint8)), f
        assert np.array_equal(desc[f, :n[f]], sd), f
        assert np.array_equal(dk[f, :n[f]].cpu().numpy().view(np.uint8).reshape(-1), sk.view(np.uint8).reshape(-1)), f
        assert np.array_equal(dd[f, :n[f]].cpu().numpy(), sd), f


def test_edge_inputs(pkg, oracle):
    """Flat images give no keypoints (the reference returns early with empty outputs, R/src/ORBextractor.cpp:1139-1146);
    dense noise, a checkerboard, one level, very few / very many features agree with the oracle; frames whose pyramid
    gets smaller than the reference's 30-pixel FAST grid are refused with an error instead of the reference's division
    by zero (:844-847)."""
    for value in (0, 255):
        ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
        k, d = ex(np.full((480, 640), value, np.uint8))
        assert len(k) == 0 and d is None      # _descriptors.release() (:1145)
    cases = [("noise", np.random.default_rng(1).integers(0, 256, (480, 640), dtype=np.uint8), 1000, 8),
             ("checker", ((np.indices((480, 640)).sum(0) // 4) % 2 * 255).astype(np.uint8), 1000, 8),
             ("one level", pkg.synth.frame(640, 480, 3), 1000, 1),
             ("few", pkg.synth.frame(640, 480, 3), 10, 8),
             ("many", pkg.synth.frame(640, 480, 3), 8000, 8),
             ("odd size", pkg.synth.frame(641, 479, 3), 1000, 8)]
    for name, img, nf, nlev in cases:
        ex = pkg.ORBextractor(nf, 1.2, nlev, 20, 7, max_width=img.shape[1], max_height=img.shape[0])
        k, d = ex(img)
        ok, od = oracle.Extractor(nf, 1.2, nlev, 20, 7)(img)
        assert len(k) == len(ok) > 0, name
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od), name
    for w, h in ((160, 120), (64, 48)):
        with pytest.raises(pkg.OrbxError, match="too small"):     # refused when the handle is created for that shape
            pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=w, max_height=h)
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=640, max_height=480)
    with pytest.raises(pkg.OrbxError, match="too small"):         # ... or when such a frame reaches a larger handle
        ex(pkg.synth.frame(160, 120, 3))
    k, d = ex(pkg.synth.frame(640, 480, 3))                       # the handle stays usable after the refusal
    assert len(k) > 400


@pytest.mark.gpu
def test_extractors_of_different_sizes_on_two_threads(pkg):
    """The stereo Frame constructor runs one extractor per eye on two threads (R/src/Frame.cpp:86-89). The opt-in
    shared-memory limit of a kernel is a property of the process, not of the thread or the handle: a small extractor on
    a second thread must not lower it under a large extractor on the first."""
    import threading
    img = pkg.synth.frame(752, 480, 21)
    big = pkg.ORBextractor(3000, 1.2, 8, 20, 7, max_width=752, max_height=480)
    k0, d0 = big(img)
    out = {}

    def worker():
        small = pkg.ORBextractor(300, 1.2, 4, 20, 7, max_width=752, max_height=480)
        out["small"] = small(img)
        small.close()

    for _ in range(2):
        t = threading.Thread(target=worker)
        t.start()
        t.join()
        k1, d1 = big(img)
        assert np.array_equal(k0.view(np.uint8), k1.view(np.uint8)) and np.array_equal(d0, d1)
    assert len(out["small"][0]) > 100
    big.close()


@pytest.mark.gpu
def test_tapered_chunks_equal_single(pkg):
    """Batches of three chunks or more are cut with a ramp (chunk/4, chunk/2, chunk ..., chunk/2, chunk/4: less exposed
    fill and drain per blocking call); every frame must still get the result of a single-frame call."""
    B = 150
    frames = np.stack([pkg.synth.frame(320, 240, 500 + i) for i in range(B)])
    ex = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240, max_batch=256)
    ex.set_pipeline(chunk_frames=32, lanes=4)   # 8 16 32 32 32 6 16 8
    cap = 700
    kps = pkg.pinned_empty((B, cap), pkg.KP_DTYPE)
    desc = pkg.pinned_empty((B, cap, 32), np.uint8)
    n = np.zeros(B, np.int32)
    single = pkg.ORBextractor(500, 1.2, 8, 20, 7, max_width=320, max_height=240)
    for _ in range(2):   # the second call reuses lanes that last held a chunk of another size
        kps[:] = 0
        ex.extract_batch_into(frames, kps, desc, n)
        for f in list(range(0, B, 7)) + [7, 8, 23, 24, 55, 56, 119, 120, 125, 126, 141, 142, 149]:
            sk, sd = single(frames[f])
            assert n[f] == len(sk)
            assert np.array_equal(kps[f, :n[f]].view(np.uint8), sk.view(np.uint8)), f
            assert np.array_equal(desc[f, :n[f]], sd), f


@pytest.mark.gpu
def test_resize_row_variants_agree(pkg):
    """The pyramid kernel runs with 8 destination rows per thread when a launch fills the GPU and with 2 when it cannot
    (one frame, or the small levels of a small batch): a 48-frame batch (8 rows on the large levels) must give every
    frame the result of a single-frame call (2 rows everywhere)."""
    B = 48
    frames = np.stack([pkg.synth.frame(640, 480, 900 + i) for i in range(B)])
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480, max_batch=B)
    ex.set_pipeline(chunk_frames=B, lanes=1)
    res = ex.extract_batch(list(frames))
    single = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=640, max_height=480)
    for f in range(0, B, 5):
        sk, sd = single(frames[f])
        assert np.array_equal(res[f][0].view(np.uint8), sk.view(np.uint8)) and np.array_equal(res[f][1], sd), f


@pytest.mark.gpu
@pytest.mark.parametrize("w,h", [(640, 480), (1241, 376)])
def test_pyramid_in_one_go_equals_per_level_copies(pkg, w, h):
    """orbx_get_pyramid (what the C++ ORBextractor fills mvImagePyramid from) against orbx_get_level, for a single frame
    and for a frame in the middle of a batch."""
    imgs = [pkg.synth.frame(w, h, 300 + i) for i in range(3)]
    ex = pkg.ORBextractor(1000, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=4)
    ex(imgs[0])
    pyr = ex.image_pyramid()
    assert pyr[0] is None
    for l in range(1, 8):
        assert np.array_equal(pyr[l], ex.image_pyramid_level(l)), l
    ex.extract_batch(imgs)
    pyr = ex.image_pyramid(frame=1)
    for l in range(1, 8):
        assert np.array_equal(pyr[l], ex.image_pyramid_level(l, frame=1)), l
    with pytest.raises(Exception):
        ex.image_pyramid(frame=3)
