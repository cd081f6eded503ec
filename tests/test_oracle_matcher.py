"""Known answers and properties of the matcher-side oracle (CPU): Frame grid, SearchForInitialization,
SearchByProjection, ComputeStereoMatches, DescriptorDistance. (These rows are pinned to the reference's own
ORBmatcher.cpp / Frame.cpp in tests/test_oracle_vs_ref_front.py; the checks here need no reference build.)"""
import numpy as np


def _feat(pkg, oracle, w, h, nf, seed, shift=(0, 0)):
    img = pkg.synth.frame(w, h, seed)
    if shift != (0, 0):
        img = np.roll(img, shift, (0, 1))
    ex = oracle.Extractor(nf, 1.2, 8, 20, 7)
    k, d = ex(img)
    return ex, img, k, d


def test_descriptor_distance_is_popcount(pkg, oracle):
    a, b = pkg.synth.descriptors(200, 1), pkg.synth.descriptors(200, 2)
    for i in range(200):
        assert oracle.descriptor_distance(a[i], b[i]) == int(np.unpackbits(a[i] ^ b[i]).sum())


def test_grid_query_equals_brute_force_as_a_set(pkg, oracle):
    _, _, k, _ = _feat(pkg, oracle, 640, 480, 1000, 3)
    b = (0.0, 640.0, 0.0, 480.0)
    g = oracle.Grid(k, b)
    rng = np.random.default_rng(0)
    for _ in range(200):
        x, y = np.float32(rng.uniform(0, 640)), np.float32(rng.uniform(0, 480))
        r = np.float32(rng.choice([5, 20, 60]))
        lo = int(rng.integers(-1, 6))
        hi = int(lo + rng.integers(0, 3)) if rng.random() < 0.7 else -1
        got = g.query(float(x), float(y), float(r), lo, hi)
        ok = (np.abs(k["x"] - x) < r) & (np.abs(k["y"] - y) < r)
        if lo > 0 or hi >= 0:
            ok &= k["octave"] >= lo
            if hi >= 0:
                ok &= k["octave"] <= hi
        # PosInGrid rounds, so a keypoint can sit one cell outside the floor/ceil cell window: the grid never
        # returns a point outside the box, and misses only points within one cell of its border
        assert set(got.tolist()) <= set(np.nonzero(ok)[0].tolist())
        inner = ok & (np.abs(k["x"] - x) < r - 10) & (np.abs(k["y"] - y) < r - 10)
        assert set(np.nonzero(inner)[0].tolist()) <= set(got.tolist())
        assert len(set(got.tolist())) == len(got)


def test_search_for_initialization_on_identical_frames(pkg, oracle):
    _, _, k, d = _feat(pkg, oracle, 640, 480, 1000, 4)
    b = (0.0, 640.0, 0.0, 480.0)
    pm = np.stack([k["x"], k["y"]], 1).astype(np.float32)
    n, m12, pm2 = oracle.search_for_initialization(k, d, k, d, b, pm, 100, np.float32(0.9))
    lvl0 = np.nonzero(k["octave"] == 0)[0]
    assert (m12[k["octave"] > 0] == -1).all()
    matched = m12[lvl0]
    assert n == (m12 >= 0).sum() and n > 0.8 * len(lvl0)
    assert (matched[matched >= 0] == lvl0[matched >= 0]).all()      # a keypoint matches itself (distance 0)
    assert np.array_equal(pm2, pm)


def test_stereo_matches_recover_a_constant_disparity(pkg, oracle):
    w, h, disp = 752, 480, 24
    left = pkg.synth.frame(w, h, 8)
    right = np.empty_like(left)
    right[:, :w - disp] = left[:, disp:]
    right[:, w - disp:] = left[:, w - 1:w]
    exl, exr = oracle.Extractor(1200, 1.2, 8, 20, 7), oracle.Extractor(1200, 1.2, 8, 20, 7)
    kl, dl = exl(left)
    kr, dr = exr(right)
    t = exl.tables()
    mb = np.float32(0.11)
    mbf = np.float32(435.2) * mb
    u, z, n = oracle.stereo_matches(kl, dl, kr, dr, [exl.level(l) for l in range(8)], [exr.level(l) for l in range(8)],
                                    t["scale"], t["inv_scale"], mbf, mb)
    ok = u >= 0
    assert n == ok.sum() and n > 0.3 * len(kl)
    d = (kl["x"] - u)[ok]
    assert np.abs(np.median(d) - disp) < 0.5 and (np.abs(d - disp) < 3 * 1.2 ** 7).all()
    assert np.allclose(z[ok], mbf / d, rtol=1e-6)
    assert (z[~ok] == -1).all()
