"""Known-answer and property tests of the extractor oracle (no reference tests exist)."""
import numpy as np


def test_ctor_tables_known_answers(oracle):
    # constants derived from R/src/ORBextractor.cpp:435-500
    t = oracle.Extractor(1000, 1.2, 8, 20, 7).tables()
    assert t["feats"].tolist() == [217, 181, 151, 126, 105, 87, 73, 60]
    assert t["umax"].tolist() == [15, 15, 15, 15, 14, 14, 14, 13, 13, 12, 11, 10, 9, 8, 6, 3]
    exp = np.array([1, 1.2000000477, 1.4400000572, 1.7280001640, 2.0736002922, 2.4883203506, 2.9859845638,
                    3.5831816196], np.float32)
    assert np.array_equal(t["scale"], exp)
    assert oracle.Extractor(2000, 1.2, 8, 20, 7).tables()["feats"].tolist() == [434, 362, 302, 251, 209, 175, 145, 122]
    assert oracle.Extractor(1200, 1.2, 8, 20, 7).tables()["feats"].tolist() == [261, 217, 181, 151, 126, 105, 87, 72]
    assert oracle.Extractor(4000, 1.2, 8, 20, 7).tables()["feats"].tolist() == [869, 724, 603, 503, 419, 349, 291, 242]


def test_level_sizes(oracle, pkg):
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    ex(pkg.synth.frame(640, 480, 0))
    sizes = [ex.level(l).shape[::-1] for l in range(8)]
    assert sizes == [(640, 480), (533, 400), (444, 333), (370, 278), (309, 231), (257, 193), (214, 161), (179, 134)]


def test_extract_properties(oracle, pkg):
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    img = pkg.synth.frame(640, 480, 9)
    k, d = ex(img)
    feats = ex.tables()["feats"]
    assert d.shape == (len(k), 32)
    assert np.all(k["class_id"] == -1)
    assert np.all(np.diff(k["octave"]) >= 0), "output is level-major"
    for l in range(8):
        kept, cands = ex.kept(l), len(ex.candidates(l))
        assert kept <= cands
        if cands >= feats[l] * 2:
            assert feats[l] <= kept <= feats[l] + 3   # stops at size>=N, a split adds at most 3
    assert np.all((k["angle"] >= 0) & (k["angle"] < 360))
    sc = ex.tables()["scale"]
    assert np.array_equal(k["size"], np.trunc(np.float32(31) * sc[k["octave"]]))
    # idempotence: same image, same result
    k2, d2 = ex(img)
    assert np.array_equal(k.view(np.uint8), k2.view(np.uint8)) and np.array_equal(d, d2)


def test_edge_inputs(oracle, pkg):
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    k, d = ex(pkg.synth.frame(640, 480, 0, "constant"))
    assert len(k) == 0
    k, d = ex(pkg.synth.frame(640, 480, 5, "lowcontrast"))   # minThFAST fallback cells
    assert 0 < len(k) < 1000


def test_tie_break_sensitivity_is_measurable(oracle, pkg):
    """The reference sorts (size, heap pointer) (R/src/ORBextractor.cpp:736); the two canonical extremes
    (creation order ascending / descending) bracket what it can produce."""
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    img = pkg.synth.frame(640, 480, 0)
    k0, _ = ex(img)
    ties = [ex.tie_sensitive(l) for l in range(8)]
    ex.set_tie_mode(1)
    k1, _ = ex(img)
    assert any(ties)
    assert abs(len(k0) - len(k1)) <= 16
    lv = [l for l in range(8) if ties[l] == 0]
    for l in lv:   # tie-insensitive levels keep the same SET under either order
        a = set(map(tuple, np.stack([k0["x"], k0["y"]], 1)[k0["octave"] == l].tolist()))
        b = set(map(tuple, np.stack([k1["x"], k1["y"]], 1)[k1["octave"] == l].tolist()))
        assert a == b


def test_distribute_standalone_matches_extractor(oracle, pkg):
    ex = oracle.Extractor(1000, 1.2, 8, 20, 7)
    k, _ = ex(pkg.synth.frame(640, 480, 3))
    c = ex.candidates(0)
    idx = oracle.distribute(c, 16, 640 - 16, 16, 480 - 16, 217)
    k0 = k[k["octave"] == 0]
    assert np.array_equal(c[idx, 0] + 16, k0["x"].astype(np.int32))
    assert np.array_equal(c[idx, 1] + 16, k0["y"].astype(np.int32))
    assert len(oracle.distribute(np.zeros((0, 3), np.int32), 16, 624, 16, 464, 217)) == 0


def test_matcher_oracle_basics(oracle, pkg):
    db = pkg.synth.descriptors(3000, 5)
    q = pkg.synth.planted_queries(db, 64, 6)
    idx, d1, d2, acc = oracle.match_bf2(q, db, np.float32(0.6), 50, nthreads=2)
    # brute-force numpy check of the definition
    x = np.unpackbits(q[:, None, :] ^ db[None, :, :], axis=2).sum(2)
    assert np.array_equal(idx, x.argmin(1))
    assert np.array_equal(d1, x.min(1))
    assert np.array_equal(d2, np.sort(x, 1)[:, 1])
    assert np.array_equal(acc.astype(bool), (d1 <= 50) & (d1.astype(np.float32) < np.float32(0.6) * d2.astype(np.float32)))
    assert oracle.descriptor_distance(np.zeros(32, np.uint8), np.full(32, 255, np.uint8)) == 256
