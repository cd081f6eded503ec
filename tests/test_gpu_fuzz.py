"""Randomised shapes and parameters: the CUDA extractor against the oracle on frame sizes, pyramid depths, scale
factors, feature budgets and FAST thresholds nobody tuned the kernels for (cell widths from 30 to 59 pixels, odd
widths, unaligned pitches, levels close to the minimum size)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cases():
    rng = np.random.default_rng(2024)
    out = []
    for i in range(18):
        w = int(rng.integers(200, 900))
        h = int(rng.integers(160, 700))
        scale = float(rng.choice([1.1, 1.2, 1.3, 1.5, 2.0]))
        if w < 0.6 * h:  # the reference's octree needs round(width/height) >= 1 (R/src/ORBextractor.cpp:577)
            w = int(0.6 * h) + 1
        nlev = int(rng.integers(1, 9))
        # keep the smallest level above the reference's 30-pixel grid minimum
        while nlev > 1 and min(w, h) / scale ** (nlev - 1) < 75:
            nlev -= 1
        # ... and the same on the border-stripped extent of every level
        while nlev > 1 and (w / scale ** (nlev - 1) - 38) < 0.55 * (h / scale ** (nlev - 1) - 38):
            nlev -= 1
        nf = int(rng.choice([50, 300, 1000, 2500]))
        ini = int(rng.choice([12, 20, 40]))
        mn = int(rng.choice([3, 7, ini]))
        kind = str(rng.choice(["shapes", "noise", "lowcontrast"]))
        out.append((w, h, scale, nlev, nf, ini, mn, kind, 500 + i))
    return out


@pytest.mark.parametrize("w,h,scale,nlev,nf,ini,mn,kind,seed", _cases())
def test_random_configuration_matches_oracle(pkg, oracle, w, h, scale, nlev, nf, ini, mn, kind, seed):
    img = pkg.synth.frame(w, h, seed, kind=kind)
    ex = pkg.ORBextractor(nf, scale, nlev, ini, mn, max_width=w, max_height=h)
    k, d = ex(img)
    ok, od = oracle.Extractor(nf, scale, nlev, ini, mn)(img)
    assert len(k) == len(ok)
    if len(ok):
        assert np.array_equal(k.view(np.uint8), ok.view(np.uint8))
        assert np.array_equal(d, od)
    # a strided view (rows of a wider buffer) must give the same result
    wide = np.zeros((h, w + 13), np.uint8)
    wide[:, 5:5 + w] = img
    k2, d2 = ex(wide[:, 5:5 + w])
    assert len(k2) == len(ok) and (len(ok) == 0 or (np.array_equal(k2.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d2, od)))


def test_portrait_frame_without_octree_roots_is_refused(pkg):
    """round(width/height) == 0 makes the reference divide by zero (R/src/ORBextractor.cpp:577-580); the library
    says so at handle creation instead of launching anything."""
    with pytest.raises(pkg.OrbxError, match="octree roots"):
        pkg.ORBextractor(1000, 1.5, 3, 20, 7, max_width=246, max_height=686)


def test_single_level_with_the_whole_budget(pkg, oracle):
    """2500 features on ONE level: the largest per-level quota the fuzz draws (node table 3400)."""
    img = pkg.synth.frame(640, 480, 77, kind="noise")
    k, d = pkg.ORBextractor(3300, 1.2, 1, 20, 7, max_width=640, max_height=480)(img)
    ok, od = oracle.Extractor(3300, 1.2, 1, 20, 7)(img)
    assert len(k) == len(ok) and len(ok) > 2500
    assert np.array_equal(k.view(np.uint8), ok.view(np.uint8)) and np.array_equal(d, od)
    with pytest.raises(pkg.OrbxError, match="node table"):
        pkg.ORBextractor(3500, 1.2, 1, 20, 7, max_width=640, max_height=480)
