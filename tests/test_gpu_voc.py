"""SURVEY.md 8f row 4 on the device: the vocabulary descent of DBoW2's transform (orbx_vocabulary_descend) with the
C++ host assembly of BowVector / FeatureVector (orb-slam2-_b200/host/orbx_bow.cc), against the oracle and, where
oracle/_ref/libref_voc.so travelled to this box, against the reference's own vendored DBoW2 on the same vocabulary
file and descriptors: words, values (bit for bit) and feature vectors identical."""
import ctypes as C
from pathlib import Path

import numpy as np
import pytest

import ref_voc_lib as rv
import voc_case

pytestmark = pytest.mark.gpu
LIB = Path(__file__).resolve().parents[1] / "orb-slam2-_b200" / "host" / "libmirror_test.so"
P = C.c_void_p


@pytest.fixture(scope="module")
def mirror():
    L = C.CDLL(str(LIB))
    L.mirror_voc_load.restype = P
    L.mirror_voc_load.argtypes = [P, C.c_char_p]
    L.mirror_voc_destroy.argtypes = [P]
    L.mirror_voc_words.argtypes = [P]
    L.mirror_voc_transform.argtypes = [P, P, C.c_int, C.c_int] + [P] * 7
    return L


def same(a, b):
    return all(x.dtype == y.dtype and np.array_equal(x.view(np.uint8), y.view(np.uint8)) for x, y in zip(a, b))


@pytest.mark.parametrize("k,L,levelsup,scoring,weighting", [(10, 4, 2, 0, 0), (10, 5, 4, 0, 0), (10, 4, 4, 0, 0), (6, 5, 3, 0, 1),
                                                            (10, 3, 1, 5, 0), (8, 4, 2, 0, 2), (8, 4, 2, 2, 3), (9, 4, 2, 1, 0),
                                                            (20, 3, 1, 0, 0)])
def test_transform(tmp_path, pkg, oracle, mirror, k, L, levelsup, scoring, weighting):
    path = tmp_path / "voc.txt"
    _, n_words = voc_case.write(path, k, L, scoring, weighting, seed=k * 10 + L, min_leaf_level=max(L - levelsup, 1) + 1)
    m = pkg.ORBmatcher(0.6)
    h = mirror.mirror_voc_load(m._h, str(path).encode())
    assert h and mirror.mirror_voc_words(h) == n_words
    orc = oracle.Vocabulary(path)
    ref = rv.Vocabulary(path) if rv.available() else None
    for n in (2000, 37, 1, 0):
        feats = voc_case.features(max(n, 1), path, seed=n + 3)[:n]
        g = rv.flat_transform(mirror.mirror_voc_transform, h, feats, levelsup)
        assert same(g, orc.transform(feats, levelsup)), n
        if ref is not None:
            assert same(g, ref.transform(feats, levelsup)), n
        if n == 2000:
            assert len(g[0]) > 300 and len(g[4]) < n
    mirror.mirror_voc_destroy(h)


def test_descent_on_extracted_descriptors(tmp_path, pkg, oracle):
    """Frame::ComputeBoW's input: the descriptors of a frame (levelsup 4 of a 6-level tree, the reference's setting)."""
    path = tmp_path / "voc6.txt"
    voc_case.write(path, 5, 6, 0, 0, seed=77, min_leaf_level=3)
    ex = pkg.ORBextractor(1500, 1.2, 8, 20, 7, max_width=640, max_height=480)
    _, desc = ex(pkg.synth.frame(640, 480, 12))
    orc = oracle.Vocabulary(path)
    parent, ndesc, weight, word, klsw = orc.export()
    v = pkg.Vocabulary(pkg.ORBmatcher(0.6), parent, ndesc, int(klsw[1]))
    leaf, level = v.descend(desc, 4)
    words, values, nodes, off, idx = orc.transform(desc, 4)
    keep = weight[leaf] > 0
    assert set(word[leaf[keep]].tolist()) == set(words.tolist())
    for a, node in enumerate(nodes):
        assert np.array_equal(np.nonzero(keep & (level == node))[0], idx[off[a]:off[a + 1]])
    with pytest.raises(pkg.OrbxError):   # a parent that does not precede its child
        pkg.Vocabulary(pkg.ORBmatcher(0.6), np.array([0, 2, 0], np.int32), np.zeros((3, 32), np.uint8), 2)
