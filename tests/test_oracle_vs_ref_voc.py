"""Pins the oracle's restatement of the vocabulary transform against the REFERENCE'S OWN vendored
DBoW2 (Thirdparty/DBoW2: the real TemplatedVocabulary.h, FORB.cpp, BowVector.cpp, FeatureVector.cpp, ScoringObject.cpp
compiled where they lie into oracle/_ref/libref_voc.so): both load the same synthetic vocabulary text file
(tests/voc_case.py; ORBvoc.txt is not shipped) and transform the same descriptors; words, values (bit for bit) and
feature vectors must be identical."""
import numpy as np
import pytest

import oracle_lib as ol
import ref_voc_lib as rv
import voc_case

pytestmark = pytest.mark.skipif(not rv.available(), reason="oracle/_ref/libref_voc.so not built")


def same(a, b):
    return all(x.dtype == y.dtype and np.array_equal(x.view(np.uint8), y.view(np.uint8)) for x, y in zip(a, b))


@pytest.mark.parametrize("k,L,levelsup,scoring,weighting", [(10, 4, 2, 0, 0), (10, 4, 1, 0, 0), (10, 4, 4, 0, 0), (6, 5, 3, 0, 1),
                                                            (10, 3, 1, 5, 0), (8, 4, 2, 0, 2), (8, 4, 2, 2, 3), (9, 4, 2, 1, 0)])
def test_transform_vs_reference(tmp_path, k, L, levelsup, scoring, weighting):
    path = tmp_path / "voc.txt"
    n_nodes, n_words = voc_case.write(path, k, L, scoring, weighting, seed=k * 10 + L, min_leaf_level=max(L - levelsup, 1) + 1)
    ref, orc = rv.Vocabulary(path), ol.Vocabulary(path)
    assert ref.words() == orc.words() == n_words
    for n in (2000, 37, 1, 0):
        feats = voc_case.features(max(n, 1), path, seed=n + 3)[:n]
        r = ref.transform(feats, levelsup)
        o = orc.transform(feats, levelsup)
        assert same(r, o), (n, [len(x) for x in r], [len(x) for x in o])
        if n == 2000:
            assert len(r[0]) > 300 and len(r[2]) > (0 if L - levelsup <= 0 else 5)
            assert len(r[4]) < n      # some features fell on zero-weight words
            if scoring == 0:
                assert abs(r[1].sum() - 1.0) < 1e-12
