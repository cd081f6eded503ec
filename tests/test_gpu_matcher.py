"""GPU parity: Hamming distance and brute-force top-2 + ratio against the CPU oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_descriptor_distance(pkg, oracle):
    a = pkg.synth.descriptors(5000, 1)
    b = pkg.synth.descriptors(5000, 2)
    b[:10] = a[:10]
    a[10:20] = 0
    b[10:20] = 255
    m = pkg.ORBmatcher()
    d = m.DescriptorDistance(a, b)
    exp = np.array([oracle.descriptor_distance(a[i], b[i]) for i in range(5000)], np.int32)
    assert np.array_equal(d, exp)
    assert d[:10].tolist() == [0] * 10 and d[10:20].tolist() == [256] * 10


@pytest.mark.parametrize("nq,ndb", [(1, 1), (7, 2), (1000, 513), (1025, 20000), (3000, 100000)])
def test_bf2_bit_exact(pkg, oracle, nq, ndb):
    db = pkg.synth.descriptors(ndb, 3)
    q = pkg.synth.planted_queries(db, nq, 4)
    if ndb > 600:
        db[300] = db[100]      # duplicate rows: first index must win, second == best
        q[0] = db[100]
    m = pkg.ORBmatcher(0.6)
    idx, d1, d2, acc = m.match_bf2(q, db)
    eidx, ed1, ed2, eacc = oracle.match_bf2(q, db, np.float32(0.6), 50)
    assert np.array_equal(idx, eidx)
    assert np.array_equal(d1, ed1)
    assert np.array_equal(d2, ed2)
    assert np.array_equal(acc, eacc)
    if ndb > 600:
        assert idx[0] == 100 and d1[0] == 0 and d2[0] == 0 and acc[0] == 0
