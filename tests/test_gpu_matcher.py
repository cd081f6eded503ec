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


@pytest.mark.parametrize("nshards", [2, 3, 8])
def test_sharded_bf2_merge_equals_whole_database(pkg, oracle, nshards):
    """Database split into contiguous row ranges (what each GPU of the box would hold), per-shard top-2 with
    idx_base, then orbx_match_merge_device: must equal the single-pass answer, duplicates included."""
    import torch
    ndb, nq = 50_000, 777
    db = pkg.synth.descriptors(ndb, 11)
    q = pkg.synth.planted_queries(db, nq, 12)
    db[ndb - 5] = db[3]
    q[0] = db[3]          # equal minimum in the first and the last shard
    m = pkg.ORBmatcher(0.6)
    dq = torch.from_numpy(q).cuda()
    ddb = torch.from_numpy(db).cuda()
    parts = torch.empty((3, nshards, nq), dtype=torch.int32, device="cuda")
    s = torch.cuda.current_stream().cuda_stream
    for r in range(nshards):
        lo, hi = pkg.sharding.block_range(ndb, r, nshards)
        m.match_bf2_device(dq.data_ptr(), nq, ddb[lo:hi].data_ptr(), hi - lo, lo, parts[0, r].data_ptr(),
                           parts[1, r].data_ptr(), parts[2, r].data_ptr(), s)
    out = torch.empty((3, nq), dtype=torch.int32, device="cuda")
    acc = torch.empty(nq, dtype=torch.uint8, device="cuda")
    m.merge_device(parts[0].data_ptr(), parts[1].data_ptr(), parts[2].data_ptr(), nshards, nq, 50,
                   out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(), acc.data_ptr(), s)
    torch.cuda.synchronize()
    eidx, ed1, ed2, eacc = oracle.match_bf2(q, db, np.float32(0.6), 50)
    o = out.cpu().numpy()
    assert np.array_equal(o[0], eidx) and np.array_equal(o[1], ed1) and np.array_equal(o[2], ed2)
    assert np.array_equal(acc.cpu().numpy(), eacc)
    assert o[0][0] == 3 and o[1][0] == 0 and o[2][0] == 0
