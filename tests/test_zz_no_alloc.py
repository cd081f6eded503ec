"""orbx_match_bf2_noalloc (static device scratch, no allocation in the call) against the existing entry points
orbx_match_bf2_device + orbx_match_merge_device, through the C ABI. Collected last on purpose (file name).
Integer results: the tolerance is equality, as in tests/test_gpu_matcher.py."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _existing(pkg, m, dq, nq, ddb, ndb, th):
    import torch
    s = torch.cuda.current_stream().cuda_stream
    part = torch.empty((3, nq), dtype=torch.int32, device="cuda")
    out = torch.empty((3, nq), dtype=torch.int32, device="cuda")
    acc = torch.empty(nq, dtype=torch.uint8, device="cuda")
    m.match_bf2_device(dq.data_ptr(), nq, ddb.data_ptr(), ndb, 0, part[0].data_ptr(), part[1].data_ptr(),
                       part[2].data_ptr(), s)
    m.merge_device(part[0].data_ptr(), part[1].data_ptr(), part[2].data_ptr(), 1, nq, th, out[0].data_ptr(),
                   out[1].data_ptr(), out[2].data_ptr(), acc.data_ptr(), s)
    torch.cuda.synchronize()
    return out.cpu().numpy(), acc.cpu().numpy()


def _noalloc(pkg, dq, nq, ddb, ndb, ratio, th):
    import torch
    from orbslam2_b200 import _capi
    s = torch.cuda.current_stream().cuda_stream
    # outputs are the caller's; poisoned so that a call that writes nothing cannot pass
    out = torch.full((3, nq), -7, dtype=torch.int32, device="cuda")
    acc = torch.full((nq,), 9, dtype=torch.uint8, device="cuda")
    _capi.check(_capi.lib().orbx_match_bf2_noalloc(0, dq.data_ptr(), nq, ddb.data_ptr(), ndb, C.c_float(ratio), th,
                                                   out[0].data_ptr(), out[1].data_ptr(), out[2].data_ptr(),
                                                   acc.data_ptr(), s))
    torch.cuda.synchronize()
    return out.cpu().numpy(), acc.cpu().numpy()


def test_noalloc_equals_existing_entry_two_calls_in_a_row(pkg):
    import torch
    m = pkg.ORBmatcher(0.6)
    # two different problems back to back: the second call finds the first one's partials in the shared array
    # (more chunks x queries in the first than in the second) and must not be affected by them
    for nq, ndb, seed in [(3000, 200_000, 21), (1025, 20_000, 23)]:
        db = pkg.synth.descriptors(ndb, seed)
        q = pkg.synth.planted_queries(db, nq, seed + 1)
        db[300] = db[100]
        q[0] = db[100]
        dq = torch.from_numpy(q).cuda()
        ddb = torch.from_numpy(db).cuda()
        exp, eacc = _existing(pkg, m, dq, nq, ddb, ndb, 50)
        got, gacc = _noalloc(pkg, dq, nq, ddb, ndb, 0.6, 50)
        assert np.array_equal(got, exp)
        assert np.array_equal(gacc, eacc)
        assert got[0][0] == 100 and got[1][0] == 0 and got[2][0] == 0 and gacc[0] == 0
        assert int(gacc.sum()) > 0
        # and the very same call again
        got2, gacc2 = _noalloc(pkg, dq, nq, ddb, ndb, 0.6, 50)
        assert np.array_equal(got2, exp) and np.array_equal(gacc2, eacc)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(pkg):
    import torch
    from orbslam2_b200 import _capi
    L = _capi.lib()
    d = torch.zeros(64, dtype=torch.uint8, device="cuda")
    o = torch.zeros(4, dtype=torch.int32, device="cuda")
    rc = L.orbx_match_bf2_noalloc(0, d.data_ptr(), 16385, d.data_ptr(), 2, C.c_float(0.6), 50, o.data_ptr(),
                                  o.data_ptr(), o.data_ptr(), None, None)
    assert rc == -1
    assert b"exceed the static scratch limit" in L.orbx_last_error()
    torch.cuda.synchronize()
