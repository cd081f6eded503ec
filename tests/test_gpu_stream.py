"""GPU tests of the sequence-level entry points: the ticket ring / lane rotor of orbx_extract_batch_submit under a long
uneven load, the stereo streaming form, the per-frame overflow report of the device-resident entry, and the
orbx_shard_* group (one process; a second device is used when the box has one)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _variants(base, count, seed):
    """count distinct frames from a few base frames: a roll of its own + +-1 noise per copy (cheap, deterministic)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    nb = base.shape[0]
    out = np.empty((count,) + base.shape[1:], np.uint8)
    for i in range(count):
        fr = np.roll(base[i % nb], (3 * (i // nb) + 1, 5 * (i // nb) + 2), (0, 1)).astype(np.int16)
        fr += rng.integers(-1, 2, fr.shape, dtype=np.int16)
        out[i] = np.clip(fr, 0, 255)
    return out


def test_soak_2000_frames_40_uneven_submissions(pkg):
    """>= 2000 distinct frames through >= 40 submissions of uneven sizes, up to 3 tickets in flight, chunks of 24
    frames over 3 lanes (so the rotor wraps at a different phase in every submission): every frame must equal the
    single-frame call (ticket ring and lane rotor of csrc/orbx_api.cu)."""
    w, h, nf = 320, 240, 500
    base = np.stack([pkg.synth.frame(w, h, 300 + i) for i in range(16)])
    total = 2048
    host = pkg.pinned_empty((total, h, w), np.uint8)
    host[:] = _variants(base, total, 1)
    cap = nf + 64
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=128)
    ex.set_pipeline(24, 3)
    kps = pkg.pinned_empty((total, cap), pkg.KP_DTYPE)
    desc = pkg.pinned_empty((total, cap, 32), np.uint8)
    n = pkg.pinned_empty(total, np.int32)
    n[:] = -5
    rng = np.random.Generator(np.random.PCG64(2))
    f0, subs, pending = 0, 0, []
    while f0 < total:
        cnt = int(min(total - f0, rng.integers(1, 100)))
        if len(pending) == 3:
            ex.extract_batch_wait(pending.pop(0))
        pending.append(ex.extract_batch_submit(host[f0:f0 + cnt], kps[f0:f0 + cnt], desc[f0:f0 + cnt], n[f0:f0 + cnt]))
        f0 += cnt
        subs += 1
    for t in pending:
        ex.extract_batch_wait(t)
    assert subs >= 40
    one = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    for f in range(total):
        k1, d1 = one(host[f])
        assert n[f] == len(k1) > 100, f
        assert np.array_equal(kps[f, :n[f]].view(np.uint8), k1.view(np.uint8)), f
        assert np.array_equal(desc[f, :n[f]], d1), f
    # waiting for a ticket twice / for one never issued is a state error, not a hang
    with pytest.raises(pkg.OrbxError):
        ex.extract_batch_wait(10_000)


def test_stereo_streaming_equals_per_pair_calls(pkg):
    """orbx_stereo_batch_submit: L R L R ... through uneven submissions; keypoints equal the single-frame call and
    mvuRight / mvDepth equal orbx_stereo_match on two extractor handles (the form the drop-in Frame constructor uses)."""
    w, h, nf, pairs = 752, 480, 1200, 9
    imgs = [pkg.synth.stereo_pair(w, h, 40 + p) for p in range(pairs)]
    B = 2 * pairs
    host = pkg.pinned_empty((B, h, w), np.uint8)
    host[:] = np.stack([im for lr in imgs for im in lr])
    cap = nf + 64
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=8)
    ex.set_pipeline(4, 3)
    kps = pkg.pinned_empty((B, cap), pkg.KP_DTYPE)
    desc = pkg.pinned_empty((B, cap, 32), np.uint8)
    n = pkg.pinned_empty(B, np.int32)
    ur = pkg.pinned_empty((pairs, cap), np.float32)
    dp = pkg.pinned_empty((pairs, cap), np.float32)
    nm = pkg.pinned_empty(pairs, np.int32)
    mbf, mb = np.float32(435.2 * 0.11), np.float32(0.0)
    tickets = []
    for p0, pc in [(0, 4), (4, 1), (5, 4)]:
        s = slice(2 * p0, 2 * (p0 + pc))
        tickets.append(ex.stereo_batch_submit(host[s], kps[s], desc[s], n[s], mbf, mb, ur[p0:p0 + pc], dp[p0:p0 + pc],
                                              nm[p0:p0 + pc]))
    for t in tickets:
        ex.extract_batch_wait(t)
    with pytest.raises(pkg.OrbxError):   # an odd number of frames is not a stereo sequence
        ex.stereo_batch_submit(host[:3], kps[:3], desc[:3], n[:3], mbf, mb, ur[:1], dp[:1], nm[:1])
    m = pkg.ORBmatcher()
    exl = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    for p in range(pairs):
        kl, dl = exl(host[2 * p])
        kr, dr = exr(host[2 * p + 1])
        assert n[2 * p] == len(kl) and n[2 * p + 1] == len(kr)
        assert np.array_equal(kps[2 * p, :len(kl)].view(np.uint8), kl.view(np.uint8))
        assert np.array_equal(desc[2 * p + 1, :len(kr)], dr)
        u, z, cnt = m.stereo_match(exl, exr, kl, dl, kr, dr, mbf, mb)
        assert cnt == nm[p] > 50
        assert np.array_equal(ur[p, :len(kl)].view(np.uint32), u.view(np.uint32))
        assert np.array_equal(dp[p, :len(kl)].view(np.uint32), z.view(np.uint32))


def test_device_entry_reports_candidate_overflow_per_frame(pkg):
    """A lattice of isolated bright pixels (tile [[0,0,255,0],[255,0,0,0]]) keeps one FAST corner per 4 pixels after
    non-maximum suppression: more than the candidate capacity of one per 8 pixels. The host-buffer entry fails with
    ORBX_ERR_CAPACITY; the device-resident entry marks exactly the overflowing frame with n = -1."""
    import torch
    w, h, nf = 640, 480, 1000
    dense = np.tile(np.array([[0, 0, 255, 0], [255, 0, 0, 0]], np.uint8), (h // 2, w // 4))
    frames = np.stack([pkg.synth.frame(w, h, 5), dense, pkg.synth.frame(w, h, 7)])
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=3)
    cap = nf + 64
    dk = torch.zeros((3, cap, 7), dtype=torch.float32, device="cuda")
    dd = torch.zeros((3, cap, 32), dtype=torch.uint8, device="cuda")
    dn = torch.zeros(3, dtype=torch.int32, device="cuda")
    df = torch.from_numpy(frames).cuda()
    ex.extract_batch_device(df.data_ptr(), 3, w, h, w, w * h, dk.data_ptr(), dd.data_ptr(), cap, dn.data_ptr())
    torch.cuda.synchronize()
    n = dn.cpu().numpy()
    assert n[1] == -1 and n[0] > 0 and n[2] > 0, n
    with pytest.raises(pkg.OrbxError) as e:
        ex(frames[1])
    assert e.value.code == -3
    k0, _ = ex(frames[0])          # the handle stays usable, and the good frames of the batch were complete
    assert len(k0) == n[0]


def _devices(pkg, want):
    if pkg.device_count() < want:
        pytest.skip(f"needs {want} GPUs")
    return list(range(want))


@pytest.mark.parametrize("ndev", [1, 2])
def test_shard_group_match_equals_single_gpu(pkg, oracle, ndev):
    """orbx_shard_match_bf2 (C++ host: per-device shards, ncclAllGather, in-place merge) == orbx_match_bf2 on the whole
    database == the oracle, duplicates across shards included."""
    devs = _devices(pkg, ndev)
    ndb, nq = 200_000, 1500
    db = pkg.synth.descriptors(ndb, 31)
    q = pkg.synth.planted_queries(db, nq, 32)
    db[ndb - 5] = db[3]
    q[0] = db[3]               # equal minimum in the first and the last shard
    g = pkg.sharding.ShardGroup(devs)
    info = g.info()
    assert info["nranks"] == ndev and info["nlocal"] == ndev
    g.set_database(db)
    got = g.match_bf2(q, 0.6, 50)
    got2 = g.match_bf2(q, 0.6, 50)
    exp = pkg.ORBmatcher(0.6).match_bf2(q, db)
    ora = oracle.match_bf2(q, db, np.float32(0.6), 50)
    for a, b, c, d in zip(got, got2, exp, ora):
        assert np.array_equal(a, c) and np.array_equal(b, c) and np.array_equal(a, d)
    assert got[0][0] == 3 and got[1][0] == 0 and got[2][0] == 0
    if ndev > 1:
        assert g.info()["nccl_version"] > 0
    g.close()


@pytest.mark.parametrize("ndev", [1, 2])
@pytest.mark.parametrize("stereo", [False, True])
def test_shard_group_sequence_in_order(pkg, ndev, stereo):
    """orbx_shard_extract_sequence: pair-aligned blocks per device, results in sequence order, equal to single calls."""
    devs = _devices(pkg, ndev)
    w, h, nf = 320, 240, 400
    if stereo:
        base = np.stack([im for p in range(4) for im in pkg.synth.stereo_pair(w, h, 60 + p)])
        count = 26
        seq = np.stack([base[(2 * (i // 2)) % 8 + (i & 1)] for i in range(count)])
        seq = np.roll(seq, 2, axis=1)
        for i in range(count):   # same roll for both frames of a pair keeps the disparity
            seq[i] = np.roll(seq[i], (i // 2) % 5, axis=0)
    else:
        count = 37
        seq = _variants(np.stack([pkg.synth.frame(w, h, 80 + i) for i in range(4)]), count, 3)
    host = pkg.pinned_empty((count, h, w), np.uint8)
    host[:] = seq
    cap = nf + 64
    kps = pkg.pinned_empty((count, cap), pkg.KP_DTYPE)
    desc = pkg.pinned_empty((count, cap, 32), np.uint8)
    n = pkg.pinned_empty(count, np.int32)
    n[:] = -5
    params = dict(nfeatures=nf, scale_factor=1.2, nlevels=8, ini_th_fast=20, min_th_fast=7, max_batch=8)
    g = pkg.sharding.ShardGroup(devs)
    st = None
    if stereo:
        ur = pkg.pinned_empty((count // 2, cap), np.float32)
        dp = pkg.pinned_empty((count // 2, cap), np.float32)
        nm = pkg.pinned_empty(count // 2, np.int32)
        st = (np.float32(40.0), np.float32(0.0), ur, dp, nm)
    g.extract_sequence(params, host, kps, desc, n, group=2 if stereo else 1, stereo=st)
    assert g.launch_count() > 0
    one = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    two = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    m = pkg.ORBmatcher()
    for f in range(count):
        k1, d1 = one(host[f])
        assert n[f] == len(k1) > 50, f
        assert np.array_equal(kps[f, :n[f]].view(np.uint8), k1.view(np.uint8)), f
        assert np.array_equal(desc[f, :n[f]], d1), f
        if stereo and f % 2 == 0:
            kr, dr = two(host[f + 1])
            u, z, cnt = m.stereo_match(one, two, k1, d1, kr, dr, st[0], st[1])
            assert cnt == nm[f // 2]
            assert np.array_equal(ur[f // 2, :len(k1)].view(np.uint32), u.view(np.uint32))
            assert np.array_equal(dp[f // 2, :len(k1)].view(np.uint32), z.view(np.uint32))
    g.close()


def test_small_chunk_route_equals_batch_route(pkg):
    """Up to 4 frames per call take the fork / join route (per-level streams, CUDA graph once the arguments repeat,
    results written straight into page-locked memory); larger batches take the single-stream route with DMA copies.
    Same frames through both, with the graph replayed several times, pinned and pageable destinations, a change of shape
    in between and a capacity that is too small."""
    w, h, nf = 640, 480, 1000
    frames = np.stack([pkg.synth.frame(w, h, 900 + i) for i in range(8)])
    cap = nf + 64
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=800, max_height=600, max_batch=8)
    bk, bd, bn = np.empty((8, cap), pkg.KP_DTYPE), np.empty((8, cap, 32), np.uint8), np.zeros(8, np.int32)
    ex.extract_batch_into(frames, bk, bd, bn)                      # 8 frames: batch route
    pk, pd = pkg.pinned_empty((1, cap), pkg.KP_DTYPE), pkg.pinned_empty((1, cap, 32), np.uint8)
    pn = np.zeros(1, np.int32)
    pimg = pkg.pinned_empty((1, h, w), np.uint8)
    for rep in range(3):                                           # same arrays every call: capture, then replays
        for f in range(8):
            pimg[0] = frames[f]
            pk[:] = 0
            ex.extract_batch_into(pimg, pk, pd, pn)
            assert pn[0] == bn[f] > 900
            assert np.array_equal(pk[0, :pn[0]].view(np.uint8), bk[f, :bn[f]].view(np.uint8)), (rep, f)
            assert np.array_equal(pd[0, :pn[0]], bd[f, :bn[f]]), (rep, f)
    # pageable destinations (lane staging) and pairs of frames
    for f in range(0, 8, 2):
        k2, d2, n2 = np.empty((2, cap), pkg.KP_DTYPE), np.empty((2, cap, 32), np.uint8), np.zeros(2, np.int32)
        ex.extract_batch_into(frames[f:f + 2], k2, d2, n2)
        for j in range(2):
            assert n2[j] == bn[f + j] and np.array_equal(k2[j, :n2[j]].view(np.uint8), bk[f + j, :n2[j]].view(np.uint8))
            assert np.array_equal(d2[j, :n2[j]], bd[f + j, :n2[j]])
    # another shape in between invalidates the captured set; coming back re-captures
    other = pkg.synth.frame(800, 600, 5)
    ko, do = ex(other)
    one = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=800, max_height=600)
    ko2, do2 = one(other)
    assert np.array_equal(ko.view(np.uint8), ko2.view(np.uint8)) and np.array_equal(do, do2)
    for _ in range(3):
        pimg[0] = frames[3]
        ex.extract_batch_into(pimg, pk, pd, pn)
        assert pn[0] == bn[3] and np.array_equal(pd[0, :pn[0]], bd[3, :bn[3]])
    # capacity too small: ORBX_ERR_CAPACITY with the needed count in n, nothing beyond the arrays written
    sk, sd = pkg.pinned_empty((1, 100), pkg.KP_DTYPE), pkg.pinned_empty((1, 100, 32), np.uint8)
    with pytest.raises(pkg.OrbxError) as e:
        ex.extract_batch_into(pimg, sk, sd, pn)
    assert e.value.code == -3 and pn[0] == bn[3]
    assert np.array_equal(sk[0].view(np.uint8), bk[3, :100].view(np.uint8))


def test_small_chunk_route_two_threads(pkg):
    """Two extractor handles driven from two host threads at once (the stereo Frame constructor), both on the graph
    route: results equal the batch route."""
    import threading
    w, h, nf = 752, 480, 1200
    left, right = pkg.synth.stereo_pair(w, h, 77)
    cap = nf + 64
    ref = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=8)
    both = np.stack([left, right] * 4)
    bk, bd, bn = np.empty((8, cap), pkg.KP_DTYPE), np.empty((8, cap, 32), np.uint8), np.zeros(8, np.int32)
    ref.extract_batch_into(both, bk, bd, bn)
    out = {}

    def work(name, img):
        ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
        k, d, n = pkg.pinned_empty((1, cap), pkg.KP_DTYPE), pkg.pinned_empty((1, cap, 32), np.uint8), np.zeros(1, np.int32)
        ok = True
        for _ in range(20):
            k[:] = 0
            ex.extract_batch_into(img[None], k, d, n)
            j = 0 if name == "L" else 1
            ok = ok and n[0] == bn[j] and np.array_equal(k[0, :n[0]].view(np.uint8), bk[j, :bn[j]].view(np.uint8)) and \
                np.array_equal(d[0, :n[0]], bd[j, :bn[j]])
        out[name] = ok
    ts = [threading.Thread(target=work, args=("L", left)), threading.Thread(target=work, args=("R", right))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert out == {"L": True, "R": True}
