"""GPU parity of the windowed search (Frame grid + GetFeaturesInArea + Hamming distances) and of
Frame::ComputeStereoMatches against the CPU oracle, through the C ABI."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _frame_features(pkg, w, h, nf, seed, shift=(0, 0)):
    img = pkg.synth.frame(w, h, seed)
    if shift != (0, 0):
        img = np.roll(img, shift, (0, 1))
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    k, d = ex(img)
    return ex, img, k, d


@pytest.mark.parametrize("w,h,nf", [(640, 480, 1000), (1241, 376, 2000)])
def test_window_candidates_match_get_features_in_area(pkg, oracle, w, h, nf):
    _, _, k, d = _frame_features(pkg, w, h, nf, 50)
    bounds = (0.0, float(w), 0.0, float(h))
    rng = np.random.default_rng(3)
    nq = 700
    q_xy = np.stack([rng.uniform(-40, w + 40, nq), rng.uniform(-40, h + 40, nq)], 1).astype(np.float32)
    q_xy[:50] = np.stack([k["x"][:50], k["y"][:50]], 1)           # exactly on keypoints
    q_xy[50:60] = [[0, 0]] * 5 + [[w, h]] * 5                      # corners of the image bounds
    q_r = rng.choice([2.5, 7.0, 15.0, 40.0, 100.0], nq).astype(np.float32)
    lo = rng.integers(-1, 8, nq).astype(np.int32)
    hi = np.where(rng.random(nq) < 0.3, -1, lo + rng.integers(0, 3, nq)).astype(np.int32)
    q_desc = pkg.synth.descriptors(nq, 9)
    q_desc[:200] = d[rng.integers(0, len(d), 200)]
    m = pkg.ORBmatcher()
    off, idx, dist = m.match_window(k, d, bounds, q_xy, q_r, lo, hi, q_desc)
    g = oracle.Grid(k, bounds)
    assert off[0] == 0 and off[-1] == len(idx)
    total = 0
    for i in range(nq):
        exp = g.query(float(q_xy[i, 0]), float(q_xy[i, 1]), float(q_r[i]), int(lo[i]), int(hi[i]))
        got = idx[off[i]:off[i + 1]]
        assert np.array_equal(got, exp), i
        for j, c in enumerate(exp[:8]):
            assert dist[off[i] + j] == oracle.descriptor_distance(q_desc[i], d[c])
        total += len(exp)
    assert total > 1000
    # empty frame and empty query set
    off, idx, dist = m.match_window(k[:0], d[:0], bounds, q_xy[:5], q_r[:5], lo[:5], hi[:5], q_desc[:5])
    assert off.tolist() == [0] * 6 and len(idx) == 0


@pytest.mark.parametrize("w,h,nf,seed", [(752, 480, 1200, 5), (1241, 376, 2000, 6)])
def test_stereo_matches_bit_exact(pkg, oracle, w, h, nf, seed):
    left, right = pkg.synth.stereo_pair(w, h, seed)
    exl = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    kl, dl = exl(left)
    kr, dr = exr(right)
    mb = np.float32(0.11)
    mbf = np.float32(435.2) * mb
    m = pkg.ORBmatcher()
    u, z, n = m.stereo_match(exl, exr, kl, dl, kr, dr, mbf, mb)
    ox = oracle.Extractor(nf, 1.2, 8, 20, 7)
    ox(left)
    pl = [ox.level(l) for l in range(8)]
    ox(right)
    pr = [ox.level(l) for l in range(8)]
    t = ox.tables()
    eu, ez, en = oracle.stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], mbf, mb)
    assert en > 50
    assert n == en
    assert np.array_equal(u.view(np.uint32), eu.view(np.uint32))
    assert np.array_equal(z.view(np.uint32), ez.view(np.uint32))
    # no right keypoints at all
    u, z, n = m.stereo_match(exl, exr, kl, dl, kr[:0], dr[:0], mbf, mb)
    assert n == 0 and (u == -1).all() and (z == -1).all()


@pytest.mark.parametrize("interleaved", [True, False])
def test_stereo_batch_device_equals_per_pair(pkg, oracle, interleaved):
    """orbx_stereo_match_batch_device on one device-resident batch of stereo pairs (L R L R ... or all lefts then all
    rights) gives every pair the result of the per-pair host call; pair 0 is also checked against the oracle."""
    import torch
    w, h, nf, pairs = 752, 480, 1200, 5
    imgs = [pkg.synth.stereo_pair(w, h, 20 + p) for p in range(pairs)]
    if interleaved:
        frames = np.stack([im for lr in imgs for im in lr])
        l0, r0, step = 0, 1, 2
    else:
        frames = np.stack([lr[0] for lr in imgs] + [lr[1] for lr in imgs])
        l0, r0, step = 0, pairs, 1
    B = 2 * pairs
    cap = nf + 64
    ex = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h, max_batch=B)
    d_frames = torch.from_numpy(frames).cuda()
    dk = torch.zeros((B, cap, 7), dtype=torch.float32, device="cuda")
    dd = torch.zeros((B, cap, 32), dtype=torch.uint8, device="cuda")
    dn = torch.zeros(B, dtype=torch.int32, device="cuda")
    du = torch.full((pairs, cap), -7.0, dtype=torch.float32, device="cuda")
    dz = torch.full((pairs, cap), -7.0, dtype=torch.float32, device="cuda")
    dm = torch.zeros(pairs, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    ex.extract_batch_device(d_frames.data_ptr(), B, w, h, w, w * h, dk.data_ptr(), dd.data_ptr(), cap, dn.data_ptr())
    m = pkg.ORBmatcher()
    mb = np.float32(0.0)                      # what the reference's stereo constructor runs with
    mbf = np.float32(435.2 * 0.11)
    m.stereo_match_batch_device(ex, pairs, l0, r0, step, dk.data_ptr(), dd.data_ptr(), dn.data_ptr(), cap, mbf, mb,
                                du.data_ptr(), dz.data_ptr(), dm.data_ptr())
    torch.cuda.synchronize()
    n = dn.cpu().numpy()
    kps = dk.cpu().numpy().view(np.uint8).reshape(B, cap, 28).view(pkg.KP_DTYPE).reshape(B, cap)
    desc = dd.cpu().numpy()
    u_all, z_all, matched = du.cpu().numpy(), dz.cpu().numpy(), dm.cpu().numpy()
    exl = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    exr = pkg.ORBextractor(nf, 1.2, 8, 20, 7, max_width=w, max_height=h)
    for p in range(pairs):
        fl, fr = l0 + p * step, r0 + p * step
        kl, dl = exl(frames[fl])
        kr, dr = exr(frames[fr])
        assert n[fl] == len(kl) and np.array_equal(kps[fl, :n[fl]].view(np.uint8), kl.view(np.uint8))
        u, z, cnt = m.stereo_match(exl, exr, kl, dl, kr, dr, mbf, mb)
        assert cnt == matched[p] > 50
        assert np.array_equal(u_all[p, :n[fl]].view(np.uint32), u.view(np.uint32))
        assert np.array_equal(z_all[p, :n[fl]].view(np.uint32), z.view(np.uint32))
        assert (u_all[p, n[fl]:] == -7.0).all()
        if p == 0:
            ox = oracle.Extractor(nf, 1.2, 8, 20, 7)
            ox(frames[fl])
            pl = [ox.level(l) for l in range(8)]
            ox(frames[fr])
            pr = [ox.level(l) for l in range(8)]
            t = ox.tables()
            eu, ez, en = oracle.stereo_matches(kl, dl, kr, dr, pl, pr, t["scale"], t["inv_scale"], mbf, mb)
            assert en == cnt and np.array_equal(u.view(np.uint32), eu.view(np.uint32)) and np.array_equal(z.view(np.uint32), ez.view(np.uint32))
