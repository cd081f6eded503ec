"""Scenario builders for the bag-of-words restricted searches and the distinctive-descriptor tests (TEST
INFRASTRUCTURE). `build` / `build_observations` create real reference Frame / KeyFrame / MapPoint objects through
tests/ref_front_lib.py; `build_arrays` produces the same plain arrays from the product extractor for boxes without
oracle/_ref."""
import importlib
from types import SimpleNamespace

import numpy as np

synth = importlib.import_module("orb-slam2-_b200.synth")

W, H, NFEAT = 640, 480, 2000
FX = FY = 500.0
CX, CY = 320.0, 240.0
BF = 40.0
SHIFT = (3, 6)   # rows, columns: image 2 = image 1 moved by +6 px in x and +3 px in y
# x1^T F12 x2 = 0 for x2 - x1 parallel to d = (6, 3, 0): F12 = [d]_x
F12 = np.array([[0, 0, 3], [0, 0, -6], [-3, 6, 0]], np.float32)
T_FAR = np.array([0.1, 0.05, 0.002], np.float32)     # epipole far outside image 2
T_CENTER = np.array([0.0, 0.0, 0.1], np.float32)     # epipole at the principal point


def node_of(desc):
    """Synthetic vocabulary node per descriptor (stands in for DBoW2's transform at levelsup 4): 7 descriptor bits, so
    that corresponding features of two views mostly share a node and every node holds ~16 of 2000 features."""
    d = np.asarray(desc, np.uint8)
    return ((d[:, 0] >> 4).astype(np.int32) << 3) | (d[:, 5] >> 5).astype(np.int32)


def images(stereo, seed=70):
    if stereo:
        left, right = synth.stereo_pair(W, H, seed)
        return (left, right), (np.roll(left, SHIFT, (0, 1)), np.roll(right, SHIFT, (0, 1)))
    img = synth.frame(W, H, seed)
    return (img,), (np.roll(img, SHIFT, (0, 1)),)


def pose(t):
    T = np.eye(4, dtype=np.float32)
    T[:3, 3] = t
    return T


def epipole_translation(kp, z=0.1):
    """Translation of camera 2 that puts the epipole of camera 1 (at the origin) on keypoint kp of image 2."""
    return np.array([(kp["x"] - CX) / FX * z, (kp["y"] - CY) / FY * z, z], np.float32)


def point_masks(n1, n2):
    """Which keypoints own a map point: two of three in key frame 1, three of four in key frame 2."""
    return (np.arange(n1) % 3) != 0, (np.arange(n2) % 4) != 0


def backproject(keys, t, depth=4.0):
    x = (keys["x"] - np.float32(CX)) / np.float32(FX) * depth
    y = (keys["y"] - np.float32(CY)) / np.float32(FY) * depth
    return np.stack([x, y, np.full(len(keys), depth, np.float32)], 1).astype(np.float32) - t


def build(rf, stereo):
    front = rf.Front(NFEAT, fx=FX, fy=FY, cx=CX, cy=CY, bf=BF)
    im1, im2 = images(stereo)
    mk = front.frame_stereo if stereo else front.frame_mono
    f1, f2, f2c = mk(*im1), mk(*im2), mk(*im2)
    f1.set_pose(pose(np.zeros(3, np.float32)))
    f2.set_pose(pose(T_FAR))
    n1, n2 = node_of(f1.desc), node_of(f2.desc)
    n2[::17] = -1            # some features carry no node; the two vectors then also differ in their node sets
    n1[n1 == 5] = -1
    n2[n2 == 9] = -1
    f1.set_featvec(n1)
    f2.set_featvec(n2)
    f2c.set_featvec(n2)
    kf1, kf2 = front.keyframe(f1), front.keyframe(f2)
    m1, m2 = point_masks(f1.N, f2.N)
    mp_of1, mp_of2 = np.full(f1.N, -1, np.int32), np.full(f2.N, -1, np.int32)
    p1, p2 = backproject(f1.keys_un, np.zeros(3, np.float32)), backproject(f2.keys_un, T_FAR)
    for i in np.nonzero(m1)[0]:
        mp_of1[i] = front.map_point(f1, int(i), p1[i])
        kf1.add_map_point(int(i), int(mp_of1[i]))
    for i in np.nonzero(m2)[0]:
        mp_of2[i] = front.map_point(f2, int(i), p2[i])
        kf2.add_map_point(int(i), int(mp_of2[i]))
    # a third pose of camera 2 whose epipole falls on a keypoint that the far-epipole search matches, so that the
    # epipole-distance rejection (:881-883) has something to reject
    _, pairs = front.search_for_triangulation(kf1, kf2, F12, False, 0.6, False)
    t_near = epipole_translation(f2.keys_un[pairs[len(pairs) // 2][1]]) if len(pairs) else T_CENTER
    f2c.set_pose(pose(t_near))
    kf2c = front.keyframe(f2c)
    for i in np.nonzero(m2)[0]:
        kf2c.add_map_point(int(i), int(mp_of2[i]))
    import oracle_lib as ol
    return SimpleNamespace(front=front, f1=f1, f2=f2, f2c=f2c, kf1=kf1, kf2=kf2, kf2_center=kf2c, mp_of1=mp_of1, mp_of2=mp_of2,
                           fv1=ol.featvec(n1), fv2=ol.featvec(n2), F12=F12, node1=n1, node2=n2, t_near=t_near)


def build_arrays(pkg, stereo):
    """The same inputs as plain arrays from the product extractor (no reference objects)."""
    import oracle_lib as ol
    im1, im2 = images(stereo)
    ex = pkg.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H)
    frames = []
    for ims in (im1, im2):
        k, d = ex(ims[0])
        ur = np.full(len(k), -1, np.float32)
        if stereo:
            exr = pkg.ORBextractor(NFEAT, 1.2, 8, 20, 7, max_width=W, max_height=H)
            k, d = ex(ims[0])
            kr, dr = exr(ims[1])
            ur, _, _ = pkg.ORBmatcher(0.6).stereo_match(ex, exr, k, d, kr, dr, BF, 0.0)
            exr.close()
        frames.append(SimpleNamespace(keys_un=k, desc=d, u_right=np.ascontiguousarray(ur, np.float32), N=len(k)))
    f1, f2 = frames
    n1, n2 = node_of(f1.desc), node_of(f2.desc)
    n2[::17] = -1
    n1[n1 == 5] = -1
    n2[n2 == 9] = -1
    m1, m2 = point_masks(f1.N, f2.N)
    mp_of1 = np.where(m1, np.arange(f1.N), -1).astype(np.int32)
    mp_of2 = np.where(m2, f1.N + np.arange(f2.N), -1).astype(np.int32)
    sf = np.asarray(ex.GetScaleFactors(), np.float32)
    ls = np.asarray(ex.GetScaleSigmaSquares(), np.float32)
    return SimpleNamespace(front=None, f1=f1, f2=f2, mp_of1=mp_of1, mp_of2=mp_of2, fv1=ol.featvec(n1), fv2=ol.featvec(n2),
                           F12=F12, node1=n1, node2=n2, scale_factors=sf, level_sigma2=ls)


def epipole_of(t):
    """(ex, ey) of SearchForTriangulation's prologue for key frame 1 at the origin and key frame 2 at [I | t], in
    float32 with the reference's operation order (fma(fx*C2x, 1/C2z, cx)); only used without oracle/_ref."""
    t = np.asarray(t, np.float32)
    invz = np.float32(1.0) / t[2]
    fma = lambda a, b, c: np.float32(np.float64(a) * np.float64(b) + np.float64(c))   # exact product, one rounding
    return fma(np.float32(FX) * t[0], invz, np.float32(CX)), fma(np.float32(FY) * t[1], invz, np.float32(CY))


def build_observations(rf, n_kf=7, n_points=60):
    """Map points observed by 1..n_kf key frames each (the input of MapPoint::ComputeDistinctiveDescriptors)."""
    front = rf.Front(500, fx=FX, fy=FY, cx=CX, cy=CY, bf=BF)
    img = synth.frame(W, H, 91)
    frames, kfs = [], []
    for j in range(n_kf):
        f = front.frame_mono(np.roll(img, (j, 2 * j), (0, 1)))
        f.set_pose(pose(np.array([0.01 * j, 0, 0], np.float32)))
        frames.append(f)
        kfs.append(front.keyframe(f))
    points = []
    for p in range(n_points):
        n_obs = 1 + p % n_kf
        j0 = p % n_kf
        idx0 = (p * 11 + j0 * 37) % frames[j0].N
        mp = front.map_point(frames[j0], idx0, backproject(frames[j0].keys_un[idx0:idx0 + 1], np.zeros(3, np.float32))[0], n_obs=0)
        for t in range(n_obs):
            j = (p + t) % n_kf
            kfs[j].add_map_point((p * 11 + j * 37) % frames[j].N, mp)
        points.append(mp)
    return SimpleNamespace(front=front, frames=frames, kfs=kfs, points=points)


T_FUSE = np.array([6 * 4.0 / FX, 3 * 4.0 / FY, 0.0], np.float32)   # points at depth 4 move by the image shift (6, 3)


def build_fuse(rf, stereo):
    """Key frame 1 with map points at depth 4 and a target key frame (the shifted view, pose T_FUSE) that owns map points
    on three of four keypoints, observed only there: the input of Fuse(pKF, vpMapPoints, th) and of
    SearchByProjection(pKF, Scw, vpPoints, vpMatched, th)."""
    front = rf.Front(NFEAT, fx=FX, fy=FY, cx=CX, cy=CY, bf=BF)
    im1, im2 = images(stereo)
    mk = front.frame_stereo if stereo else front.frame_mono
    f1, f2 = mk(*im1), mk(*im2)
    f1.set_pose(pose(np.zeros(3, np.float32)))
    f2.set_pose(pose(T_FUSE))
    kf1, kf2 = front.keyframe(f1), front.keyframe(f2)
    p1, p2 = backproject(f1.keys_un, np.zeros(3, np.float32)), backproject(f2.keys_un, T_FUSE)
    src = np.array([front.map_point(f1, int(i), p1[i]) for i in range(0, f1.N, 2)], np.int32)
    for k, i in enumerate(range(0, f1.N, 2)):
        kf1.add_map_point(int(i), int(src[k]))
    mp_of2 = np.full(f2.N, -1, np.int32)
    for i in range(f2.N):
        if i % 4:
            mp_of2[i] = front.map_point(f2, int(i), p2[i], n_obs=(3 if i % 8 < 4 else 0))
            kf2.add_map_point(int(i), int(mp_of2[i]))
    return SimpleNamespace(front=front, f1=f1, f2=f2, kf1=kf1, kf2=kf2, src=src, mp_of2=mp_of2)


def simulate_fuse(state, obs, src, valid, best_idx, best_dist, stereo_slot, th_low=50):
    """What Fuse does with the search results (R/src/ORBmatcher.cpp:1112-1136), on index arrays: state[idx] = map point
    owning keypoint idx of the target key frame, obs[mp] = Observations(). Returns nFused."""
    n = 0
    for i in np.nonzero(valid)[0]:
        if best_dist[i] > th_low:
            continue
        b, mp = int(best_idx[i]), int(src[i])
        if state[b] >= 0:
            if obs[state[b]] > obs[mp]:
                pass                      # pMP->Replace(pMPinKF): the key frame keeps its point
            else:                         # pMPinKF->Replace(pMP): its single observation (this slot) moves to pMP
                obs[mp] += 2 if stereo_slot[b] else 1
                state[b] = mp
        else:
            obs[mp] += 2 if stereo_slot[b] else 1
            state[b] = mp
        n += 1
    return n
