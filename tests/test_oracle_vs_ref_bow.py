"""Pins the oracle's restatements of SURVEY.md 8f rows 1-2 against the REFERENCE'S OWN CODE: ORBmatcher::SearchByBoW
(R/src/ORBmatcher.cpp:220-365 and :632-772), ORBmatcher::SearchForTriangulation (:785-983) and
MapPoint::ComputeDistinctiveDescriptors (R/src/MapPoint.cpp:306-385), run on real KeyFrame / Frame / MapPoint objects
(R/src/KeyFrame.cpp compiled where it lies into oracle/_ref/libref_front.so).

The vocabulary itself (DBoW2 transform, ORBvoc.txt) is outside the path and not shipped: feature vectors are filled
through DBoW2::FeatureVector::addFeature with a synthetic node id per feature (bow_case.node_of)."""
import numpy as np
import pytest

import bow_case
import oracle_lib as ol
import ref_front_lib as rf

pytestmark = pytest.mark.skipif(not rf.available(), reason="oracle/_ref/libref_front.so not built")


@pytest.fixture(scope="module")
def mono():
    return bow_case.build(rf, stereo=False)


@pytest.fixture(scope="module")
def stereo():
    return bow_case.build(rf, stereo=True)


@pytest.mark.parametrize("check_ori", [True, False])
@pytest.mark.parametrize("ratio", [0.7, 0.9])
def test_search_by_bow_keyframe_frame_vs_reference(mono, check_ori, ratio):
    c = mono
    rn, rmp = c.front.search_by_bow_kf_frame(c.kf1, c.f2, ratio, check_ori)
    en, em = ol.search_by_bow_kf_frame(c.f1.desc, c.f1.keys_un["angle"], c.mp_of1 >= 0, c.fv1, c.f2.desc,
                                       c.f2.keys_un["angle"], c.fv2, np.float32(ratio), check_ori)
    assert rn > 40
    assert rn == en
    assert np.array_equal(rmp, np.where(em >= 0, c.mp_of1[np.maximum(em, 0)], -1))


@pytest.mark.parametrize("check_ori", [True, False])
@pytest.mark.parametrize("ratio", [0.75, 0.9])
def test_search_by_bow_keyframes_vs_reference(mono, check_ori, ratio):
    c = mono
    rn, rmp = c.front.search_by_bow_kf_kf(c.kf1, c.kf2, ratio, check_ori)
    en, em = ol.search_by_bow_kf_kf(c.f1.desc, c.f1.keys_un["angle"], c.mp_of1 >= 0, c.fv1, c.f2.desc,
                                    c.f2.keys_un["angle"], c.mp_of2 >= 0, c.fv2, np.float32(ratio), check_ori)
    assert rn > 20
    assert rn == en
    assert np.array_equal(rmp, np.where(em >= 0, c.mp_of2[np.maximum(em, 0)], -1))


@pytest.mark.parametrize("case", ["mono", "stereo"])
@pytest.mark.parametrize("only_stereo", [False, True])
@pytest.mark.parametrize("check_ori", [False, True])
@pytest.mark.parametrize("epipole_inside", [False, True])
def test_search_for_triangulation_vs_reference(request, case, only_stereo, check_ori, epipole_inside):
    c = request.getfixturevalue(case)
    kf2 = c.kf2_center if epipole_inside else c.kf2
    ex, ey = c.front.epipole(c.kf1, kf2)
    rn, rpairs = c.front.search_for_triangulation(c.kf1, kf2, c.F12, only_stereo, 0.6, check_ori)
    sf, ls = kf2.tables()
    en, em = ol.search_for_triangulation(c.f1.keys_un, c.f1.desc, c.mp_of1 >= 0, c.f1.u_right, c.fv1, c.f2.keys_un, c.f2.desc,
                                         c.mp_of2 >= 0, c.f2.u_right, c.fv2, c.F12, ex, ey, sf, ls, only_stereo, check_ori)
    i1 = np.nonzero(em >= 0)[0]
    assert rn == en == len(i1)
    assert np.array_equal(rpairs, np.stack([i1, em[i1]], 1))
    if case == "mono":
        assert rn == 0 if only_stereo else rn > 20
        if epipole_inside and not only_stereo:   # the epipole test (:881-883) must have removed candidates
            n_far, _ = ol.search_for_triangulation(c.f1.keys_un, c.f1.desc, c.mp_of1 >= 0, c.f1.u_right, c.fv1, c.f2.keys_un,
                                                   c.f2.desc, c.mp_of2 >= 0, c.f2.u_right, c.fv2, c.F12, 1e6, 1e6, sf, ls,
                                                   only_stereo, check_ori)
            assert n_far > rn
    else:
        assert rn > 5


def test_compute_distinctive_descriptors_vs_reference():
    c = bow_case.build_observations(rf)
    rows, off = [], [0]
    for mp in c.points:
        r = c.front.observed_rows(mp)
        rows.append(r)
        off.append(off[-1] + len(r))
    desc = np.concatenate(rows)
    best, med = ol.distinctive(desc, np.array(off, np.int32))
    sizes = np.diff(off)
    assert sizes.min() == 1 and sizes.max() >= 6
    for k, mp in enumerate(c.points):
        chosen = c.front.distinctive(mp)   # runs the reference's function on the real MapPoint
        assert np.array_equal(chosen, rows[k][best[k]]), (k, sizes[k])
    # the median is the sorted row's element 0.5*(N-1), self distance included
    for k in range(len(c.points)):
        r = rows[k].astype(np.uint8)
        d = np.unpackbits(r[:, None, :] ^ r[None, :, :], axis=2).sum(2)
        m = np.sort(d, axis=1)[:, int(0.5 * (len(r) - 1))]
        assert med[k] == m.min() and best[k] == int(np.argmin(m))


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [3.0, 8.0])
def test_fuse_vs_reference(stereo, th):
    """Fuse(pKF, vpMapPoints, th): the oracle's windowed best-match core on the reference's own projection prologue, then
    the reference's bookkeeping rule on index arrays, must leave the key frame with the same map points as the reference's
    Fuse run on the real objects."""
    c = bow_case.build_fuse(rf, stereo)
    valid, uv, ur, radius, level, desc = c.front.fuse_queries(c.kf2, c.src, th)
    bi, bd = ol.search_best_in_window(c.f2.keys_un, c.f2.desc, c.f2.u_right, c.f2.bounds, c.kf2.inv_level_sigma2(), uv[valid],
                                      ur[valid], radius[valid], level[valid], desc[valid], None, 50)
    best_idx, best_dist = np.full(len(c.src), -1, np.int32), np.full(len(c.src), 256, np.int32)
    best_idx[valid], best_dist[valid] = bi, bd
    n_points = int(max(c.src.max(), c.mp_of2.max())) + 1
    obs = np.array([c.front.observations(k) for k in range(n_points)])
    state = c.mp_of2.copy()
    en = bow_case.simulate_fuse(state, obs, c.src, valid, best_idx, best_dist, c.f2.u_right >= 0)
    rn = c.front.fuse(c.kf2, c.src, th)
    assert valid.sum() > 500 and rn > (20 if stereo else 300)
    assert rn == en
    assert np.array_equal(c.kf2.map_points(), state)
    replaced = (state != c.mp_of2) & (c.mp_of2 >= 0)
    assert replaced.any() and ((state == c.mp_of2) & (c.mp_of2 >= 0)).any() and ((c.mp_of2 < 0) & (state >= 0)).any()


@pytest.mark.parametrize("stereo", [False, True])
@pytest.mark.parametrize("th", [4, 10])
def test_search_by_projection_keyframe_scw_vs_reference(stereo, th):
    c = bow_case.build_fuse(rf, stereo)
    Scw = bow_case.pose(bow_case.T_FUSE) * np.float32(1.0)
    Scw[:3, :] *= np.float32(1.25)    # a similarity: the function divides the scale out again (:378-382)
    matched = np.full(c.f2.N, -1, np.int32)
    matched[5::9] = c.src[np.arange(len(matched[5::9])) % len(c.src)]   # keypoints loop closing has matched already
    valid, uv, radius, level, desc = c.front.scw_queries(c.kf2, Scw, c.src, matched, th)
    taken = (matched >= 0).astype(np.uint8)
    bi, bd = ol.search_best_in_window(c.f2.keys_un, c.f2.desc, None, c.f2.bounds, None, uv[valid], None, radius[valid],
                                      level[valid], desc[valid], taken, 50)
    expect = matched.copy()
    hit = bd <= 50
    expect[bi[hit]] = c.src[valid][hit]
    rn, rmatched = c.front.search_by_projection_kf_scw(c.kf2, Scw, c.src, matched, th)
    assert rn == int(hit.sum()) and rn > 200
    assert np.array_equal(rmatched, expect)
    assert (~valid).sum() >= (matched >= 0).sum() - 5    # the already-found points were skipped
