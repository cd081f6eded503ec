// TEST INFRASTRUCTURE (oracle) -- CPU restatement of the ORBmatcher / Frame hot-path pieces.
// Follows R/src/ORBmatcher.cpp (DescriptorDistance :1901-1917, top-2 + ratio loops :266-300,
// SearchForInitialization :499-617, SearchByProjection :63-163 / :1564-1718, ComputeThreeMaxima
// :1854-1895) and R/src/Frame.cpp (grid :244-260, :387-452; ComputeStereoMatches :551-770).
// PINNED: tests/test_oracle_vs_ref_front.py runs the reference's own Frame.cpp / ORBmatcher.cpp / MapPoint.cpp
// (oracle/_ref/libref_front.so, compiled where they lie against oracle/ref_shim) on real Frame and MapPoint
// objects and requires bit-identical results from every search and from the stereo matches.
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "orb_oracle.h"

namespace orb_oracle {

// R/src/ORBmatcher.cpp:1901-1917 -- 8 x 32-bit SWAR population count of a XOR b.
static inline int descriptor_distance(const uint8_t* a, const uint8_t* b) {
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t x, y;
    std::memcpy(&x, a + 4 * i, 4);
    std::memcpy(&y, b + 4 * i, 4);
    uint32_t v = x ^ y;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}


// ---- Frame grid: R/include/Frame.h:37-38,181; R/src/Frame.cpp AssignFeaturesToGrid :244-260,
//      PosInGrid :442-452, GetFeaturesInArea :387-440 -------------------------------------------
constexpr int kGridCols = 64, kGridRows = 48;

struct Grid {
  std::vector<orb_oracle_keypoint> kps;  // mvKeysUn
  float minX, minY, invW, invH;          // mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv
  std::vector<int> cells[kGridCols][kGridRows];

  void build(const orb_oracle_keypoint* k, int n, float minX_, float maxX_, float minY_, float maxY_) {
    kps.assign(k, k + n);
    minX = minX_; minY = minY_;
    invW = static_cast<float>(kGridCols) / (maxX_ - minX_);   // Frame.cpp:168-169
    invH = static_cast<float>(kGridRows) / (maxY_ - minY_);
    for (int i = 0; i < n; i++) {
      const int px = (int)std::round((kps[i].x - minX) * invW);  // :444-445 (round half away from zero)
      const int py = (int)std::round((kps[i].y - minY) * invH);
      if (px < 0 || px >= kGridCols || py < 0 || py >= kGridRows) continue;
      cells[px][py].push_back(i);
    }
  }

  void query(float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) const {
    out.clear();
    const int nMinCellX = std::max(0, (int)std::floor((x - minX - r) * invW));
    if (nMinCellX >= kGridCols) return;
    const int nMaxCellX = std::min(kGridCols - 1, (int)std::ceil((x - minX + r) * invW));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - minY - r) * invH));
    if (nMinCellY >= kGridRows) return;
    const int nMaxCellY = std::min(kGridRows - 1, (int)std::ceil((y - minY + r) * invH));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++)
      for (int iy = nMinCellY; iy <= nMaxCellY; iy++)
        for (int idx : cells[ix][iy]) {
          const orb_oracle_keypoint& kp = kps[idx];
          if (bCheckLevels) {
            if (kp.octave < minLevel) continue;
            if (maxLevel >= 0 && kp.octave > maxLevel) continue;
          }
          const float distx = kp.x - x, disty = kp.y - y;
          if (std::fabs(distx) < r && std::fabs(disty) < r) out.push_back(idx);
        }
  }
};

// R/src/ORBmatcher.cpp:1854-1895
static void compute_three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
  int max1 = 0, max2 = 0, max3 = 0;
  for (int i = 0; i < L; i++) {
    const int s = (int)histo[i].size();
    if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
    else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
    else if (s > max3) { max3 = s; ind3 = i; }
  }
  if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
  else if (max3 < 0.1f * (float)max1) { ind3 = -1; }
}

constexpr int kHistoLength = 30, kThLow = 50, kThHigh = 100;  // R/src/ORBmatcher.cpp:37-39

static inline int rot_bin(float a1, float a2) {  // :569-575, :1684-1690
  const float factor = kHistoLength / 360.0f;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)std::round(rot * factor);
  if (bin == kHistoLength) bin = 0;
  return bin;
}

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {

int orb_oracle_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// Brute-force top-2 + ratio: the inner-loop pattern of R/src/ORBmatcher.cpp:266-300 / :526-556
// applied to every database row in index order.
void orb_oracle_match_bf2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, float ratio, int th,
                          int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  auto work = [&](int q0, int q1) {
    for (int i = q0; i < q1; i++) {
      const uint8_t* dq = q + (size_t)i * 32;
      int best = INT_MAX, best2 = INT_MAX, bidx = -1;
      for (int64_t j = 0; j < ndb; j++) {
        const int dist = descriptor_distance(dq, db + (size_t)j * 32);
        if (dist < best) { best2 = best; best = dist; bidx = (int)j; }
        else if (dist < best2) { best2 = dist; }
      }
      idx[i] = bidx; d1[i] = best; d2[i] = best2;
      if (accepted) accepted[i] = (best <= th && (float)best < ratio * (float)best2) ? 1 : 0;
    }
  };
  std::vector<std::thread> ts;
  for (int t = 0; t < nthreads; t++) ts.emplace_back(work, (int)((int64_t)nq * t / nthreads), (int)((int64_t)nq * (t + 1) / nthreads));
  for (auto& t : ts) t.join();
}

}  // extern "C"

// ---- windowed searches ----------------------------------------------------------------------------
struct orb_oracle_grid { Grid g; };

extern "C" {

orb_oracle_grid* orb_oracle_grid_create(const orb_oracle_keypoint* kps_un, int n, float min_x, float max_x,
                                        float min_y, float max_y) {
  orb_oracle_grid* h = new orb_oracle_grid();
  h->g.build(kps_un, n, min_x, max_x, min_y, max_y);
  return h;
}
void orb_oracle_grid_destroy(orb_oracle_grid* h) { delete h; }
int orb_oracle_grid_query(orb_oracle_grid* h, float x, float y, float r, int min_level, int max_level,
                          int32_t* out_idx, int cap) {
  std::vector<int> v;
  h->g.query(x, y, r, min_level, max_level, v);
  for (int i = 0; i < (int)v.size() && i < cap; i++) out_idx[i] = v[i];
  return (int)v.size();
}

// R/src/ORBmatcher.cpp:499-617
int orb_oracle_search_for_initialization(const orb_oracle_keypoint* kps1, const uint8_t* desc1, int n1,
                                         const orb_oracle_keypoint* kps2, const uint8_t* desc2, int n2,
                                         float min_x, float max_x, float min_y, float max_y,
                                         float* prev_matched_xy, int32_t* matches12, int window, float nnratio,
                                         int check_orientation) {
  Grid g2;
  g2.build(kps2, n2, min_x, max_x, min_y, max_y);
  int nmatches = 0;
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<int> rotHist[kHistoLength];
  std::vector<int> vMatchedDistance(n2, INT_MAX), vnMatches21(n2, -1), vIndices2;
  for (int i1 = 0; i1 < n1; i1++) {
    const int level1 = kps1[i1].octave;
    if (level1 > 0) continue;
    g2.query(prev_matched_xy[2 * i1], prev_matched_xy[2 * i1 + 1], (float)window, level1, level1, vIndices2);
    if (vIndices2.empty()) continue;
    const uint8_t* d1 = desc1 + (size_t)i1 * 32;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      const int dist = descriptor_distance(d1, desc2 + (size_t)i2 * 32);
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) { bestDist2 = dist; }
    }
    if (bestDist <= kThLow) {
      if (bestDist < (float)bestDist2 * nnratio) {
        if (vnMatches21[bestIdx2] >= 0) { matches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        matches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (check_orientation) rotHist[rot_bin(kps1[i1].angle, kps2[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, kHistoLength, ind1, ind2, ind3);
    for (int i = 0; i < kHistoLength; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int idx1 : rotHist[i])
        if (matches12[idx1] >= 0) { matches12[idx1] = -1; nmatches--; }
    }
  }
  for (int i1 = 0; i1 < n1; i1++)
    if (matches12[i1] >= 0) {
      prev_matched_xy[2 * i1] = kps2[matches12[i1]].x;
      prev_matched_xy[2 * i1 + 1] = kps2[matches12[i1]].y;
    }
  return nmatches;
}

// R/src/ORBmatcher.cpp:63-163. The map-point side is given as arrays: projection (u,v), search radius
// r = RadiusByViewingCos(viewCos)*th (:85-89), predicted level, projected right coordinate, descriptor.
// has_mp[idx] != 0 <=> F.mvpMapPoints[idx] && Observations()>0 (:118-121); mp_match[iMP] = keypoint or -1.
// mp_has_obs[iMP] != 0 <=> pMP->Observations()>0 (null: every point has observations): a keypoint taken by a point
// WITHOUT observations stays available to later points (:118-121 test the slot's point, :157 overwrites it).
int orb_oracle_search_by_projection_mp(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n,
                                       const float* uright, const uint8_t* has_mp, float min_x, float max_x,
                                       float min_y, float max_y, const float* mp_uv, const float* mp_r,
                                       const int32_t* mp_level, const float* mp_ur, const uint8_t* mp_desc, int nmp,
                                       const float* scale_factors, float nnratio, int32_t* mp_match,
                                       const uint8_t* mp_has_obs) {
  Grid g;
  g.build(kps_un, n, min_x, max_x, min_y, max_y);
  std::vector<uint8_t> owned(has_mp, has_mp + n);
  std::vector<int> vIndices;
  int nmatches = 0;
  for (int iMP = 0; iMP < nmp; iMP++) {
    mp_match[iMP] = -1;
    const int nPredictedLevel = mp_level[iMP];
    const float r = mp_r[iMP];
    g.query(mp_uv[2 * iMP], mp_uv[2 * iMP + 1], r * scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel,
            vIndices);
    if (vIndices.empty()) continue;
    const uint8_t* dMP = mp_desc + (size_t)iMP * 32;
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int idx : vIndices) {
      if (owned[idx]) continue;
      if (uright && uright[idx] > 0) {
        const float er = std::fabs(mp_ur[iMP] - uright[idx]);
        if (er > r * scale_factors[nPredictedLevel]) continue;
      }
      const int dist = descriptor_distance(dMP, desc + (size_t)idx * 32);
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = kps_un[idx].octave; bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = kps_un[idx].octave; bestDist2 = dist;
      }
    }
    if (bestDist <= kThHigh) {
      if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
      mp_match[iMP] = bestIdx;
      owned[bestIdx] = mp_has_obs ? (mp_has_obs[iMP] != 0) : 1;  // F.mvpMapPoints[bestIdx]=pMP (:157)
      nmatches++;
    }
  }
  return nmatches;
}

// R/src/ORBmatcher.cpp:1564-1718 after the projection prologue (:1600-1634): per last-frame map point the
// projection (u,v), radius th*scale[octave], level window, projected right coordinate ur = u - mbf*invzc
// (:1655), the last frame's keypoint angle and descriptor. cur_match[i2] = query index or -1.
int orb_oracle_search_by_projection_frame(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n,
                                          const float* uright, const uint8_t* has_mp, float min_x, float max_x,
                                          float min_y, float max_y, const float* q_uv, const float* q_r,
                                          const int32_t* q_min_level, const int32_t* q_max_level, const float* q_ur,
                                          const float* q_angle, const uint8_t* q_desc, int nq, int th_dist,
                                          int check_orientation, int32_t* cur_match, const uint8_t* q_has_obs) {
  Grid g;
  g.build(kps_un, n, min_x, max_x, min_y, max_y);
  std::vector<uint8_t> owned(has_mp, has_mp + n);
  for (int i = 0; i < n; i++) cur_match[i] = -1;
  std::vector<int> rotHist[kHistoLength], vIndices2;
  int nmatches = 0;
  for (int i = 0; i < nq; i++) {
    const float radius = q_r[i];
    g.query(q_uv[2 * i], q_uv[2 * i + 1], radius, q_min_level[i], q_max_level[i], vIndices2);
    if (vIndices2.empty()) continue;
    const uint8_t* dMP = q_desc + (size_t)i * 32;
    int bestDist = 256, bestIdx2 = -1;
    for (int i2 : vIndices2) {
      if (owned[i2]) continue;
      if (uright && uright[i2] > 0) {
        const float er = std::fabs(q_ur[i] - uright[i2]);
        if (er > radius) continue;
      }
      const int dist = descriptor_distance(dMP, desc + (size_t)i2 * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= th_dist) {
      cur_match[bestIdx2] = i;
      owned[bestIdx2] = q_has_obs ? (q_has_obs[i] != 0) : 1;  // the slot now holds pMP (:1675)
      nmatches++;
      if (check_orientation) rotHist[rot_bin(q_angle[i], kps_un[bestIdx2].angle)].push_back(bestIdx2);
    }
  }
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, kHistoLength, ind1, ind2, ind3);
    for (int i = 0; i < kHistoLength; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (int idx : rotHist[i]) { cur_match[idx] = -1; nmatches--; }
  }
  return nmatches;
}

}  // extern "C"

// ---- Frame::ComputeStereoMatches (R/src/Frame.cpp:551-770) ----------------------------------------
extern "C" int orb_oracle_stereo_matches(const orb_oracle_keypoint* kps_l, const uint8_t* desc_l, int nl,
                                         const orb_oracle_keypoint* kps_r, const uint8_t* desc_r, int nr,
                                         const uint8_t* const* pyr_l, const uint8_t* const* pyr_r,
                                         const int32_t* lvl_w, const int32_t* lvl_h, const int32_t* lvl_stride,
                                         int nlevels, const float* scale_factors, const float* inv_scale_factors,
                                         float mbf, float mb, float* u_right, float* depth) {
  (void)nlevels;
  for (int i = 0; i < nl; i++) { u_right[i] = -1.0f; depth[i] = -1.0f; }
  const int thOrbDist = (kThHigh + kThLow) / 2;
  const int nRows = lvl_h[0];
  std::vector<std::vector<int>> vRowIndices(nRows);
  for (int iR = 0; iR < nr; iR++) {
    const float kpY = kps_r[iR].y;
    const float r = 2.0f * scale_factors[kps_r[iR].octave];
    const int maxr = (int)std::ceil(kpY + r);
    const int minr = (int)std::floor(kpY - r);
    for (int yi = minr; yi <= maxr; yi++)
      if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // (the reference indexes unchecked)
  }
  const float minZ = mb, minD = 0, maxD = mbf / minZ;
  std::vector<std::pair<int, int>> vDistIdx;
  for (int iL = 0; iL < nl; iL++) {
    const orb_oracle_keypoint& kpL = kps_l[iL];
    const int levelL = kpL.octave;
    const float vL = kpL.y, uL = kpL.x;
    const std::vector<int>& vCandidates = vRowIndices[(size_t)vL];
    if (vCandidates.empty()) continue;
    const float minU = uL - maxD, maxU = uL - minD;
    if (maxU < 0) continue;
    int bestDist = kThHigh;
    int bestIdxR = 0;
    const uint8_t* dL = desc_l + (size_t)iL * 32;
    for (int iR : vCandidates) {
      const orb_oracle_keypoint& kpR = kps_r[iR];
      if (kpR.octave < levelL - 1 || kpR.octave > levelL + 1) continue;
      const float uR = kpR.x;
      if (uR >= minU && uR <= maxU) {
        const int dist = descriptor_distance(dL, desc_r + (size_t)iR * 32);
        if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
      }
    }
    if (bestDist < thOrbDist) {
      const float uR0 = kps_r[bestIdxR].x;
      const float scaleFactor = inv_scale_factors[kpL.octave];
      const float scaleduL = std::round(kpL.x * scaleFactor);
      const float scaledvL = std::round(kpL.y * scaleFactor);
      const float scaleduR0 = std::round(uR0 * scaleFactor);
      const int w = 5, L = 5;
      const int lv = kpL.octave;
      const uint8_t* IL = pyr_l[lv];
      const uint8_t* IR = pyr_r[lv];
      const int sL = lvl_stride[lv];
      const int y0 = (int)(scaledvL - w), xL0 = (int)(scaleduL - w);
      const float iniu = scaleduR0 + L - w, endu = scaleduR0 + L + w + 1;
      if (iniu < 0 || endu >= lvl_w[lv]) continue;
      int bestSad = INT_MAX, bestincR = 0;
      float vDists[2 * 5 + 1];
      const float cL = (float)IL[(size_t)(y0 + w) * sL + xL0 + w];
      for (int incR = -L; incR <= +L; incR++) {
        const int xR0 = (int)(scaleduR0 + incR - w);
        const float cR = (float)IR[(size_t)(y0 + w) * sL + xR0 + w];
        double acc = 0;  // cv::norm(NORM_L1) of exact small integers: exact in any accumulation order
        for (int yy = 0; yy < 2 * w + 1; yy++)
          for (int xx = 0; xx < 2 * w + 1; xx++) {
            const float a = (float)IL[(size_t)(y0 + yy) * sL + xL0 + xx] - cL;
            const float b = (float)IR[(size_t)(y0 + yy) * sL + xR0 + xx] - cR;
            acc += std::fabs(a - b);
          }
        const float dist = (float)acc;
        if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
        vDists[L + incR] = dist;
      }
      if (bestincR == -L || bestincR == L) continue;
      const float dist1 = vDists[L + bestincR - 1], dist2 = vDists[L + bestincR], dist3 = vDists[L + bestincR + 1];
      const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
      if (deltaR < -1 || deltaR > 1) continue;
      float bestuR = scale_factors[kpL.octave] * ((float)scaleduR0 + (float)bestincR + deltaR);
      float disparity = (uL - bestuR);
      if (disparity >= minD && disparity < maxD) {
        if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uL - 0.01); }  // :742-743: double expression, narrowed once
        depth[iL] = mbf / disparity;
        u_right[iL] = bestuR;
        vDistIdx.push_back(std::make_pair(bestSad, iL));
      }
    }
  }
  if (vDistIdx.empty()) return 0;  // (the reference reads vDistIdx[0] unconditionally)
  std::sort(vDistIdx.begin(), vDistIdx.end());
  const float median = (float)vDistIdx[vDistIdx.size() / 2].first;
  const float thDist = 1.5f * 1.4f * median;
  int kept = (int)vDistIdx.size();
  for (int i = (int)vDistIdx.size() - 1; i >= 0; i--) {
    if (vDistIdx[i].first < thDist) break;
    u_right[vDistIdx[i].second] = -1;
    depth[vDistIdx[i].second] = -1;
    kept--;
  }
  return kept;
}

// ======================================================================================================
// bag-of-words restricted searches and the distinctive descriptor of a map point.
// PINNED: tests/test_oracle_vs_ref_front.py runs the reference's own ORBmatcher::SearchByBoW (both overloads),
// SearchForTriangulation and MapPoint::ComputeDistinctiveDescriptors on real KeyFrame / MapPoint objects
// (R/src/KeyFrame.cpp compiled where it lies into oracle/_ref/libref_front.so) and requires identical results.
//
// A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) is passed flattened: node ids ascending,
// CSR offsets, feature indices in insertion order -- the iteration order of the map.
// ======================================================================================================
namespace orb_oracle {

struct FeatVec { const int32_t *nodes, *off, *idx; int n; };

// the merge-join both SearchByBoW overloads and SearchForTriangulation use (:232, :331-340, :654, :758-767):
// calls f(i1, i2) for every node id both vectors hold, ascending
template <class F>
static void for_common_nodes(const FeatVec& a, const FeatVec& b, F f) {
  int i = 0, j = 0;
  while (i < a.n && j < b.n) {
    if (a.nodes[i] == b.nodes[j]) { f(i, j); i++; j++; }
    else if (a.nodes[i] < b.nodes[j]) i = (int)(std::lower_bound(a.nodes, a.nodes + a.n, b.nodes[j]) - a.nodes);
    else j = (int)(std::lower_bound(b.nodes, b.nodes + b.n, a.nodes[i]) - b.nodes);
  }
}

}  // namespace orb_oracle

extern "C" {

/* block distances, same layout as orbx_match_blocks */
void orb_oracle_block_distances(const uint8_t* desc_a, const uint8_t* desc_b, const int32_t* a_off, const int32_t* a_idx,
                                const int32_t* b_off, const int32_t* b_idx, int nblocks, int64_t* dist_off, uint16_t* dist) {
  int64_t t = 0;
  dist_off[0] = 0;
  for (int k = 0; k < nblocks; k++) {
    for (int ia = a_off[k]; ia < a_off[k + 1]; ia++)
      for (int ib = b_off[k]; ib < b_off[k + 1]; ib++)
        dist[t++] = (uint16_t)descriptor_distance(desc_a + (size_t)a_idx[ia] * 32, desc_b + (size_t)b_idx[ib] * 32);
    dist_off[k + 1] = t;
  }
}

/* R/src/ORBmatcher.cpp:220-365. kf_usable[i] = vpMapPointsKF[i] && !isBad(). match_f[iF] = key-frame feature whose
 * map point was assigned to frame keypoint iF, or -1. Returns nmatches. */
int orb_oracle_search_by_bow_kf_frame(const uint8_t* desc_kf, const float* angle_kf, const uint8_t* kf_usable,
                                      const int32_t* kf_nodes, const int32_t* kf_off, const int32_t* kf_idx, int kf_nn,
                                      const uint8_t* desc_f, const float* angle_f, int n_f, const int32_t* f_nodes,
                                      const int32_t* f_off, const int32_t* f_idx, int f_nn, float nnratio,
                                      int check_orientation, int32_t* match_f) {
  for (int i = 0; i < n_f; i++) match_f[i] = -1;   // vpMapPointMatches = vector<MapPoint*>(F.N, NULL) (:224)
  std::vector<int> rotHist[kHistoLength];
  int nmatches = 0;
  const FeatVec kf{kf_nodes, kf_off, kf_idx, kf_nn}, f{f_nodes, f_off, f_idx, f_nn};
  for_common_nodes(kf, f, [&](int nk, int nf) {
    for (int iKF = kf_off[nk]; iKF < kf_off[nk + 1]; iKF++) {
      const int realIdxKF = kf_idx[iKF];
      if (!kf_usable[realIdxKF]) continue;                       // :247-253
      const uint8_t* dKF = desc_kf + (size_t)realIdxKF * 32;
      int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
      for (int iF = f_off[nf]; iF < f_off[nf + 1]; iF++) {
        const int realIdxF = f_idx[iF];
        if (match_f[realIdxF] >= 0) continue;                    // :265
        const int dist = descriptor_distance(dKF, desc_f + (size_t)realIdxF * 32);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 <= kThLow && (float)bestDist1 < nnratio * (float)bestDist2) {   // :285-287
        match_f[bestIdxF] = realIdxKF;
        if (check_orientation) rotHist[rot_bin(angle_kf[realIdxKF], angle_f[bestIdxF])].push_back(bestIdxF);
        nmatches++;
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, kHistoLength, ind1, ind2, ind3);
    for (int i = 0; i < kHistoLength; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int j : rotHist[i]) { match_f[j] = -1; nmatches--; }
    }
  }
  return nmatches;
}

/* R/src/ORBmatcher.cpp:632-772. usable1/2[i] = vpMapPoints[i] && !isBad(). matches12[i1] = i2 or -1. */
int orb_oracle_search_by_bow_kf_kf(const uint8_t* desc1, const float* angle1, const uint8_t* usable1, int n1,
                                   const int32_t* nodes1, const int32_t* off1, const int32_t* idx1, int nn1,
                                   const uint8_t* desc2, const float* angle2, const uint8_t* usable2, int n2,
                                   const int32_t* nodes2, const int32_t* off2, const int32_t* idx2, int nn2, float nnratio,
                                   int check_orientation, int32_t* matches12) {
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<uint8_t> vbMatched2((size_t)n2, 0);
  std::vector<int> rotHist[kHistoLength];
  int nmatches = 0;
  const FeatVec f1{nodes1, off1, idx1, nn1}, f2{nodes2, off2, idx2, nn2};
  for_common_nodes(f1, f2, [&](int k1, int k2) {
    for (int e1 = off1[k1]; e1 < off1[k1 + 1]; e1++) {
      const int i1 = idx1[e1];
      if (!usable1[i1]) continue;
      const uint8_t* d1 = desc1 + (size_t)i1 * 32;
      int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
      for (int e2 = off2[k2]; e2 < off2[k2 + 1]; e2++) {
        const int i2 = idx2[e2];
        if (vbMatched2[i2] || !usable2[i2]) continue;            // :682-688
        const int dist = descriptor_distance(d1, desc2 + (size_t)i2 * 32);
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = i2; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 < kThLow && (float)bestDist1 < nnratio * (float)bestDist2) {    // :707-709 (strict <)
        matches12[i1] = bestIdx2;
        vbMatched2[bestIdx2] = 1;
        if (check_orientation) rotHist[rot_bin(angle1[i1], angle2[bestIdx2])].push_back(i1);
        nmatches++;
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, kHistoLength, ind1, ind2, ind3);
    for (int i = 0; i < kHistoLength; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int j : rotHist[i]) { matches12[j] = -1; nmatches--; }
    }
  }
  return nmatches;
}

/* R/src/ORBmatcher.cpp:785-983 after its prologue (:790-803: epipole ex, ey of camera 1 in image 2).
 * CheckDistEpipolarLine :175-203 with the contractions GCC -O3 emits for it (vfmadd231ss: a = fma(x,F00,y*F10)+F20,
 * den = fma(a,a,b*b), num = fma(b,y2,a*x2) + c; checked in the disassembly of the reference's object file).
 * has_mp1/2[i] = GetMapPoint(i) != NULL; uright1/2 = mvuRight; kps = mvKeysUn. matches12[i1] = i2 or -1. */
int orb_oracle_search_for_triangulation(const orb_oracle_keypoint* kps1, const uint8_t* desc1, const uint8_t* has_mp1,
                                        const float* uright1, int n1, const int32_t* nodes1, const int32_t* off1,
                                        const int32_t* idx1, int nn1, const orb_oracle_keypoint* kps2, const uint8_t* desc2,
                                        const uint8_t* has_mp2, const float* uright2, int n2, const int32_t* nodes2,
                                        const int32_t* off2, const int32_t* idx2, int nn2, const float* F12 /* 3x3 */,
                                        float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                        int only_stereo, int check_orientation, int32_t* matches12) {
  for (int i = 0; i < n1; i++) matches12[i] = -1;
  std::vector<uint8_t> vbMatched2((size_t)n2, 0);
  std::vector<int> rotHist[kHistoLength];
  int nmatches = 0;
  auto epipolar_ok = [&](const orb_oracle_keypoint& kp1, const orb_oracle_keypoint& kp2) {
    const float a = std::fma(kp1.x, F12[0], kp1.y * F12[3]) + F12[6];
    const float b = std::fma(kp1.x, F12[1], kp1.y * F12[4]) + F12[7];
    const float c = std::fma(kp1.y, F12[5], kp1.x * F12[2]) + F12[8];
    const float num = c + std::fma(b, kp2.y, a * kp2.x);
    const float den = std::fma(a, a, b * b);
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return (double)dsqr < 3.84 * (double)level_sigma2_2[kp2.octave];
  };
  const FeatVec f1{nodes1, off1, idx1, nn1}, f2{nodes2, off2, idx2, nn2};
  for_common_nodes(f1, f2, [&](int k1, int k2) {
    for (int e1 = off1[k1]; e1 < off1[k1 + 1]; e1++) {
      const int i1 = idx1[e1];
      if (has_mp1[i1]) continue;                                  // :833-836
      const bool bStereo1 = uright1[i1] >= 0;
      if (only_stereo && !bStereo1) continue;
      const orb_oracle_keypoint& kp1 = kps1[i1];
      const uint8_t* d1 = desc1 + (size_t)i1 * 32;
      int bestDist = kThLow, bestIdx2 = -1;
      for (int e2 = off2[k2]; e2 < off2[k2 + 1]; e2++) {
        const int i2 = idx2[e2];
        if (vbMatched2[i2] || has_mp2[i2]) continue;              // :858-859
        const bool bStereo2 = uright2[i2] >= 0;
        if (only_stereo && !bStereo2) continue;
        const int dist = descriptor_distance(d1, desc2 + (size_t)i2 * 32);
        if (dist > kThLow || dist > bestDist) continue;           // :872
        const orb_oracle_keypoint& kp2 = kps2[i2];
        if (!bStereo1 && !bStereo2) {
          const float distex = ex - kp2.x, distey = ey - kp2.y;
          if (std::fma(distex, distex, distey * distey) < 100 * scale_factors2[kp2.octave]) continue;   // :881-883
        }
        if (epipolar_ok(kp1, kp2)) { bestIdx2 = i2; bestDist = dist; }
      }
      if (bestIdx2 >= 0) {
        matches12[i1] = bestIdx2;
        vbMatched2[bestIdx2] = 1;
        nmatches++;
        if (check_orientation) rotHist[rot_bin(kp1.angle, kps2[bestIdx2].angle)].push_back(i1);
      }
    }
  });
  if (check_orientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    compute_three_maxima(rotHist, kHistoLength, ind1, ind2, ind3);
    for (int i = 0; i < kHistoLength; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (int j : rotHist[i]) { matches12[j] = -1; nmatches--; }
    }
  }
  return nmatches;
}

/* R/src/MapPoint.cpp:306-385 for npoints points: rows desc[obs_off[p] .. obs_off[p+1]) in observation order */
void orb_oracle_distinctive(const uint8_t* desc, const int32_t* obs_off, int npoints, int32_t* best, int32_t* best_median) {
  for (int p = 0; p < npoints; p++) {
    const uint8_t* D = desc + (size_t)obs_off[p] * 32;
    const size_t N = (size_t)(obs_off[p + 1] - obs_off[p]);
    if (N == 0) { best[p] = -1; best_median[p] = -1; continue; }
    std::vector<std::vector<float>> Distances(N, std::vector<float>(N, 0));
    for (size_t i = 0; i < N; i++) {
      Distances[i][i] = 0;
      for (size_t j = i + 1; j < N; j++) {
        const int distij = descriptor_distance(D + i * 32, D + j * 32);
        Distances[i][j] = (float)distij;
        Distances[j][i] = (float)distij;
      }
    }
    int BestMedian = INT_MAX, BestIdx = 0;
    for (size_t i = 0; i < N; i++) {
      std::vector<int> vDists(Distances[i].begin(), Distances[i].end());
      std::sort(vDists.begin(), vDists.end());
      const int median = vDists[(size_t)(0.5 * (double)(N - 1))];
      if (median < BestMedian) { BestMedian = median; BestIdx = (int)i; }
    }
    best[p] = BestIdx;
    best_median[p] = BestMedian;
  }
}

}  // extern "C"

extern "C" {

/* The windowed best-match loop shared by ORBmatcher::Fuse (R/src/ORBmatcher.cpp:995-1160, with its chi-square
 * gates :1070-1099; :1164-1296 without), SearchByProjection(KeyFrame*, Scw, ...) (:370-497, `taken` = vpMatched)
 * and both directions of SearchBySim3 (:1305-1547), after their projection prologues: per query (u, v), radius =
 * th * mvScaleFactors[level], predicted level, ur = u - bf*invz (gate only), descriptor. Candidates =
 * KeyFrame::GetFeaturesInArea(u, v, radius) (R/src/KeyFrame.cpp:702-743) with octave in [level-1, level].
 * inv_level_sigma2 != NULL switches the gates on: e2 = ex*ex + ey*ey (+ er*er when mvuRight[idx] >= 0) with the
 * contractions of the reference's object code (fma(ex,ex,ey*ey), fma(er,er,.)), rejected when e2*invSigma2 > 5.99
 * (7.8 stereo). taken != NULL: keypoints with taken[idx] are skipped and a query that ends with best <= th takes its
 * keypoint (:453-454, :479-483). PINNED: tests/test_oracle_vs_ref_bow.py (Fuse and SearchByProjection on real objects). */
void orb_oracle_search_best_in_window(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n, const float* uright,
                                      float min_x, float max_x, float min_y, float max_y, const float* inv_level_sigma2,
                                      const float* q_uv, const float* q_ur, const float* q_radius, const int32_t* q_level,
                                      const uint8_t* q_desc, int nq, uint8_t* taken, int th, int32_t* best_idx,
                                      int32_t* best_dist) {
  Grid g;
  g.build(kps_un, n, min_x, max_x, min_y, max_y);
  std::vector<int> vIndices;
  for (int i = 0; i < nq; i++) {
    const float u = q_uv[2 * i], v = q_uv[2 * i + 1];
    const int nPredictedLevel = q_level[i];
    g.query(u, v, q_radius[i], -1, -1, vIndices);
    int bestDist = 256, bestIdx = -1;
    for (int idx : vIndices) {
      if (taken && taken[idx]) continue;
      const orb_oracle_keypoint& kp = kps_un[idx];
      const int kpLevel = kp.octave;
      if (kpLevel < nPredictedLevel - 1 || kpLevel > nPredictedLevel) continue;
      if (inv_level_sigma2) {
        const float ex = u - kp.x, ey = v - kp.y;
        const float e2m = std::fma(ex, ex, ey * ey);
        if (uright && uright[idx] >= 0) {
          const float er = q_ur[i] - uright[idx];
          const float e2 = std::fma(er, er, e2m);
          if ((double)(e2 * inv_level_sigma2[kpLevel]) > 7.8) continue;
        } else {
          if ((double)(e2m * inv_level_sigma2[kpLevel]) > 5.99) continue;
        }
      }
      const int dist = descriptor_distance(q_desc + (size_t)i * 32, desc + (size_t)idx * 32);
      if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
    }
    best_idx[i] = bestIdx;
    best_dist[i] = bestDist;
    if (taken && bestDist <= th) taken[bestIdx] = 1;
  }
}

}  // extern "C"
