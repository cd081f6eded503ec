// TEST INFRASTRUCTURE (oracle) -- CPU restatement of ORB_SLAM2::ORBextractor.
//
// Follows R/src/ORBextractor.cpp: ctor :418-502, ComputePyramid :1197-1229,
// ComputeKeyPointsOctTree :819-921, DistributeOctTree :571-817, ExtractorNode::DivideNode :513-569,
// IC_Angle :79-108, computeOrbDescriptor :113-155, operator() :1120-1188. The OpenCV calls go to
// the restated primitives in cv_prims.cpp. Float expressions are written with the types and
// operation order of the reference; where the reference's build (-O3 -march=native, GCC's default
// -ffp-contract=fast) fuses a multiply-add, std::fma is written explicitly (this file is compiled
// with -ffp-contract=off so nothing else fuses).
#include <algorithm>
#include <cmath>
#include <cstring>
#include <list>
#include <utility>
#include <vector>

#include "cv_prims.h"
#include "orb_oracle.h"

namespace orb_oracle {

static const int PATCH_SIZE = 31;
static const int HALF_PATCH_SIZE = 15;
static const int EDGE_THRESHOLD = 19;

static const int8_t kPattern[1024] = {
#include "../orb-slam2-_b200/csrc/orb_pattern.inc"
};

struct Plane {
  int w = 0, h = 0;
  std::vector<uint8_t> px;
  const uint8_t* row(int y) const { return px.data() + (size_t)y * w; }
  uint8_t* row(int y) { return px.data() + (size_t)y * w; }
};

struct Cand { int x, y, score; };  // level coordinates relative to minBorder; FAST response

struct KP {
  float x, y, size, angle, response;
  int octave;
};

// ---- DistributeOctTree (R/src/ORBextractor.cpp:571-817) ---------------------------------------
struct Node {
  std::vector<int> keys;  // indices into the candidate array, in insertion order
  int ULx, ULy, URx, URy, BLx, BLy, BRx, BRy;
  std::list<Node>::iterator lit;
  bool noMore = false;
  long seq = 0;  // creation sequence number: the canonical replacement for the heap address
};

static void divide_node(const Node& p, const std::vector<Cand>& c, Node& n1, Node& n2, Node& n3, Node& n4) {
  // :515-516
  const int halfX = (int)std::ceil(static_cast<float>(p.URx - p.ULx) / 2);
  const int halfY = (int)std::ceil(static_cast<float>(p.BRy - p.ULy) / 2);
  n1.ULx = p.ULx; n1.ULy = p.ULy;
  n1.URx = p.ULx + halfX; n1.URy = p.ULy;
  n1.BLx = p.ULx; n1.BLy = p.ULy + halfY;
  n1.BRx = p.ULx + halfX; n1.BRy = p.ULy + halfY;

  n2.ULx = n1.URx; n2.ULy = n1.URy;
  n2.URx = p.URx; n2.URy = p.URy;
  n2.BLx = n1.BRx; n2.BLy = n1.BRy;
  n2.BRx = p.URx; n2.BRy = p.ULy + halfY;

  n3.ULx = n1.BLx; n3.ULy = n1.BLy;
  n3.URx = n1.BRx; n3.URy = n1.BRy;
  n3.BLx = p.BLx; n3.BLy = p.BLy;
  n3.BRx = n1.BRx; n3.BRy = p.BLy;

  n4.ULx = n3.URx; n4.ULy = n3.URy;
  n4.URx = n2.BRx; n4.URy = n2.BRy;
  n4.BLx = n3.BRx; n4.BLy = n3.BRy;
  n4.BRx = p.BRx; n4.BRy = p.BRy;

  for (int idx : p.keys) {  // :544-558; pt.x/pt.y are integer-valued floats, compared with ints
    const float px = (float)c[idx].x, py = (float)c[idx].y;
    if (px < n1.URx) {
      if (py < n1.BRy) n1.keys.push_back(idx);
      else n3.keys.push_back(idx);
    } else if (py < n1.BRy) {
      n2.keys.push_back(idx);
    } else {
      n4.keys.push_back(idx);
    }
  }
  if (n1.keys.size() == 1) n1.noMore = true;
  if (n2.keys.size() == 1) n2.noMore = true;
  if (n3.keys.size() == 1) n3.noMore = true;
  if (n4.keys.size() == 1) n4.noMore = true;
}

struct SortEntry {
  int size;
  long key;  // canonical: seq (tie_mode 0) or -seq (tie_mode 1)
  Node* node;
  bool operator<(const SortEntry& o) const { return size != o.size ? size < o.size : key < o.key; }
};

// Returns indices of the kept candidates in list order. *tie_sensitive counts phase-2 rounds that
// stopped (break at size>=N) inside a run of equal-size nodes, i.e. where the reference's pointer
// tie-break could have picked a different node.
static std::vector<int> distribute_oct_tree(const std::vector<Cand>& cands, int minX, int maxX, int minY,
                                            int maxY, int N, int tie_mode, int* tie_sensitive) {
  if (tie_sensitive) *tie_sensitive = 0;
  std::vector<int> result;
  if (cands.empty()) return result;  // reference would walk an empty list and return nothing
  long seq = 0;
  const int nIni = (int)std::round(static_cast<float>(maxX - minX) / (maxY - minY));  // :577
  const float hX = static_cast<float>(maxX - minX) / nIni;                             // :580

  std::list<Node> lNodes;
  std::vector<Node*> vpIni(nIni > 0 ? nIni : 0);
  for (int i = 0; i < nIni; i++) {
    Node ni;
    ni.ULx = (int)(hX * static_cast<float>(i)); ni.ULy = 0;
    ni.URx = (int)(hX * static_cast<float>(i + 1)); ni.URy = 0;
    ni.BLx = ni.ULx; ni.BLy = maxY - minY;
    ni.BRx = ni.URx; ni.BRy = maxY - minY;
    ni.seq = seq++;
    lNodes.push_back(ni);
    vpIni[i] = &lNodes.back();
  }
  for (size_t i = 0; i < cands.size(); i++) {
    const float px = (float)cands[i].x;
    vpIni[(int)(px / hX)]->keys.push_back((int)i);  // :611
  }
  for (auto lit = lNodes.begin(); lit != lNodes.end();) {
    if (lit->keys.size() == 1) { lit->noMore = true; ++lit; }
    else if (lit->keys.empty()) lit = lNodes.erase(lit);
    else ++lit;
  }

  bool finish = false;
  std::vector<SortEntry> vSize;
  auto push_child = [&](Node& n, int* nToExpand) {
    if (n.keys.empty()) return;
    n.seq = seq++;
    lNodes.push_front(n);
    if (n.keys.size() > 1) {
      if (nToExpand) (*nToExpand)++;
      Node* p = &lNodes.front();
      vSize.push_back({(int)n.keys.size(), tie_mode == 0 ? p->seq : -p->seq, p});
      lNodes.front().lit = lNodes.begin();
    }
  };

  while (!finish) {
    int prevSize = (int)lNodes.size();
    auto lit = lNodes.begin();
    int nToExpand = 0;
    vSize.clear();
    while (lit != lNodes.end()) {
      if (lit->noMore) { ++lit; continue; }
      Node n1, n2, n3, n4;
      divide_node(*lit, cands, n1, n2, n3, n4);
      push_child(n1, &nToExpand);
      push_child(n2, &nToExpand);
      push_child(n3, &nToExpand);
      push_child(n4, &nToExpand);
      lit = lNodes.erase(lit);
    }
    if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) {
      finish = true;
    } else if (((int)lNodes.size() + nToExpand * 3) > N) {
      while (!finish) {
        prevSize = (int)lNodes.size();
        std::vector<SortEntry> vPrev = vSize;
        vSize.clear();
        std::sort(vPrev.begin(), vPrev.end());
        for (int j = (int)vPrev.size() - 1; j >= 0; j--) {
          Node n1, n2, n3, n4;
          divide_node(*vPrev[j].node, cands, n1, n2, n3, n4);
          push_child(n1, nullptr);
          push_child(n2, nullptr);
          push_child(n3, nullptr);
          push_child(n4, nullptr);
          lNodes.erase(vPrev[j].node->lit);
          if ((int)lNodes.size() >= N) {
            // stopped at j: tie-sensitive if an unsplit entry has the same size as a split one
            if (tie_sensitive && j > 0 && vPrev[j - 1].size == vPrev[j].size) (*tie_sensitive)++;
            break;
          }
        }
        if ((int)lNodes.size() >= N || (int)lNodes.size() == prevSize) finish = true;
      }
    }
  }

  result.reserve(lNodes.size());
  for (auto& n : lNodes) {  // :798-814 first point with the maximum response
    int best = n.keys[0];
    float maxResponse = (float)cands[best].score;
    for (size_t k = 1; k < n.keys.size(); k++) {
      if ((float)cands[n.keys[k]].score > maxResponse) {
        best = n.keys[k];
        maxResponse = (float)cands[best].score;
      }
    }
    result.push_back(best);
  }
  return result;
}

// ---- the extractor ----------------------------------------------------------------------------
struct Extractor {
  int nfeatures;
  double scaleFactor;  // the reference stores the float ctor argument in a double member (.h:106)
  int nlevels, iniThFAST, minThFAST;
  int tie_mode = 0;
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  std::vector<int> mnFeaturesPerLevel, umax;
  std::vector<Plane> pyramid, blurred;
  std::vector<std::vector<Cand>> levelCands;
  std::vector<int> levelKept, levelTie;

  Extractor(int nf, float sf, int nl, int ini, int mn)
      : nfeatures(nf), scaleFactor(sf), nlevels(nl), iniThFAST(ini), minThFAST(mn) {
    mvScaleFactor.resize(nlevels);
    mvLevelSigma2.resize(nlevels);
    mvScaleFactor[0] = 1.0f;
    mvLevelSigma2[0] = 1.0f;
    for (int i = 1; i < nlevels; i++) {
      mvScaleFactor[i] = (float)(mvScaleFactor[i - 1] * scaleFactor);  // float*double -> float
      mvLevelSigma2[i] = mvScaleFactor[i] * mvScaleFactor[i];
    }
    mvInvScaleFactor.resize(nlevels);
    mvInvLevelSigma2.resize(nlevels);
    for (int i = 0; i < nlevels; i++) {
      mvInvScaleFactor[i] = 1.0f / mvScaleFactor[i];
      mvInvLevelSigma2[i] = 1.0f / mvLevelSigma2[i];
    }
    pyramid.resize(nlevels);
    blurred.resize(nlevels);
    levelCands.resize(nlevels);
    levelKept.assign(nlevels, 0);
    levelTie.assign(nlevels, 0);

    mnFeaturesPerLevel.resize(nlevels);
    float factor = (float)(1.0f / scaleFactor);
    float nDesired = nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)nlevels));
    int sumFeatures = 0;
    for (int level = 0; level < nlevels - 1; level++) {
      mnFeaturesPerLevel[level] = cv_round_f(nDesired);
      sumFeatures += mnFeaturesPerLevel[level];
      nDesired *= factor;
    }
    mnFeaturesPerLevel[nlevels - 1] = std::max(nfeatures - sumFeatures, 0);

    // :480-500 end of each row of the circular patch
    umax.resize(HALF_PATCH_SIZE + 1);
    int v, v0, vmax = (int)std::floor(HALF_PATCH_SIZE * std::sqrt(2.f) / 2 + 1);
    int vmin = (int)std::ceil(HALF_PATCH_SIZE * std::sqrt(2.f) / 2);
    const double hp2 = HALF_PATCH_SIZE * HALF_PATCH_SIZE;
    for (v = 0; v <= vmax; ++v) umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (v = HALF_PATCH_SIZE, v0 = 0; v >= vmin; --v) {
      while (umax[v0] == umax[v0 + 1]) ++v0;
      umax[v] = v0;
      ++v0;
    }
  }

  void compute_pyramid(const uint8_t* img, int w, int h, int stride) {  // :1197-1229
    for (int level = 0; level < nlevels; ++level) {
      float scale = mvInvScaleFactor[level];
      int lw = cv_round_f((float)w * scale), lh = cv_round_f((float)h * scale);
      Plane& P = pyramid[level];
      P.w = lw; P.h = lh;
      P.px.resize((size_t)lw * lh);
      if (level != 0) {
        const Plane& S = pyramid[level - 1];
        resize_linear_u8(S.px.data(), S.w, S.h, S.w, P.px.data(), lw, lh, lw);
      } else {
        for (int y = 0; y < h; y++) std::memcpy(P.row(y), img + (size_t)y * stride, w);  // :1223
      }
      // copyMakeBorder results are never read downstream: omitted.
    }
  }

  float ic_angle(const Plane& P, float ptx, float pty) const {  // :79-108
    int m_01 = 0, m_10 = 0;
    const int step = P.w;
    const uint8_t* center = P.row(cv_round_f(pty)) + cv_round_f(ptx);
    for (int u = -HALF_PATCH_SIZE; u <= HALF_PATCH_SIZE; ++u) m_10 += u * center[u];
    for (int v = 1; v <= HALF_PATCH_SIZE; ++v) {
      int v_sum = 0, d = umax[v];
      for (int u = -d; u <= d; ++u) {
        int val_plus = center[u + v * step], val_minus = center[u - v * step];
        v_sum += (val_plus - val_minus);
        m_10 += u * (val_plus + val_minus);
      }
      m_01 += v * v_sum;
    }
    return fast_atan2((float)m_01, (float)m_10);
  }

  void compute_keypoints(std::vector<std::vector<KP>>& all) {  // :819-921
    all.assign(nlevels, {});
    const float W = 30;
    for (int level = 0; level < nlevels; ++level) {
      const Plane& P = pyramid[level];
      const int minBorderX = EDGE_THRESHOLD - 3, minBorderY = minBorderX;
      const int maxBorderX = P.w - EDGE_THRESHOLD + 3, maxBorderY = P.h - EDGE_THRESHOLD + 3;
      std::vector<Cand>& toDistribute = levelCands[level];
      toDistribute.clear();
      levelKept[level] = 0;
      levelTie[level] = 0;

      const float width = (float)(maxBorderX - minBorderX), height = (float)(maxBorderY - minBorderY);
      const int nCols = (int)(width / W), nRows = (int)(height / W);
      if (nCols <= 0 || nRows <= 0) continue;  // reference divides by zero here; out of contract
      const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);

      std::vector<FastPoint> cell;
      for (int i = 0; i < nRows; i++) {
        const float iniY = (float)(minBorderY + i * hCell);
        float maxY = iniY + hCell + 6;
        if (iniY >= maxBorderY - 3) continue;
        if (maxY > maxBorderY) maxY = (float)maxBorderY;
        for (int j = 0; j < nCols; j++) {
          const float iniX = (float)(minBorderX + j * wCell);
          float maxX = iniX + wCell + 6;
          if (iniX >= maxBorderX - 6) continue;
          if (maxX > maxBorderX) maxX = (float)maxBorderX;
          const int x0 = (int)iniX, x1 = (int)maxX, y0 = (int)iniY, y1 = (int)maxY;
          const uint8_t* roi = P.row(y0) + x0;
          fast9_16(roi, x1 - x0, y1 - y0, P.w, iniThFAST, true, cell);
          if (cell.empty()) fast9_16(roi, x1 - x0, y1 - y0, P.w, minThFAST, true, cell);
          for (const FastPoint& fp : cell) toDistribute.push_back({fp.x + j * wCell, fp.y + i * hCell, fp.score});
        }
      }

      int tie = 0;
      std::vector<int> kept = distribute_oct_tree(toDistribute, minBorderX, maxBorderX, minBorderY, maxBorderY,
                                                  mnFeaturesPerLevel[level], tie_mode, &tie);
      levelKept[level] = (int)kept.size();
      levelTie[level] = tie;
      const int scaledPatchSize = (int)(PATCH_SIZE * mvScaleFactor[level]);
      std::vector<KP>& kps = all[level];
      kps.reserve(kept.size());
      for (int idx : kept) {
        KP k;
        k.x = (float)toDistribute[idx].x + minBorderX;
        k.y = (float)toDistribute[idx].y + minBorderY;
        k.size = (float)scaledPatchSize;
        k.angle = -1.f;
        k.response = (float)toDistribute[idx].score;
        k.octave = level;
        kps.push_back(k);
      }
    }
    for (int level = 0; level < nlevels; ++level)
      for (KP& k : all[level]) k.angle = ic_angle(pyramid[level], k.x, k.y);
  }

  void describe(const Plane& B, const KP& k, uint8_t* desc) const {  // :113-155
    const float factorPI = (float)(M_PI / 180.f);
    float angle = k.angle * factorPI;
    float a, b;
    ::sincosf(angle, &b, &a);  // GCC merges the reference's cos()/sin() pair into one sincosf call
    const int step = B.w;
    const uint8_t* center = B.row(cv_round_f(k.y)) + cv_round_f(k.x);
    const int8_t* pat = kPattern;
    auto get = [&](int idx) -> int {
      const float px = (float)pat[2 * idx], py = (float)pat[2 * idx + 1];
      // reference: cvRound(x*b + y*a), cvRound(x*a - y*b); -O3 -march=native contracts these to
      // vfmadd231ss / vfmsub132ss: fma(x,b,y*a), fma(x,a,-(y*b))
      const int r = cv_round_f(std::fma(px, b, py * a));
      const int c = cv_round_f(std::fma(px, a, -(py * b)));
      return center[r * step + c];
    };
    for (int i = 0; i < 32; ++i, pat += 32) {
      int val = 0;
      for (int bit = 0; bit < 8; bit++) {
        int t0 = get(2 * bit), t1 = get(2 * bit + 1);
        val |= (t0 < t1) << bit;
      }
      desc[i] = (uint8_t)val;
    }
  }

  int extract(const uint8_t* img, int w, int h, int stride, orb_oracle_keypoint* out, uint8_t* desc, int cap) {
    if (!img || w <= 0 || h <= 0) return 0;
    compute_pyramid(img, w, h, stride);
    std::vector<std::vector<KP>> all;
    compute_keypoints(all);
    int total = 0;
    for (auto& v : all) total += (int)v.size();
    if (total > cap) return -total;
    int offset = 0;
    for (int level = 0; level < nlevels; ++level) {
      std::vector<KP>& kps = all[level];
      Plane& B = blurred[level];
      if (kps.empty()) { B.w = B.h = 0; B.px.clear(); continue; }
      const Plane& P = pyramid[level];
      B.w = P.w; B.h = P.h;
      B.px.resize(P.px.size());
      gaussian_blur7_u8(P.px.data(), P.w, P.h, P.w, B.px.data(), B.w);
      for (size_t i = 0; i < kps.size(); i++) describe(B, kps[i], desc + (size_t)(offset + i) * 32);
      if (level != 0) {
        float scale = mvScaleFactor[level];
        for (KP& k : kps) { k.x *= scale; k.y *= scale; }
      }
      for (size_t i = 0; i < kps.size(); i++) {
        orb_oracle_keypoint& o = out[offset + i];
        o.x = kps[i].x; o.y = kps[i].y; o.size = kps[i].size; o.angle = kps[i].angle;
        o.response = kps[i].response; o.octave = kps[i].octave; o.class_id = -1;
      }
      offset += (int)kps.size();
    }
    return total;
  }
};

}  // namespace orb_oracle

using namespace orb_oracle;

struct orb_oracle_extractor { Extractor ex; };

extern "C" {

void orb_oracle_resize(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride) {
  resize_linear_u8(src, sw, sh, sstride, dst, dw, dh, dstride);
}
void orb_oracle_blur7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
  gaussian_blur7_u8(src, w, h, sstride, dst, dstride);
}
int orb_oracle_fast(const uint8_t* img, int w, int h, int stride, int threshold, int nms, int32_t* xys, int cap) {
  std::vector<FastPoint> out;
  fast9_16(img, w, h, stride, threshold, nms != 0, out);
  for (int i = 0; i < (int)out.size() && i < cap; i++) {
    xys[3 * i] = out[i].x; xys[3 * i + 1] = out[i].y; xys[3 * i + 2] = out[i].score;
  }
  return (int)out.size();
}
float orb_oracle_fast_atan2(float y, float x) { return fast_atan2(y, x); }
void orb_oracle_sincosf_port(float a, float* s, float* c) { sincosf_port(a, s, c); }
void orb_oracle_sincosf_libm(float a, float* s, float* c) { ::sincosf(a, s, c); }
int orb_oracle_cv_round(double v) { return cv_round(v); }

orb_oracle_extractor* orb_oracle_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
  if (nfeatures <= 0 || nlevels <= 0 || nlevels > 32 || !(scale_factor > 1.f)) return nullptr;
  return new orb_oracle_extractor{Extractor(nfeatures, scale_factor, nlevels, ini_th, min_th)};
}
void orb_oracle_extractor_destroy(orb_oracle_extractor* e) { delete e; }
void orb_oracle_extractor_set_tie_mode(orb_oracle_extractor* e, int m) { e->ex.tie_mode = m; }
int orb_oracle_extract(orb_oracle_extractor* e, const uint8_t* img, int w, int h, int stride,
                       orb_oracle_keypoint* kps, uint8_t* desc, int cap) {
  return e->ex.extract(img, w, h, stride, kps, desc, cap);
}
void orb_oracle_extractor_tables(orb_oracle_extractor* e, float* scale, float* inv_scale, float* sigma2,
                                 float* inv_sigma2, int32_t* feats, int32_t* umax16) {
  const Extractor& x = e->ex;
  for (int i = 0; i < x.nlevels; i++) {
    if (scale) scale[i] = x.mvScaleFactor[i];
    if (inv_scale) inv_scale[i] = x.mvInvScaleFactor[i];
    if (sigma2) sigma2[i] = x.mvLevelSigma2[i];
    if (inv_sigma2) inv_sigma2[i] = x.mvInvLevelSigma2[i];
    if (feats) feats[i] = x.mnFeaturesPerLevel[i];
  }
  if (umax16) for (int i = 0; i < 16; i++) umax16[i] = x.umax[i];
}
void orb_oracle_level_size(orb_oracle_extractor* e, int level, int* w, int* h) {
  *w = e->ex.pyramid[level].w; *h = e->ex.pyramid[level].h;
}
void orb_oracle_get_level(orb_oracle_extractor* e, int level, uint8_t* dst, int dstride) {
  const Plane& P = e->ex.pyramid[level];
  for (int y = 0; y < P.h; y++) std::memcpy(dst + (size_t)y * dstride, P.row(y), P.w);
}
void orb_oracle_get_blurred_level(orb_oracle_extractor* e, int level, uint8_t* dst, int dstride) {
  const Plane& P = e->ex.blurred[level];
  for (int y = 0; y < P.h; y++) std::memcpy(dst + (size_t)y * dstride, P.row(y), P.w);
}
int orb_oracle_level_candidates(orb_oracle_extractor* e, int level, int32_t* xys, int cap) {
  const auto& c = e->ex.levelCands[level];
  for (int i = 0; i < (int)c.size() && i < cap; i++) {
    xys[3 * i] = c[i].x; xys[3 * i + 1] = c[i].y; xys[3 * i + 2] = c[i].score;
  }
  return (int)c.size();
}
int orb_oracle_level_kept(orb_oracle_extractor* e, int level) { return e->ex.levelKept[level]; }
int orb_oracle_level_tie_sensitive(orb_oracle_extractor* e, int level) { return e->ex.levelTie[level]; }

int orb_oracle_distribute(const int32_t* xys, int n, int minX, int maxX, int minY, int maxY, int N,
                          int tie_mode, int32_t* out_idx, int cap) {
  std::vector<Cand> c(n);
  for (int i = 0; i < n; i++) c[i] = {xys[3 * i], xys[3 * i + 1], xys[3 * i + 2]};
  std::vector<int> kept = distribute_oct_tree(c, minX, maxX, minY, maxY, N, tie_mode, nullptr);
  for (int i = 0; i < (int)kept.size() && i < cap; i++) out_idx[i] = kept[i];
  return (int)kept.size();
}

}  // extern "C"
