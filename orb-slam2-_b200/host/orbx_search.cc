// See orbx_search.h. Control flow follows R/src/ORBmatcher.cpp line by line; the candidate lists and the
// Hamming distances come from the device.
#include "orbx_search.h"

#include <algorithm>
#include <climits>
#include <cmath>
#include <cstdio>
#include <cstdlib>

namespace orbx_host {

namespace {

struct Candidates {
  std::vector<int> off, idx, dist;
};

void Fail(const char* what) {
  std::fprintf(stderr, "orbx_host: %s: %s\n", what, orbx_last_error());
  std::abort();  // no CPU fallback
}

void WindowSearch(orbx_matcher* m, const FrameView& F, const std::vector<float>& qxy, const std::vector<float>& qr,
                  const std::vector<int>& qmin, const std::vector<int>& qmax, const unsigned char* qdesc, int nq,
                  Candidates* c) {
  c->off.assign(nq + 1, 0);
  if (nq == 0) return;
  int cap = 64 * nq, need = 0;
  for (;;) {
    c->idx.resize(cap);
    c->dist.resize(cap);
    const int rc = orbx_match_window(m, F.keysUn, F.descriptors, F.N, F.minX, F.maxX, F.minY, F.maxY, &qxy[0], &qr[0],
                                     &qmin[0], &qmax[0], qdesc, nq, &c->off[0], &c->idx[0], &c->dist[0], cap, &need);
    if (rc == ORBX_ERR_CAPACITY && need > cap) { cap = need; continue; }
    if (rc != ORBX_OK) Fail("orbx_match_window");
    return;
  }
}

int RotBin(float a1, float a2) {  // :569-575
  const float factor = HISTO_LENGTH / 360.0f;
  float rot = a1 - a2;
  if (rot < 0.0) rot += 360.0f;
  int bin = (int)roundf(rot * factor);
  if (bin == HISTO_LENGTH) bin = 0;
  return bin;
}

}  // namespace

void ComputeThreeMaxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
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

int SearchForInitialization(orbx_matcher* m, const FrameView& F1, const FrameView& F2, float* vbPrevMatched,
                            int* vnMatches12, int windowSize, float nnratio, bool checkOrientation) {
  int nmatches = 0;
  for (int i = 0; i < F1.N; i++) vnMatches12[i] = -1;
  // queries: the level-0 keypoints of F1 (:515-517), window around vbPrevMatched (:519)
  std::vector<int> qOf;  // query -> i1
  std::vector<float> qxy, qr;
  std::vector<int> qmin, qmax;
  std::vector<unsigned char> qdesc;
  for (int i1 = 0; i1 < F1.N; i1++) {
    if (F1.keysUn[i1].octave > 0) continue;
    qOf.push_back(i1);
    qxy.push_back(vbPrevMatched[2 * i1]);
    qxy.push_back(vbPrevMatched[2 * i1 + 1]);
    qr.push_back((float)windowSize);
    qmin.push_back(F1.keysUn[i1].octave);
    qmax.push_back(F1.keysUn[i1].octave);
    qdesc.insert(qdesc.end(), F1.descriptors + (size_t)i1 * 32, F1.descriptors + (size_t)i1 * 32 + 32);
  }
  Candidates c;
  WindowSearch(m, F2, qxy, qr, qmin, qmax, qdesc.empty() ? NULL : &qdesc[0], (int)qOf.size(), &c);

  std::vector<int> rotHist[HISTO_LENGTH];
  std::vector<int> vMatchedDistance(F2.N, INT_MAX), vnMatches21(F2.N, -1);
  for (size_t q = 0; q < qOf.size(); q++) {
    const int i1 = qOf[q];
    if (c.off[q] == c.off[q + 1]) continue;
    int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
    for (int k = c.off[q]; k < c.off[q + 1]; k++) {
      const int i2 = c.idx[k], dist = c.dist[k];
      if (vMatchedDistance[i2] <= dist) continue;
      if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
      else if (dist < bestDist2) { bestDist2 = dist; }
    }
    if (bestDist <= TH_LOW) {
      if (bestDist < (float)bestDist2 * nnratio) {
        if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; nmatches--; }
        vnMatches12[i1] = bestIdx2;
        vnMatches21[bestIdx2] = i1;
        vMatchedDistance[bestIdx2] = bestDist;
        nmatches++;
        if (checkOrientation) rotHist[RotBin(F1.keysUn[i1].angle, F2.keysUn[bestIdx2].angle)].push_back(i1);
      }
    }
  }
  if (checkOrientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++) {
      if (i == ind1 || i == ind2 || i == ind3) continue;
      for (size_t j = 0; j < rotHist[i].size(); j++) {
        const int idx1 = rotHist[i][j];
        if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; nmatches--; }
      }
    }
  }
  for (int i1 = 0; i1 < F1.N; i1++)
    if (vnMatches12[i1] >= 0) {
      vbPrevMatched[2 * i1] = F2.keysUn[vnMatches12[i1]].x;
      vbPrevMatched[2 * i1 + 1] = F2.keysUn[vnMatches12[i1]].y;
    }
  return nmatches;
}

int SearchByProjection(orbx_matcher* m, const FrameView& F, unsigned char* hasMapPoint, const float* mpUV,
                       const float* mpRadius, const int* mpLevel, const float* mpUR, const unsigned char* mpDesc,
                       int nMP, const float* scaleFactors, float nnratio, int* mpMatch, const unsigned char* mpHasObs) {
  std::vector<float> qxy(mpUV, mpUV + 2 * (size_t)nMP), qr(nMP);
  std::vector<int> qmin(nMP), qmax(nMP);
  for (int i = 0; i < nMP; i++) {
    qr[i] = mpRadius[i] * scaleFactors[mpLevel[i]];  // :93
    qmin[i] = mpLevel[i] - 1;
    qmax[i] = mpLevel[i];
  }
  Candidates c;
  WindowSearch(m, F, qxy, qr, qmin, qmax, mpDesc, nMP, &c);
  int nmatches = 0;
  for (int iMP = 0; iMP < nMP; iMP++) {
    mpMatch[iMP] = -1;
    if (c.off[iMP] == c.off[iMP + 1]) continue;
    const int nPredictedLevel = mpLevel[iMP];
    const float r = mpRadius[iMP];
    int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
    for (int k = c.off[iMP]; k < c.off[iMP + 1]; k++) {
      const int idx = c.idx[k];
      if (hasMapPoint[idx]) continue;  // :118-121
      if (F.uRight && F.uRight[idx] > 0) {
        const float er = fabsf(mpUR[iMP] - F.uRight[idx]);
        if (er > r * scaleFactors[nPredictedLevel]) continue;
      }
      const int dist = c.dist[k];
      if (dist < bestDist) {
        bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = F.keysUn[idx].octave; bestIdx = idx;
      } else if (dist < bestDist2) {
        bestLevel2 = F.keysUn[idx].octave; bestDist2 = dist;
      }
    }
    if (bestDist <= TH_HIGH) {
      if (bestLevel == bestLevel2 && bestDist > nnratio * bestDist2) continue;
      mpMatch[iMP] = bestIdx;
      hasMapPoint[bestIdx] = mpHasObs ? (mpHasObs[iMP] != 0) : 1;  // F.mvpMapPoints[bestIdx]=pMP (:157), tested by :118-121
      nmatches++;
    }
  }
  return nmatches;
}

int SearchByProjection(orbx_matcher* m, const FrameView& Cur, unsigned char* hasMapPoint, const float* qUV,
                       const float* qRadius, const int* qMinLevel, const int* qMaxLevel, const float* qUR,
                       const float* qAngle, const unsigned char* qDesc, int nQ, int thDist, bool checkOrientation,
                       int* curMatch, const unsigned char* qHasObs, std::vector<int>* cleared) {
  std::vector<float> qxy(qUV, qUV + 2 * (size_t)nQ), qr(qRadius, qRadius + nQ);
  std::vector<int> qmin(qMinLevel, qMinLevel + nQ), qmax(qMaxLevel, qMaxLevel + nQ);
  Candidates c;
  WindowSearch(m, Cur, qxy, qr, qmin, qmax, qDesc, nQ, &c);
  for (int i = 0; i < Cur.N; i++) curMatch[i] = -1;
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (int i = 0; i < nQ; i++) {
    if (c.off[i] == c.off[i + 1]) continue;
    const float radius = qRadius[i];
    int bestDist = 256, bestIdx2 = -1;
    for (int k = c.off[i]; k < c.off[i + 1]; k++) {
      const int i2 = c.idx[k];
      if (hasMapPoint[i2]) continue;  // :1648-1650
      if (Cur.uRight && Cur.uRight[i2] > 0) {
        const float er = fabsf(qUR[i] - Cur.uRight[i2]);
        if (er > radius) continue;
      }
      const int dist = c.dist[k];
      if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
    }
    if (bestDist <= thDist) {
      curMatch[bestIdx2] = i;
      hasMapPoint[bestIdx2] = qHasObs ? (qHasObs[i] != 0) : 1;  // CurrentFrame.mvpMapPoints[bestIdx2]=pMP (:1675)
      nmatches++;
      if (checkOrientation) rotHist[RotBin(qAngle[i], Cur.keysUn[bestIdx2].angle)].push_back(bestIdx2);
    }
  }
  if (checkOrientation) {
    int ind1 = -1, ind2 = -1, ind3 = -1;
    ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
    for (int i = 0; i < HISTO_LENGTH; i++)
      if (i != ind1 && i != ind2 && i != ind3)
        for (size_t j = 0; j < rotHist[i].size(); j++) {
          curMatch[rotHist[i][j]] = -1;
          hasMapPoint[rotHist[i][j]] = 0;  // mvpMapPoints[...] = NULL (:1711)
          if (cleared) cleared->push_back(rotHist[i][j]);
          nmatches--;
        }
  }
  return nmatches;
}

// ---- bag-of-words restricted searches ------------------------------------------------------------------------
namespace {

// The vocabulary nodes both feature vectors hold, in ascending order (the while / lower_bound walk of :232-340), and the
// distance matrix of every such node from the device.
struct NodeBlocks {
  std::vector<int> aOff, aIdx, bOff, bIdx;
  std::vector<long long> dOff;
  std::vector<unsigned short> dist;
  int n() const { return (int)aOff.size() - 1; }
};

void CommonNodeDistances(orbx_matcher* m, const BowView& A, const BowView& B, NodeBlocks* nb) {
  nb->aOff.assign(1, 0);
  nb->bOff.assign(1, 0);
  nb->aIdx.clear();
  nb->bIdx.clear();
  const FeatVecView &fa = A.featVec, &fb = B.featVec;
  int i = 0, j = 0;
  while (i < fa.n && j < fb.n) {
    if (fa.nodes[i] == fb.nodes[j]) {
      nb->aIdx.insert(nb->aIdx.end(), fa.idx + fa.off[i], fa.idx + fa.off[i + 1]);
      nb->bIdx.insert(nb->bIdx.end(), fb.idx + fb.off[j], fb.idx + fb.off[j + 1]);
      nb->aOff.push_back((int)nb->aIdx.size());
      nb->bOff.push_back((int)nb->bIdx.size());
      i++; j++;
    } else if (fa.nodes[i] < fb.nodes[j]) {
      i = (int)(std::lower_bound(fa.nodes, fa.nodes + fa.n, fb.nodes[j]) - fa.nodes);
    } else {
      j = (int)(std::lower_bound(fb.nodes, fb.nodes + fb.n, fa.nodes[i]) - fb.nodes);
    }
  }
  const int K = nb->n();
  nb->dOff.assign(K + 1, 0);
  long long total = 0;
  for (int k = 0; k < K; k++) total += (long long)(nb->aOff[k + 1] - nb->aOff[k]) * (nb->bOff[k + 1] - nb->bOff[k]);
  nb->dist.resize((size_t)std::max<long long>(total, 1));
  int64_t need = 0;
  static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
  const int rc = orbx_match_blocks(m, A.descriptors, A.N, B.descriptors, B.N, &nb->aOff[0], nb->aIdx.empty() ? nullptr : &nb->aIdx[0],
                                   &nb->bOff[0], nb->bIdx.empty() ? nullptr : &nb->bIdx[0], K,
                                   reinterpret_cast<int64_t*>(&nb->dOff[0]), &nb->dist[0], (int64_t)nb->dist.size(), &need);
  if (rc != ORBX_OK) Fail("orbx_match_blocks");
}

// removes the matches outside the three dominant rotation bins (:342-362, :744-768, :947-967); returns how many
template <class Clear>
int PruneRotation(const std::vector<int>* rotHist, Clear clear) {
  int ind1 = -1, ind2 = -1, ind3 = -1, removed = 0;
  ComputeThreeMaxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
  for (int i = 0; i < HISTO_LENGTH; i++) {
    if (i == ind1 || i == ind2 || i == ind3) continue;
    for (size_t j = 0; j < rotHist[i].size(); j++) { clear(rotHist[i][j]); removed++; }
  }
  return removed;
}

}  // namespace

int SearchByBoW(orbx_matcher* m, const BowView& KF, const BowView& F, float nnratio, bool checkOrientation, int* matchF) {
  for (int i = 0; i < F.N; i++) matchF[i] = -1;
  NodeBlocks nb;
  CommonNodeDistances(m, KF, F, &nb);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (int k = 0; k < nb.n(); k++) {
    const int nF = nb.bOff[k + 1] - nb.bOff[k];
    const int* idxF = nF ? &nb.bIdx[nb.bOff[k]] : nullptr;
    for (int e = nb.aOff[k]; e < nb.aOff[k + 1]; e++) {
      const int realIdxKF = nb.aIdx[e];
      if (!KF.flag[realIdxKF]) continue;                               // :247-253
      const unsigned short* d = &nb.dist[nb.dOff[k] + (long long)(e - nb.aOff[k]) * nF];
      int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
      for (int iF = 0; iF < nF; iF++) {
        const int realIdxF = idxF[iF];
        if (matchF[realIdxF] >= 0) continue;                           // :265
        const int dist = d[iF];
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = realIdxF; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 <= TH_LOW && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {   // :285-287
        matchF[bestIdxF] = realIdxKF;
        if (checkOrientation) rotHist[RotBin(KF.keysUn[realIdxKF].angle, F.keysUn[bestIdxF].angle)].push_back(bestIdxF);
        nmatches++;
      }
    }
  }
  if (checkOrientation) nmatches -= PruneRotation(rotHist, [&](int iF) { matchF[iF] = -1; });
  return nmatches;
}

int SearchByBoWKeyFrames(orbx_matcher* m, const BowView& KF1, const BowView& KF2, float nnratio, bool checkOrientation,
                         int* matches12) {
  for (int i = 0; i < KF1.N; i++) matches12[i] = -1;
  NodeBlocks nb;
  CommonNodeDistances(m, KF1, KF2, &nb);
  std::vector<unsigned char> vbMatched2((size_t)std::max(KF2.N, 1), 0);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (int k = 0; k < nb.n(); k++) {
    const int n2 = nb.bOff[k + 1] - nb.bOff[k];
    const int* list2 = n2 ? &nb.bIdx[nb.bOff[k]] : nullptr;
    for (int e = nb.aOff[k]; e < nb.aOff[k + 1]; e++) {
      const int idx1 = nb.aIdx[e];
      if (!KF1.flag[idx1]) continue;                                   // :667-673
      const unsigned short* d = &nb.dist[nb.dOff[k] + (long long)(e - nb.aOff[k]) * n2];
      int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
      for (int i2 = 0; i2 < n2; i2++) {
        const int idx2 = list2[i2];
        if (vbMatched2[idx2] || !KF2.flag[idx2]) continue;             // :682-688
        const int dist = d[i2];
        if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
        else if (dist < bestDist2) bestDist2 = dist;
      }
      if (bestDist1 < TH_LOW && static_cast<float>(bestDist1) < nnratio * static_cast<float>(bestDist2)) {   // :707-709
        matches12[idx1] = bestIdx2;
        vbMatched2[bestIdx2] = 1;
        if (checkOrientation) rotHist[RotBin(KF1.keysUn[idx1].angle, KF2.keysUn[bestIdx2].angle)].push_back(idx1);
        nmatches++;
      }
    }
  }
  if (checkOrientation) nmatches -= PruneRotation(rotHist, [&](int i1) { matches12[i1] = -1; });
  return nmatches;
}

namespace {
// R/src/ORBmatcher.cpp:175-203. The reference is built with -O3 and GCC's default -ffp-contract=fast; the fused
// operations below are the ones its object code holds, written out so that this file gives the
// same bits under any contraction setting.
bool CheckDistEpipolarLine(const orbx_keypoint& kp1, const orbx_keypoint& kp2, const float* F12, const float* levelSigma2_2) {
  const float a = fmaf(kp1.x, F12[0], kp1.y * F12[3]) + F12[6];
  const float b = fmaf(kp1.x, F12[1], kp1.y * F12[4]) + F12[7];
  const float c = fmaf(kp1.y, F12[5], kp1.x * F12[2]) + F12[8];
  const float num = c + fmaf(b, kp2.y, a * kp2.x);
  const float den = fmaf(a, a, b * b);
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * levelSigma2_2[kp2.octave];
}
}  // namespace

int SearchForTriangulation(orbx_matcher* m, const BowView& KF1, const BowView& KF2, const float* F12, float ex, float ey,
                           const float* scaleFactors2, const float* levelSigma2_2, bool onlyStereo, bool checkOrientation,
                           std::vector<std::pair<int, int> >* vMatchedPairs) {
  NodeBlocks nb;
  CommonNodeDistances(m, KF1, KF2, &nb);
  std::vector<unsigned char> vbMatched2((size_t)std::max(KF2.N, 1), 0);
  std::vector<int> vMatches12((size_t)KF1.N, -1);
  std::vector<int> rotHist[HISTO_LENGTH];
  int nmatches = 0;
  for (int k = 0; k < nb.n(); k++) {
    const int n2 = nb.bOff[k + 1] - nb.bOff[k];
    const int* list2 = n2 ? &nb.bIdx[nb.bOff[k]] : nullptr;
    for (int e = nb.aOff[k]; e < nb.aOff[k + 1]; e++) {
      const int idx1 = nb.aIdx[e];
      if (KF1.flag[idx1]) continue;                                    // :833-836: already has a map point
      const bool bStereo1 = KF1.uRight && KF1.uRight[idx1] >= 0;
      if (onlyStereo && !bStereo1) continue;
      const orbx_keypoint& kp1 = KF1.keysUn[idx1];
      const unsigned short* d = &nb.dist[nb.dOff[k] + (long long)(e - nb.aOff[k]) * n2];
      int bestDist = TH_LOW, bestIdx2 = -1;
      for (int i2 = 0; i2 < n2; i2++) {
        const int idx2 = list2[i2];
        if (vbMatched2[idx2] || KF2.flag[idx2]) continue;              // :858-859
        const bool bStereo2 = KF2.uRight && KF2.uRight[idx2] >= 0;
        if (onlyStereo && !bStereo2) continue;
        const int dist = d[i2];
        if (dist > TH_LOW || dist > bestDist) continue;                // :872
        const orbx_keypoint& kp2 = KF2.keysUn[idx2];
        if (!bStereo1 && !bStereo2) {
          const float distex = ex - kp2.x, distey = ey - kp2.y;
          if (fmaf(distex, distex, distey * distey) < 100 * scaleFactors2[kp2.octave]) continue;   // :881-883
        }
        if (CheckDistEpipolarLine(kp1, kp2, F12, levelSigma2_2)) { bestIdx2 = idx2; bestDist = dist; }
      }
      if (bestIdx2 >= 0) {
        vMatches12[idx1] = bestIdx2;
        vbMatched2[bestIdx2] = 1;
        nmatches++;
        if (checkOrientation) rotHist[RotBin(kp1.angle, KF2.keysUn[bestIdx2].angle)].push_back(idx1);
      }
    }
  }
  if (checkOrientation) nmatches -= PruneRotation(rotHist, [&](int i1) { vMatches12[i1] = -1; });
  vMatchedPairs->clear();
  vMatchedPairs->reserve(nmatches);
  for (int i = 0; i < KF1.N; i++)
    if (vMatches12[i] >= 0) vMatchedPairs->push_back(std::make_pair(i, vMatches12[i]));
  return nmatches;
}

void SearchBestInWindow(orbx_matcher* m, const FrameView& KF, const float* invLevelSigma2, const float* qUV, const float* qUR,
                        const float* qRadius, const int* qLevel, const unsigned char* qDesc, int nQ, unsigned char* taken,
                        int th, int* bestIdx, int* bestDist) {
  std::vector<float> qxy(qUV, qUV + 2 * (size_t)nQ), qr(qRadius, qRadius + nQ);
  std::vector<int> qmin((size_t)nQ), qmax((size_t)nQ);
  for (int i = 0; i < nQ; i++) { qmin[i] = qLevel[i] - 1; qmax[i] = qLevel[i]; }   // the octave test of the loops, done on the device
  Candidates c;
  WindowSearch(m, KF, qxy, qr, qmin, qmax, qDesc, nQ, &c);
  for (int i = 0; i < nQ; i++) {
    const float u = qUV[2 * i], v = qUV[2 * i + 1];
    int best = 256, idxBest = -1;
    for (int k = c.off[i]; k < c.off[i + 1]; k++) {
      const int idx = c.idx[k];
      if (taken && taken[idx]) continue;
      if (invLevelSigma2) {
        // fused like the reference's object code (-O3, -ffp-contract=fast), spelled out so that any build agrees
        const orbx_keypoint& kp = KF.keysUn[idx];
        const float ex = u - kp.x, ey = v - kp.y;
        const float e2m = fmaf(ex, ex, ey * ey);
        if (KF.uRight && KF.uRight[idx] >= 0) {
          const float er = qUR[i] - KF.uRight[idx];
          const float e2 = fmaf(er, er, e2m);
          if (e2 * invLevelSigma2[kp.octave] > 7.8) continue;
        } else {
          if (e2m * invLevelSigma2[kp.octave] > 5.99) continue;
        }
      }
      if (c.dist[k] < best) { best = c.dist[k]; idxBest = idx; }
    }
    bestIdx[i] = idxBest;
    bestDist[i] = best;
    if (taken && best <= th) taken[idxBest] = 1;
  }
}

}  // namespace orbx_host
