// Host half of ORBmatcher's windowed searches over liborbx.so.
//
// The device (orbx_match_window) returns, per query, the candidates Frame::GetFeaturesInArea would return
// -- in the reference's order -- with their Hamming distances. What remains is the ORDER-DEPENDENT part
// of the reference's loops (a keypoint matched by an earlier query changes what a later query may take,
// best / second-best bookkeeping, rotation-histogram consistency), which is kept here with the reference's
// control flow. The functions take plain arrays, so a replacement ORBmatcher.cpp calls them after flattening
// its Frame / MapPoint objects.
#ifndef ORBX_SEARCH_H_
#define ORBX_SEARCH_H_

#include <utility>
#include <vector>

#include "orbx.h"

namespace orbx_host {

const int TH_LOW = 50, TH_HIGH = 100, HISTO_LENGTH = 30;   // R/src/ORBmatcher.cpp:37-39

// A Frame as the matcher sees it: mvKeysUn, mDescriptors, the image bounds that define its grid
// (mnMinX..mnMaxY), and mvuRight (NULL for monocular frames).
struct FrameView {
  const orbx_keypoint* keysUn;
  const unsigned char* descriptors;
  int N;
  float minX, maxX, minY, maxY;
  const float* uRight;
};

// R/src/ORBmatcher.cpp:1854-1895
void ComputeThreeMaxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3);

// R/src/ORBmatcher.cpp:499-617. vbPrevMatched: F1.N (x,y) pairs, updated like the reference's :612-614;
// vnMatches12: F1.N entries. Returns nmatches.
int SearchForInitialization(orbx_matcher* m, const FrameView& F1, const FrameView& F2, float* vbPrevMatched,
                            int* vnMatches12, int windowSize, float nnratio, bool checkOrientation);

// R/src/ORBmatcher.cpp:63-163 for the map points that are in view: projection (u,v), r =
// RadiusByViewingCos(viewCos)*th, predicted level, projected right coordinate, representative descriptor.
// hasMapPoint[i] != 0 <=> F.mvpMapPoints[i] with Observations()>0; it is updated as points are assigned.
// mpMatch[iMP] = keypoint index or -1. Returns nmatches.
int SearchByProjection(orbx_matcher* m, const FrameView& F, unsigned char* hasMapPoint, const float* mpUV,
                       const float* mpRadius, const int* mpLevel, const float* mpUR, const unsigned char* mpDesc,
                       int nMP, const float* scaleFactors, float nnratio, int* mpMatch,
                       const unsigned char* mpHasObs = nullptr);

// R/src/ORBmatcher.cpp:1564-1718 (last frame -> current frame) and :1723-1851 (keyframe -> frame) after
// their projection prologues: per query (u,v), radius, level window, ur = u - mbf*invzc, the source
// keypoint's angle and descriptor. curMatch[i2] = query index or -1. Returns nmatches.
// cleared (optional): the keypoints whose mvpMapPoints entry the rotation check resets to NULL (:1704-1714), which a
// caller holding pointers cannot tell from untouched entries by curMatch alone.
int SearchByProjection(orbx_matcher* m, const FrameView& Cur, unsigned char* hasMapPoint, const float* qUV,
                       const float* qRadius, const int* qMinLevel, const int* qMaxLevel, const float* qUR,
                       const float* qAngle, const unsigned char* qDesc, int nQ, int thDist, bool checkOrientation,
                       int* curMatch, const unsigned char* qHasObs = nullptr, std::vector<int>* cleared = nullptr);

// ---- bag-of-words restricted searches ---------------------------------------------------
// A DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) flattened in map order: node ids ascending,
// CSR offsets, feature indices. A replacement ORBmatcher.cpp builds it with one pass over mFeatVec.
struct FeatVecView {
  const int* nodes;
  const int* off;
  const int* idx;
  int n;
};

// A KeyFrame (or Frame) as these searches see it: mvKeysUn, mDescriptors, mFeatVec, mvuRight (NULL: monocular) and,
// per keypoint, a flag whose meaning the function states (it flattens GetMapPointMatches() / GetMapPoint(i)).
struct BowView {
  const orbx_keypoint* keysUn;
  const unsigned char* descriptors;
  int N;
  FeatVecView featVec;
  const unsigned char* flag;
  const float* uRight;
};

// R/src/ORBmatcher.cpp:220-365. KF.flag[i] != 0 <=> vpMapPointsKF[i] && !isBad(). F.flag is not read.
// matchF[iF] (F.N entries) = key-frame feature whose map point goes to vpMapPointMatches[iF], or -1. Returns nmatches.
int SearchByBoW(orbx_matcher* m, const BowView& KF, const BowView& F, float nnratio, bool checkOrientation, int* matchF);

// R/src/ORBmatcher.cpp:632-772. flag[i] != 0 <=> vpMapPoints[i] && !isBad() in both key frames.
// matches12[i1] (KF1.N entries) = feature of key frame 2 whose map point goes to vpMatches12[i1], or -1.
int SearchByBoWKeyFrames(orbx_matcher* m, const BowView& KF1, const BowView& KF2, float nnratio, bool checkOrientation,
                         int* matches12);

// R/src/ORBmatcher.cpp:785-983 after its prologue (:790-803, cv::Mat algebra, stays with the caller): (ex, ey) = the
// epipole of camera 1 in image 2; F12 = 3x3 row-major; flag[i] != 0 <=> GetMapPoint(i) != NULL; scaleFactors2 /
// levelSigma2_2 = pKF2->mvScaleFactors / mvLevelSigma2. vMatchedPairs = (idx1, idx2) in idx1 order. Returns nmatches.
int SearchForTriangulation(orbx_matcher* m, const BowView& KF1, const BowView& KF2, const float* F12, float ex, float ey,
                           const float* scaleFactors2, const float* levelSigma2_2, bool onlyStereo, bool checkOrientation,
                           std::vector<std::pair<int, int> >* vMatchedPairs);

// ---- windowed best match (Fuse x2, SearchByProjection(KeyFrame*, Scw), SearchBySim3) --------
// The loop those four functions share after their projection prologues (R/src/ORBmatcher.cpp:1046-1110, :1248-1270,
// :441-470, :1379-1406 / :1459-1486): candidates = KeyFrame::GetFeaturesInArea(u, v, radius) (R/src/KeyFrame.cpp:702-743,
// the Frame grid without a level argument) with octave in [level-1, level], best Hamming distance, first minimum wins.
//   invLevelSigma2 != NULL: Fuse's chi-square gates (:1070-1099) on (u - kpx, v - kpy [, ur - mvuRight]) -- 5.99 / 7.8;
//   taken != NULL: vpMatched of SearchByProjection(KeyFrame*, Scw, ...) (:453-454): skipped when set, set for the best
//   keypoint of a query that ends with bestDist <= th (:479-483).
// bestIdx[i] = keypoint or -1, bestDist[i] = its distance or 256. What the caller does with a hit (Replace /
// AddObservation / vpReplacePoint / the mutual check of SearchBySim3 :1494-1508) is pointer-graph work and stays there.
void SearchBestInWindow(orbx_matcher* m, const FrameView& KF, const float* invLevelSigma2, const float* qUV, const float* qUR,
                        const float* qRadius, const int* qLevel, const unsigned char* qDesc, int nQ, unsigned char* taken,
                        int th, int* bestIdx, int* bestDist);

}  // namespace orbx_host

#endif
