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
int SearchByProjection(orbx_matcher* m, const FrameView& Cur, unsigned char* hasMapPoint, const float* qUV,
                       const float* qRadius, const int* qMinLevel, const int* qMaxLevel, const float* qUR,
                       const float* qAngle, const unsigned char* qDesc, int nQ, int thDist, bool checkOrientation,
                       int* curMatch, const unsigned char* qHasObs = nullptr);

}  // namespace orbx_host

#endif
