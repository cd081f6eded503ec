// Drop-in replacement of R/src/ORBmatcher.cpp over liborbx.so: the same class (R/include/ORBmatcher.h is compiled
// unchanged), the same member functions, the same results. Tracking, LocalMapping and LoopClosing keep calling
// ORBmatcher::SearchByProjection / SearchByBoW / Fuse ... on their Frame / KeyFrame / MapPoint objects.
//
// Every member has three parts:
//   gather  -- walk the pointer graph once, apply the reference's skip rules and geometric prologue (cv::Mat
//              algebra stays on the host), and flatten what is left into query arrays;
//   search  -- one call into orb-slam2-_b200/host/orbx_search.h, whose candidate lists and Hamming distances come
//              from the device (orbx_match_window / orbx_match_blocks);
//   apply   -- write the matches back into the caller's vectors and objects.
// The scalar float expressions of the prologues are written with the fused operations the reference's -O3 object
// code contains, spelled out with fmaf so that this file gives the same bits under any
// -ffp-contract setting.
//
// tests/test_gpu_dropin_matcher.py links this file with the reference's own Frame.cpp / KeyFrame.cpp / MapPoint.cpp
// (oracle/_ref/libdropin_front.so) and compares every member against the reference's ORBmatcher.cpp on the same
// objects.
#include "ORBmatcher.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <set>

#include "orbx.h"
#include "orbx_search.h"
#include "orbx_thread.h"

namespace ORB_SLAM2 {

const int ORBmatcher::TH_HIGH = 100;       // R/src/ORBmatcher.cpp:37-39
const int ORBmatcher::TH_LOW = 50;
const int ORBmatcher::HISTO_LENGTH = 30;

namespace {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte record of include/orbx.h");

orbx_matcher* Handle() { return orbx_host::ThreadMatcher(); }   // one device handle per host thread (orbx_thread.h)

const orbx_keypoint* Keys(const std::vector<cv::KeyPoint>& v) {
  return v.empty() ? nullptr : reinterpret_cast<const orbx_keypoint*>(&v[0]);
}

// mDescriptors is one continuous N x 32 CV_8U matrix (R/src/ORBextractor.cpp:1148)
const unsigned char* Rows(const cv::Mat& d) { return d.empty() ? nullptr : d.ptr<unsigned char>(0); }

orbx_host::FrameView ViewOf(const Frame& F) {
  orbx_host::FrameView v;
  v.keysUn = Keys(F.mvKeysUn); v.descriptors = Rows(F.mDescriptors); v.N = F.N;
  v.minX = Frame::mnMinX; v.maxX = Frame::mnMaxX; v.minY = Frame::mnMinY; v.maxY = Frame::mnMaxY;
  v.uRight = F.mvuRight.empty() ? nullptr : &F.mvuRight[0];
  return v;
}

orbx_host::FrameView ViewOf(const KeyFrame* K) {
  orbx_host::FrameView v;
  v.keysUn = Keys(K->mvKeysUn); v.descriptors = Rows(K->mDescriptors); v.N = K->N;
  v.minX = (float)K->mnMinX; v.maxX = (float)K->mnMaxX; v.minY = (float)K->mnMinY; v.maxY = (float)K->mnMaxY;
  v.uRight = K->mvuRight.empty() ? nullptr : &K->mvuRight[0];
  return v;
}

// DBoW2::FeatureVector (std::map<NodeId, std::vector<unsigned>>) in map order
struct FlatFeatVec {
  std::vector<int> nodes, off, idx;
  explicit FlatFeatVec(const DBoW2::FeatureVector& fv) {
    off.push_back(0);
    for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it) {
      nodes.push_back((int)it->first);
      idx.insert(idx.end(), it->second.begin(), it->second.end());
      off.push_back((int)idx.size());
    }
  }
  orbx_host::FeatVecView view() const {
    orbx_host::FeatVecView v;
    v.nodes = nodes.empty() ? nullptr : &nodes[0]; v.off = &off[0]; v.idx = idx.empty() ? nullptr : &idx[0];
    v.n = (int)nodes.size();
    return v;
  }
};

orbx_host::BowView BowOf(const std::vector<cv::KeyPoint>& keysUn, const cv::Mat& desc, int N, const FlatFeatVec& fv,
                         const std::vector<unsigned char>* flag, const std::vector<float>& uRight) {
  orbx_host::BowView v;
  v.keysUn = Keys(keysUn); v.descriptors = Rows(desc); v.N = N; v.featVec = fv.view();
  v.flag = flag && !flag->empty() ? &(*flag)[0] : nullptr;
  v.uRight = uRight.empty() ? nullptr : &uRight[0];
  return v;
}

void AppendRow(std::vector<unsigned char>* rows, const cv::Mat& d) {
  const unsigned char* p = d.ptr<unsigned char>(0);
  rows->insert(rows->end(), p, p + 32);
}

// the queries of the windowed best-match loop (orbx_host::SearchBestInWindow)
struct WindowQueries {
  std::vector<float> uv, ur, radius;
  std::vector<int> level, owner;   // owner: index into the caller's point vector
  std::vector<unsigned char> desc;
  int n() const { return (int)radius.size(); }
  void add(int who, float u, float v, float r, float rad, int lvl, const cv::Mat& d) {
    owner.push_back(who); uv.push_back(u); uv.push_back(v); ur.push_back(r); radius.push_back(rad); level.push_back(lvl);
    AppendRow(&desc, d);
  }
};

}  // namespace

ORBmatcher::ORBmatcher(float nnratio, bool checkOri) : mfNNratio(nnratio), mbCheckOrientation(checkOri) {}

// R/src/ORBmatcher.cpp:1901-1917. Single pairs stay on the host: eight popcounts are cheaper than a launch.
int ORBmatcher::DescriptorDistance(const cv::Mat& a, const cv::Mat& b) {
  const unsigned char *pa = a.ptr<unsigned char>(0), *pb = b.ptr<unsigned char>(0);
  int dist = 0;
  for (int i = 0; i < 4; i++) {
    unsigned long long x, y;
    std::memcpy(&x, pa + 8 * i, 8);
    std::memcpy(&y, pb + 8 * i, 8);
    dist += __builtin_popcountll(x ^ y);
  }
  return dist;
}

float ORBmatcher::RadiusByViewingCos(const float& viewCos) { return viewCos > 0.998 ? 2.5f : 4.0f; }   // :166-172

void ORBmatcher::ComputeThreeMaxima(std::vector<int>* histo, const int L, int& ind1, int& ind2, int& ind3) {
  orbx_host::ComputeThreeMaxima(histo, L, ind1, ind2, ind3);
}

// :175-203 (kept for callers inside the class hierarchy; SearchForTriangulation evaluates it inside orbx_search.cc)
bool ORBmatcher::CheckDistEpipolarLine(const cv::KeyPoint& kp1, const cv::KeyPoint& kp2, const cv::Mat& F12, const KeyFrame* pKF2) {
  const float a = fmaf(kp1.pt.x, F12.at<float>(0, 0), kp1.pt.y * F12.at<float>(1, 0)) + F12.at<float>(2, 0);
  const float b = fmaf(kp1.pt.x, F12.at<float>(0, 1), kp1.pt.y * F12.at<float>(1, 1)) + F12.at<float>(2, 1);
  const float c = fmaf(kp1.pt.y, F12.at<float>(1, 2), kp1.pt.x * F12.at<float>(0, 2)) + F12.at<float>(2, 2);
  const float num = c + fmaf(b, kp2.pt.y, a * kp2.pt.x);
  const float den = fmaf(a, a, b * b);
  if (den == 0) return false;
  const float dsqr = num * num / den;
  return dsqr < 3.84 * pKF2->mvLevelSigma2[kp2.octave];
}

// ---- Tracking::SearchLocalPoints -> SearchByProjection(F, vpMapPoints, th) (R/src/ORBmatcher.cpp:63-163) ----------
int ORBmatcher::SearchByProjection(Frame& F, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  const bool scaled = th != 1.0;
  // gather: the points Frame::isInFrustum marked (mbTrackInView) carry their projection and predicted level
  std::vector<int> owner, level;
  std::vector<float> uv, radius, ur;
  std::vector<unsigned char> desc, hasObs;
  for (size_t k = 0; k < vpMapPoints.size(); k++) {
    MapPoint* p = vpMapPoints[k];
    if (!p->mbTrackInView || p->isBad()) continue;
    float r = RadiusByViewingCos(p->mTrackViewCos);
    if (scaled) r *= th;
    owner.push_back((int)k); level.push_back(p->mnTrackScaleLevel); radius.push_back(r);
    uv.push_back(p->mTrackProjX); uv.push_back(p->mTrackProjY); ur.push_back(p->mTrackProjXR);
    AppendRow(&desc, p->GetDescriptor());
    hasObs.push_back(p->Observations() > 0);
  }
  const int nQ = (int)owner.size();
  if (nQ == 0 || F.N == 0) return 0;
  std::vector<unsigned char> occupied(F.N);   // :118-121
  for (int i = 0; i < F.N; i++) occupied[i] = F.mvpMapPoints[i] && F.mvpMapPoints[i]->Observations() > 0;
  // search
  std::vector<int> match(nQ);
  const int nmatches = orbx_host::SearchByProjection(Handle(), ViewOf(F), &occupied[0], &uv[0], &radius[0], &level[0], &ur[0],
                                                     &desc[0], nQ, &F.mvScaleFactors[0], mfNNratio, &match[0], &hasObs[0]);
  // apply, in query order (a later point may take the keypoint of an earlier one that has no observation yet)
  for (int q = 0; q < nQ; q++)
    if (match[q] >= 0) F.mvpMapPoints[match[q]] = vpMapPoints[owner[q]];
  return nmatches;
}

namespace {

// what the two frame-to-frame projections below share after their gather
int ProjectAndAssign(Frame& Cur, const orbx_host::FrameView& view, std::vector<unsigned char>& occupied,
                     const std::vector<MapPoint*>& qPoint, const std::vector<float>& uv, const std::vector<float>& radius,
                     const std::vector<int>& minLevel, const std::vector<int>& maxLevel, const std::vector<float>& ur,
                     const std::vector<float>& angle, const std::vector<unsigned char>& desc,
                     const std::vector<unsigned char>* hasObs, int thDist, bool checkOrientation) {
  const int nQ = (int)qPoint.size();
  if (nQ == 0 || Cur.N == 0) return 0;
  std::vector<int> curMatch(Cur.N), cleared;
  const int nmatches = orbx_host::SearchByProjection(Handle(), view, &occupied[0], &uv[0], &radius[0], &minLevel[0], &maxLevel[0],
                                                     &ur[0], &angle[0], &desc[0], nQ, thDist, checkOrientation, &curMatch[0],
                                                     hasObs ? &(*hasObs)[0] : nullptr, &cleared);
  for (int i2 = 0; i2 < Cur.N; i2++)
    if (curMatch[i2] >= 0) Cur.mvpMapPoints[i2] = qPoint[curMatch[i2]];
  for (size_t k = 0; k < cleared.size(); k++) Cur.mvpMapPoints[cleared[k]] = static_cast<MapPoint*>(NULL);
  return nmatches;
}

}  // namespace

// ---- Tracking::TrackWithMotionModel -> SearchByProjection(CurrentFrame, LastFrame, th, bMono) (:1564-1718) --------
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, const Frame& LastFrame, const float th, const bool bMono) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;   // the current camera centre seen from the last frame
  const bool forward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  const bool backward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;

  std::vector<MapPoint*> qPoint;
  std::vector<float> uv, radius, ur, angle;
  std::vector<int> minLevel, maxLevel;
  std::vector<unsigned char> desc, hasObs;
  for (int i = 0; i < LastFrame.N; i++) {
    MapPoint* p = LastFrame.mvpMapPoints[i];
    if (!p || LastFrame.mvbOutlier[i]) continue;
    const cv::Mat x3Dc = Rcw * p->GetWorldPos() + tcw;
    const float invzc = 1.0 / x3Dc.at<float>(2);
    if (invzc < 0) continue;
    const float u = fmaf(CurrentFrame.fx * x3Dc.at<float>(0), invzc, CurrentFrame.cx);
    const float v = fmaf(CurrentFrame.fy * x3Dc.at<float>(1), invzc, CurrentFrame.cy);
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX || v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    const int octave = LastFrame.mvKeys[i].octave;
    qPoint.push_back(p);
    uv.push_back(u); uv.push_back(v);
    radius.push_back(th * CurrentFrame.mvScaleFactors[octave]);
    minLevel.push_back(forward ? octave : backward ? 0 : octave - 1);          // :1629-1634
    maxLevel.push_back(forward ? -1 : backward ? octave : octave + 1);         // -1: no upper bound (R/src/Frame.cpp:408)
    ur.push_back(fmaf(-CurrentFrame.mbf, invzc, u));
    angle.push_back(LastFrame.mvKeysUn[i].angle);
    AppendRow(&desc, p->GetDescriptor());
    hasObs.push_back(p->Observations() > 0);
  }
  std::vector<unsigned char> occupied(CurrentFrame.N);   // :1649-1651
  for (int i = 0; i < CurrentFrame.N; i++)
    occupied[i] = CurrentFrame.mvpMapPoints[i] && CurrentFrame.mvpMapPoints[i]->Observations() > 0;
  return ProjectAndAssign(CurrentFrame, ViewOf(CurrentFrame), occupied, qPoint, uv, radius, minLevel, maxLevel, ur, angle, desc,
                          &hasObs, TH_HIGH, mbCheckOrientation);
}

// ---- Tracking::Relocalization -> SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1723-1851) ----
int ORBmatcher::SearchByProjection(Frame& CurrentFrame, KeyFrame* pKF, const std::set<MapPoint*>& sAlreadyFound, const float th,
                                   const int ORBdist) {
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat Ow = -Rcw.t() * tcw;
  const std::vector<MapPoint*> vpMPs = pKF->GetMapPointMatches();

  std::vector<MapPoint*> qPoint;
  std::vector<float> uv, radius, ur, angle;
  std::vector<int> minLevel, maxLevel;
  std::vector<unsigned char> desc;
  for (size_t i = 0; i < vpMPs.size(); i++) {
    MapPoint* p = vpMPs[i];
    if (!p || p->isBad() || sAlreadyFound.count(p)) continue;
    const cv::Mat x3Dw = p->GetWorldPos();
    const cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float invzc = 1.0 / x3Dc.at<float>(2);
    const float u = fmaf(CurrentFrame.fx * x3Dc.at<float>(0), invzc, CurrentFrame.cx);
    const float v = fmaf(CurrentFrame.fy * x3Dc.at<float>(1), invzc, CurrentFrame.cy);
    if (u < CurrentFrame.mnMinX || u > CurrentFrame.mnMaxX || v < CurrentFrame.mnMinY || v > CurrentFrame.mnMaxY) continue;
    const float dist3D = cv::norm(x3Dw - Ow);
    if (dist3D < p->GetMinDistanceInvariance() || dist3D > p->GetMaxDistanceInvariance()) continue;
    const int lvl = p->PredictScale(dist3D, &CurrentFrame);
    qPoint.push_back(p);
    uv.push_back(u); uv.push_back(v);
    radius.push_back(th * CurrentFrame.mvScaleFactors[lvl]);
    minLevel.push_back(lvl - 1); maxLevel.push_back(lvl + 1);
    ur.push_back(0.f);   // this search has no right-image test
    angle.push_back(pKF->mvKeysUn[i].angle);
    AppendRow(&desc, p->GetDescriptor());
  }
  std::vector<unsigned char> occupied(CurrentFrame.N);   // :1792: any pointer counts here
  for (int i = 0; i < CurrentFrame.N; i++) occupied[i] = CurrentFrame.mvpMapPoints[i] != NULL;
  orbx_host::FrameView view = ViewOf(CurrentFrame);
  view.uRight = nullptr;
  return ProjectAndAssign(CurrentFrame, view, occupied, qPoint, uv, radius, minLevel, maxLevel, ur, angle, desc, nullptr, ORBdist,
                          mbCheckOrientation);
}

// ---- Tracking::MonocularInitialization -> SearchForInitialization (:499-617) -----------------------------------------
int ORBmatcher::SearchForInitialization(Frame& F1, Frame& F2, std::vector<cv::Point2f>& vbPrevMatched, std::vector<int>& vnMatches12,
                                        int windowSize) {
  vnMatches12.assign(F1.mvKeysUn.size(), -1);
  if (F1.N == 0 || F2.N == 0) return 0;
  static_assert(sizeof(cv::Point2f) == 2 * sizeof(float), "cv::Point2f must be two packed floats");
  return orbx_host::SearchForInitialization(Handle(), ViewOf(F1), ViewOf(F2), reinterpret_cast<float*>(&vbPrevMatched[0]),
                                            &vnMatches12[0], windowSize, mfNNratio, mbCheckOrientation);
}

namespace {

// flag[i] = the key frame holds a usable map point on keypoint i (pointer set and !isBad())
std::vector<unsigned char> GoodPointFlags(const std::vector<MapPoint*>& pts) {
  std::vector<unsigned char> f(pts.size());
  for (size_t i = 0; i < pts.size(); i++) f[i] = pts[i] && !pts[i]->isBad();
  return f;
}

}  // namespace

// ---- Tracking::TrackReferenceKeyFrame / Relocalization -> SearchByBoW(pKF, F, vpMapPointMatches) (:220-365) --------
int ORBmatcher::SearchByBoW(KeyFrame* pKF, Frame& F, std::vector<MapPoint*>& vpMapPointMatches) {
  const std::vector<MapPoint*> pointsKF = pKF->GetMapPointMatches();
  vpMapPointMatches.assign(F.N, static_cast<MapPoint*>(NULL));
  if (F.N == 0 || pKF->N == 0) return 0;
  const std::vector<unsigned char> good = GoodPointFlags(pointsKF);
  const FlatFeatVec fvKF(pKF->mFeatVec), fvF(F.mFeatVec);
  static const std::vector<float> none;
  std::vector<int> matchF(F.N);
  // the rotation bin is taken against F.mvKeys (:310), whose angles are those of mvKeysUn
  const int nmatches = orbx_host::SearchByBoW(Handle(), BowOf(pKF->mvKeysUn, pKF->mDescriptors, pKF->N, fvKF, &good, none),
                                              BowOf(F.mvKeys, F.mDescriptors, F.N, fvF, nullptr, none), mfNNratio,
                                              mbCheckOrientation, &matchF[0]);
  for (int i = 0; i < F.N; i++)
    if (matchF[i] >= 0) vpMapPointMatches[i] = pointsKF[matchF[i]];
  return nmatches;
}

// ---- LoopClosing::ComputeSim3 -> SearchByBoW(pKF1, pKF2, vpMatches12) (:632-772) ------------------------------------
int ORBmatcher::SearchByBoW(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12) {
  const std::vector<MapPoint*> points1 = pKF1->GetMapPointMatches(), points2 = pKF2->GetMapPointMatches();
  vpMatches12.assign(points1.size(), static_cast<MapPoint*>(NULL));
  if (pKF1->N == 0 || pKF2->N == 0) return 0;
  const std::vector<unsigned char> good1 = GoodPointFlags(points1), good2 = GoodPointFlags(points2);
  const FlatFeatVec fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
  static const std::vector<float> none;
  std::vector<int> match12(points1.size());
  const int nmatches = orbx_host::SearchByBoWKeyFrames(Handle(), BowOf(pKF1->mvKeysUn, pKF1->mDescriptors, pKF1->N, fv1, &good1, none),
                                                       BowOf(pKF2->mvKeysUn, pKF2->mDescriptors, pKF2->N, fv2, &good2, none),
                                                       mfNNratio, mbCheckOrientation, &match12[0]);
  for (size_t i = 0; i < points1.size(); i++)
    if (match12[i] >= 0) vpMatches12[i] = points2[match12[i]];
  return nmatches;
}

// ---- LocalMapping::CreateNewMapPoints -> SearchForTriangulation (:785-983) -------------------------------------------
int ORBmatcher::SearchForTriangulation(KeyFrame* pKF1, KeyFrame* pKF2, cv::Mat F12,
                                       std::vector<std::pair<size_t, size_t> >& vMatchedPairs, const bool bOnlyStereo) {
  // the epipole of camera 1 in image 2 (:790-803)
  const cv::Mat C2 = pKF2->GetRotation() * pKF1->GetCameraCenter() + pKF2->GetTranslation();
  const float invz = 1.0f / C2.at<float>(2);
  const float ex = fmaf(pKF2->fx * C2.at<float>(0), invz, pKF2->cx);
  const float ey = fmaf(pKF2->fy * C2.at<float>(1), invz, pKF2->cy);

  // only keypoints without a map point take part (:838-842, :861-863)
  std::vector<unsigned char> has1(pKF1->N), has2(pKF2->N);
  for (int i = 0; i < pKF1->N; i++) has1[i] = pKF1->GetMapPoint(i) != NULL;
  for (int i = 0; i < pKF2->N; i++) has2[i] = pKF2->GetMapPoint(i) != NULL;
  const FlatFeatVec fv1(pKF1->mFeatVec), fv2(pKF2->mFeatVec);
  float f12[9];
  for (int r = 0; r < 3; r++)
    for (int c = 0; c < 3; c++) f12[3 * r + c] = F12.at<float>(r, c);

  std::vector<std::pair<int, int> > pairs;
  vMatchedPairs.clear();
  if (pKF1->N == 0 || pKF2->N == 0) return 0;
  const int nmatches = orbx_host::SearchForTriangulation(
      Handle(), BowOf(pKF1->mvKeysUn, pKF1->mDescriptors, pKF1->N, fv1, &has1, pKF1->mvuRight),
      BowOf(pKF2->mvKeysUn, pKF2->mDescriptors, pKF2->N, fv2, &has2, pKF2->mvuRight), f12, ex, ey, &pKF2->mvScaleFactors[0],
      &pKF2->mvLevelSigma2[0], bOnlyStereo, mbCheckOrientation, &pairs);
  vMatchedPairs.reserve(pairs.size());
  for (size_t k = 0; k < pairs.size(); k++) vMatchedPairs.push_back(std::make_pair((size_t)pairs[k].first, (size_t)pairs[k].second));
  return nmatches;
}

// ---- windowed best match in a key frame: Fuse x2, SearchByProjection(pKF, Scw, ...), SearchBySim3 --------------------
namespace {

// A camera pose with the similarity scale divided out (:376-382, :1170-1174)
struct Pose {
  cv::Mat Rcw, tcw, Ow;
  static Pose OfKeyFrame(KeyFrame* pKF) {
    Pose p;
    p.Rcw = pKF->GetRotation(); p.tcw = pKF->GetTranslation(); p.Ow = pKF->GetCameraCenter();
    return p;
  }
  static Pose OfSim3(const cv::Mat& Scw) {
    Pose p;
    const cv::Mat sRcw = Scw.rowRange(0, 3).colRange(0, 3);
    const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
    p.Rcw = sRcw / scw;
    p.tcw = Scw.rowRange(0, 3).col(3) / scw;
    p.Ow = -p.Rcw.t() * p.tcw;
    return p;
  }
};

// Pinhole projection of a camera-frame point into pKF with the reference's rounding (x = X/z first, then fma(x, fx, cx));
// false when behind the camera or outside the image (:1029-1044)
bool ProjectInto(const KeyFrame* pKF, const cv::Mat& p3Dc, float* u, float* v, float* invz) {
  if (p3Dc.at<float>(2) < 0.0f) return false;
  *invz = 1 / p3Dc.at<float>(2);
  *u = fmaf(p3Dc.at<float>(0) * *invz, pKF->fx, pKF->cx);
  *v = fmaf(p3Dc.at<float>(1) * *invz, pKF->fy, pKF->cy);
  return pKF->IsInImage(*u, *v);
}

// The per-point prologue of Fuse and SearchByProjection(pKF, Scw): projection, distance range, viewing angle under 60
// degrees, predicted level (:1019-1061, :401-443). Adds a query when the point passes.
void AddPoseQuery(WindowQueries* q, int who, KeyFrame* pKF, const Pose& pose, MapPoint* pMP, float th) {
  const cv::Mat p3Dw = pMP->GetWorldPos();
  float u = 0.f, v = 0.f, invz = 0.f;
  if (!ProjectInto(pKF, pose.Rcw * p3Dw + pose.tcw, &u, &v, &invz)) return;
  const cv::Mat PO = p3Dw - pose.Ow;
  const float dist3D = cv::norm(PO);
  if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) return;
  if (PO.dot(pMP->GetNormal()) < 0.5 * dist3D) return;
  const int level = pMP->PredictScale(dist3D, pKF);
  q->add(who, u, v, fmaf(-pKF->mbf, invz, u), th * pKF->mvScaleFactors[level], level, pMP->GetDescriptor());
}

struct WindowResult {
  std::vector<int> idx, dist;
};

WindowResult BestInWindow(const orbx_host::FrameView& view, const float* invLevelSigma2, const WindowQueries& q,
                          unsigned char* taken, int th) {
  WindowResult r;
  r.idx.assign(q.n(), -1);
  r.dist.assign(q.n(), 256);
  if (q.n() && view.N)
    orbx_host::SearchBestInWindow(Handle(), view, invLevelSigma2, &q.uv[0], invLevelSigma2 ? &q.ur[0] : nullptr, &q.radius[0],
                                  &q.level[0], &q.desc[0], q.n(), taken, th, &r.idx[0], &r.dist[0]);
  return r;
}

}  // namespace

// ---- LocalMapping::SearchInNeighbors -> Fuse(pKF, vpMapPoints, th) (:995-1160) ---------------------------------------
int ORBmatcher::Fuse(KeyFrame* pKF, const std::vector<MapPoint*>& vpMapPoints, const float th) {
  const Pose pose = Pose::OfKeyFrame(pKF);
  WindowQueries q;
  for (size_t i = 0; i < vpMapPoints.size(); i++) {
    MapPoint* pMP = vpMapPoints[i];
    if (!pMP || pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    AddPoseQuery(&q, (int)i, pKF, pose, pMP, th);
  }
  const WindowResult best = BestInWindow(ViewOf(pKF), &pKF->mvInvLevelSigma2[0], q, nullptr, TH_LOW);
  // apply in point order. The search of a point does not depend on what earlier points did to the map, its skip rule
  // does (a point replaced or added by an earlier iteration is bad or in the key frame by now): test it again here.
  int nFused = 0;
  for (int k = 0; k < q.n(); k++) {
    if (best.dist[k] > TH_LOW) continue;
    MapPoint* pMP = vpMapPoints[q.owner[k]];
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    MapPoint* pMPinKF = pKF->GetMapPoint(best.idx[k]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) {   // keep the point seen from more key frames (:1116-1125)
        if (pMPinKF->Observations() > pMP->Observations()) pMP->Replace(pMPinKF);
        else pMPinKF->Replace(pMP);
      }
    } else {
      pMP->AddObservation(pKF, best.idx[k]);
      pKF->AddMapPoint(pMP, best.idx[k]);
    }
    nFused++;
  }
  return nFused;
}

// ---- LoopClosing::SearchAndFuse -> Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1164-1290) --------------------------
int ORBmatcher::Fuse(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints, float th,
                     std::vector<MapPoint*>& vpReplacePoint) {
  const Pose pose = Pose::OfSim3(Scw);
  const std::set<MapPoint*> alreadyFound = pKF->GetMapPoints();
  WindowQueries q;
  for (size_t i = 0; i < vpPoints.size(); i++) {
    MapPoint* pMP = vpPoints[i];
    if (pMP->isBad() || alreadyFound.count(pMP)) continue;
    AddPoseQuery(&q, (int)i, pKF, pose, pMP, th);
  }
  orbx_host::FrameView view = ViewOf(pKF);
  view.uRight = nullptr;   // no chi-square gate in this variant
  const WindowResult best = BestInWindow(view, nullptr, q, nullptr, TH_LOW);
  int nFused = 0;
  for (int k = 0; k < q.n(); k++) {
    if (best.dist[k] > TH_LOW) continue;
    MapPoint* pMP = vpPoints[q.owner[k]];
    MapPoint* pMPinKF = pKF->GetMapPoint(best.idx[k]);
    if (pMPinKF) {
      if (!pMPinKF->isBad()) vpReplacePoint[q.owner[k]] = pMPinKF;   // the caller replaces it under the map mutex
    } else {
      pMP->AddObservation(pKF, best.idx[k]);
      pKF->AddMapPoint(pMP, best.idx[k]);
    }
    nFused++;
  }
  return nFused;
}

// ---- LoopClosing::ComputeSim3 -> SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:370-497) ---------------------
int ORBmatcher::SearchByProjection(KeyFrame* pKF, cv::Mat Scw, const std::vector<MapPoint*>& vpPoints,
                                   std::vector<MapPoint*>& vpMatched, int th) {
  const Pose pose = Pose::OfSim3(Scw);
  std::set<MapPoint*> alreadyFound(vpMatched.begin(), vpMatched.end());
  alreadyFound.erase(static_cast<MapPoint*>(NULL));
  WindowQueries q;
  for (size_t i = 0; i < vpPoints.size(); i++) {
    MapPoint* pMP = vpPoints[i];
    if (pMP->isBad() || alreadyFound.count(pMP)) continue;
    AddPoseQuery(&q, (int)i, pKF, pose, pMP, (float)th);
  }
  std::vector<unsigned char> taken(vpMatched.size());   // :453-454; the search sets it as keypoints are assigned (:479-483)
  for (size_t i = 0; i < vpMatched.size(); i++) taken[i] = vpMatched[i] != NULL;
  orbx_host::FrameView view = ViewOf(pKF);
  view.uRight = nullptr;
  const WindowResult best = BestInWindow(view, nullptr, q, taken.empty() ? nullptr : &taken[0], TH_LOW);
  int nmatches = 0;
  for (int k = 0; k < q.n(); k++)
    if (best.dist[k] <= TH_LOW) {
      vpMatched[best.idx[k]] = vpPoints[q.owner[k]];
      nmatches++;
    }
  return nmatches;
}

// ---- LoopClosing::ComputeSim3 -> SearchBySim3 (:1305-1560) -------------------------------------------------------------
int ORBmatcher::SearchBySim3(KeyFrame* pKF1, KeyFrame* pKF2, std::vector<MapPoint*>& vpMatches12, const float& s12,
                             const cv::Mat& R12, const cv::Mat& t12, const float th) {
  const cv::Mat R1w = pKF1->GetRotation(), t1w = pKF1->GetTranslation();
  const cv::Mat R2w = pKF2->GetRotation(), t2w = pKF2->GetTranslation();
  const cv::Mat sR12 = s12 * R12;
  const cv::Mat sR21 = (1.0 / s12) * R12.t();
  const cv::Mat t21 = -sR21 * t12;
  const std::vector<MapPoint*> points1 = pKF1->GetMapPointMatches(), points2 = pKF2->GetMapPointMatches();
  const int N1 = (int)points1.size(), N2 = (int)points2.size();

  // matches the caller already holds stay out of both directions (:1337-1350)
  std::vector<unsigned char> done1(N1, 0), done2(N2, 0);
  for (int i = 0; i < N1; i++)
    if (MapPoint* p = vpMatches12[i]) {
      done1[i] = 1;
      const int idx2 = p->GetIndexInKeyFrame(pKF2);
      if (idx2 >= 0 && idx2 < N2) done2[idx2] = 1;
    }

  // one direction: points of `from` through (R, t) of their own camera, then the similarity (sR, ts) into `to`
  // (:1357-1440, :1444-1526); both directions project with the intrinsics of pKF1, as the reference does
  struct Direction {
    static void Gather(WindowQueries* q, const std::vector<MapPoint*>& pts, const std::vector<unsigned char>& done,
                       const cv::Mat& R, const cv::Mat& t, const cv::Mat& sR, const cv::Mat& ts, KeyFrame* to, const KeyFrame* K,
                       float th) {
      for (size_t i = 0; i < pts.size(); i++) {
        MapPoint* pMP = pts[i];
        if (!pMP || done[i] || pMP->isBad()) continue;
        const cv::Mat inOwn = R * pMP->GetWorldPos() + t;
        const cv::Mat inOther = sR * inOwn + ts;
        if (inOther.at<float>(2) < 0.0) continue;
        const float invz = 1.0 / inOther.at<float>(2);
        const float u = fmaf(inOther.at<float>(0) * invz, K->fx, K->cx);
        const float v = fmaf(inOther.at<float>(1) * invz, K->fy, K->cy);
        if (!to->IsInImage(u, v)) continue;
        const float dist3D = cv::norm(inOther);
        if (dist3D < pMP->GetMinDistanceInvariance() || dist3D > pMP->GetMaxDistanceInvariance()) continue;
        const int level = pMP->PredictScale(dist3D, to);
        q->add((int)i, u, v, 0.f, th * to->mvScaleFactors[level], level, pMP->GetDescriptor());
      }
    }
  };
  WindowQueries q12, q21;
  Direction::Gather(&q12, points1, done1, R1w, t1w, sR21, t21, pKF2, pKF1, th);
  Direction::Gather(&q21, points2, done2, R2w, t2w, sR12, t12, pKF1, pKF1, th);
  orbx_host::FrameView view1 = ViewOf(pKF1), view2 = ViewOf(pKF2);
  view1.uRight = view2.uRight = nullptr;
  const WindowResult in2 = BestInWindow(view2, nullptr, q12, nullptr, TH_HIGH);
  const WindowResult in1 = BestInWindow(view1, nullptr, q21, nullptr, TH_HIGH);

  std::vector<int> match1(N1, -1), match2(N2, -1);
  for (int k = 0; k < q12.n(); k++)
    if (in2.dist[k] <= TH_HIGH) match1[q12.owner[k]] = in2.idx[k];
  for (int k = 0; k < q21.n(); k++)
    if (in1.dist[k] <= TH_HIGH) match2[q21.owner[k]] = in1.idx[k];
  int nFound = 0;   // keep what both directions agree on (:1530-1545)
  for (int i1 = 0; i1 < N1; i1++)
    if (match1[i1] >= 0 && match2[match1[i1]] == i1) {
      vpMatches12[i1] = points2[match1[i1]];
      nFound++;
    }
  return nFound;
}

}  // namespace ORB_SLAM2
