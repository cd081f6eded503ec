// TEST INFRASTRUCTURE (oracle/_ref) -- C entry points around the REFERENCE's own Frame / ORBmatcher / MapPoint
// code: R/src/Frame.cpp, R/src/ORBmatcher.cpp, R/src/MapPoint.cpp, R/src/Map.cpp and R/src/ORBextractor.cpp
// compiled where they lie against the cv shim in this directory (Converter.h and the DBoW2 vocabulary headers
// are shadowed by declarations-only stubs: Eigen is not installed and bag-of-words is out of the hot path).
//
// Used only by tests/ to pin the matcher / stereo rows of SURVEY.md 8a against the reference's code itself:
//   Frame::Frame (mono, stereo)            R/src/Frame.cpp:67-128, 187-250   (extraction, grid, ComputeStereoMatches)
//   ORBmatcher::SearchForInitialization    R/src/ORBmatcher.cpp:499-617
//   ORBmatcher::SearchByProjection(F, vpMapPoints, th)           :63-163
//   ORBmatcher::SearchByProjection(CurrentFrame, LastFrame, ...)  :1564-1718
// The harness builds real Frame and MapPoint objects, calls the reference's member functions and copies plain
// arrays out. Nothing here re-implements the functions under test.
#include <cstdint>
#include <cmath>
#include <cstring>
#include <set>
#include <vector>

#include "Frame.h"
#include "KeyFrame.h"
#include "Map.h"
#include "MapPoint.h"
#include "ORBextractor.h"
#include "ORBmatcher.h"

using namespace ORB_SLAM2;

struct ref_keypoint { float x, y, size, angle, response; int32_t octave, class_id; };

namespace {

struct Front {
  ORBextractor *left, *right;
  cv::Mat K, distCoef;
  float bf, thDepth;
  Map map;
  std::vector<MapPoint*> points;  // owned
  std::vector<KeyFrame*> keyFrames;  // owned
};

int point_index(const Front* f, const MapPoint* p) {
  if (!p) return -1;
  for (size_t k = 0; k < f->points.size(); k++) if (f->points[k] == p) return (int)k;
  return -2;
}

void put_keys(const std::vector<cv::KeyPoint>& keys, ref_keypoint* out) {
  if (!out) return;
  for (size_t i = 0; i < keys.size(); i++) {
    out[i].x = keys[i].pt.x; out[i].y = keys[i].pt.y; out[i].size = keys[i].size; out[i].angle = keys[i].angle;
    out[i].response = keys[i].response; out[i].octave = keys[i].octave; out[i].class_id = keys[i].class_id;
  }
}

cv::Mat wrap(const uint8_t* img, int w, int h, int stride) {
  return cv::Mat(h, w, CV_8UC1, const_cast<uint8_t*>(img), (size_t)stride);
}

}  // namespace

extern "C" {

void* ref_front_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh, float fx, float fy, float cx,
                       float cy, float bf, float thDepth) {
  Front* f = new Front();
  f->left = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  f->right = new ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
  f->K = cv::Mat::eye(3, 3, CV_32F);
  f->K.at<float>(0, 0) = fx; f->K.at<float>(1, 1) = fy; f->K.at<float>(0, 2) = cx; f->K.at<float>(1, 2) = cy;
  f->distCoef = cv::Mat::zeros(4, 1, CV_32F);  // the hot path works on the image as given
  f->bf = bf; f->thDepth = thDepth;
  // image bounds and grid cell sizes are process-wide statics set by the first Frame (R/src/Frame.cpp:222-240)
  Frame::mbInitialComputations = true;
  return f;
}

void ref_front_destroy(void* h) {
  Front* f = static_cast<Front*>(h);
  for (MapPoint* p : f->points) delete p;
  for (KeyFrame* k : f->keyFrames) delete k;
  delete f->left; delete f->right;
  delete f;
}

// R/src/Frame.cpp:187-250: extraction, UndistortKeyPoints (identity here), image bounds, AssignFeaturesToGrid
void* ref_frame_mono(void* h, const uint8_t* img, int w, int hgt, int stride) {
  Front* f = static_cast<Front*>(h);
  return new Frame(wrap(img, w, hgt, stride), 0.0, f->left, static_cast<ORBVocabulary*>(nullptr), f->K, f->distCoef, f->bf,
                   f->thDepth);
}

// R/src/Frame.cpp:67-128: two extraction threads, then ComputeStereoMatches (:551-770)
void* ref_frame_stereo(void* h, const uint8_t* imgL, const uint8_t* imgR, int w, int hgt, int stride) {
  Front* f = static_cast<Front*>(h);
  return new Frame(wrap(imgL, w, hgt, stride), wrap(imgR, w, hgt, stride), 0.0, f->left, f->right,
                   static_cast<ORBVocabulary*>(nullptr), f->K, f->distCoef, f->bf, f->thDepth);
}

void ref_frame_destroy(void* fr) { delete static_cast<Frame*>(fr); }
int ref_frame_n(void* fr) { return static_cast<Frame*>(fr)->N; }
int ref_frame_n_right(void* fr) { return (int)static_cast<Frame*>(fr)->mvKeysRight.size(); }

// any output may be null; bounds = mnMinX, mnMaxX, mnMinY, mnMaxY; mb = baseline
void ref_frame_get(void* fr, ref_keypoint* keys, ref_keypoint* keysUn, uint8_t* desc, float* uRight, float* depth,
                   float* bounds, float* mb) {
  Frame& F = *static_cast<Frame*>(fr);
  put_keys(F.mvKeys, keys);
  put_keys(F.mvKeysUn, keysUn);
  if (desc) for (int i = 0; i < F.N; i++) std::memcpy(desc + (size_t)i * 32, F.mDescriptors.ptr(i), 32);
  if (uRight) for (int i = 0; i < F.N; i++) uRight[i] = F.mvuRight[i];
  if (depth) for (int i = 0; i < F.N; i++) depth[i] = F.mvDepth[i];
  if (bounds) { bounds[0] = Frame::mnMinX; bounds[1] = Frame::mnMaxX; bounds[2] = Frame::mnMinY; bounds[3] = Frame::mnMaxY; }
  if (mb) *mb = F.mb;
}

void ref_frame_get_right(void* fr, ref_keypoint* keys, uint8_t* desc) {
  Frame& F = *static_cast<Frame*>(fr);
  put_keys(F.mvKeysRight, keys);
  if (desc) for (size_t i = 0; i < F.mvKeysRight.size(); i++) std::memcpy(desc + i * 32, F.mDescriptorsRight.ptr((int)i), 32);
}

// level `level` of the pyramid the frame's left (side 0) / right (side 1) extractor holds
void ref_front_level(void* h, int side, int level, int* w, int* hgt, uint8_t* dst, int dstride) {
  Front* f = static_cast<Front*>(h);
  const cv::Mat& m = (side ? f->right : f->left)->mvImagePyramid[level];
  if (w) *w = m.cols;
  if (hgt) *hgt = m.rows;
  if (dst) for (int r = 0; r < m.rows; r++) std::memcpy(dst + (size_t)r * dstride, m.ptr(r), (size_t)m.cols);
}

// Tcw: 4x4 row-major world-to-camera pose (Frame::SetPose -> UpdatePoseMatrices, R/src/Frame.cpp:276-295)
void ref_frame_set_pose(void* fr, const float* Tcw) {
  cv::Mat T(4, 4, CV_32F);
  for (int i = 0; i < 16; i++) T.at<float>(i / 4, i % 4) = Tcw[i];
  static_cast<Frame*>(fr)->SetPose(T);
}

// vbPrevMatched (n1 x 2 floats) is read and updated like the reference's argument; returns the match count
int ref_search_for_initialization(void* fr1, void* fr2, float* prevMatched, int* matches12, int windowSize, float nnratio,
                                  int checkOri) {
  Frame &F1 = *static_cast<Frame*>(fr1), &F2 = *static_cast<Frame*>(fr2);
  std::vector<cv::Point2f> prev(F1.mvKeysUn.size());
  for (size_t i = 0; i < prev.size(); i++) prev[i] = cv::Point2f(prevMatched[2 * i], prevMatched[2 * i + 1]);
  std::vector<int> m12;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int n = matcher.SearchForInitialization(F1, F2, prev, m12, windowSize);
  for (size_t i = 0; i < prev.size(); i++) { prevMatched[2 * i] = prev[i].x; prevMatched[2 * i + 1] = prev[i].y; }
  for (size_t i = 0; i < m12.size(); i++) matches12[i] = m12[i];
  return n;
}

// ---- map points ----------------------------------------------------------------------------------------
// MapPoint(Pos, pMap, pFrame, idxF) (R/src/MapPoint.cpp:64-93): descriptor = row idxF of the frame's descriptors.
// The frame needs a pose (ref_frame_set_pose) first. Returns an index into the front's point table.
int ref_map_point_create(void* h, void* fr, int idx, const float* pos) {
  Front* f = static_cast<Front*>(h);
  cv::Mat P(3, 1, CV_32F);
  for (int i = 0; i < 3; i++) P.at<float>(i) = pos[i];
  f->points.push_back(new MapPoint(P, &f->map, static_cast<Frame*>(fr), idx));
  return (int)f->points.size() - 1;
}

// the fields Frame::isInFrustum leaves on a point for SearchByProjection (R/src/Frame.cpp:360-366)
void ref_map_point_set_track(void* h, int mp, int inView, float u, float v, float uR, int level, float viewCos) {
  MapPoint* p = static_cast<Front*>(h)->points[mp];
  p->mbTrackInView = inView != 0;
  p->mTrackProjX = u; p->mTrackProjY = v; p->mTrackProjXR = uR;
  p->mnTrackScaleLevel = level;
  p->mTrackViewCos = viewCos;
}

// nObs is what Observations() returns (normally maintained by AddObservation with key frames, which are outside
// this harness)
void ref_map_point_set_nobs(void* h, int mp, int nObs) { static_cast<Front*>(h)->points[mp]->nObs = nObs; }

// mp < 0 clears the slot
void ref_frame_set_map_point(void* h, void* fr, int idx, int mp) {
  static_cast<Frame*>(fr)->mvpMapPoints[idx] = mp < 0 ? static_cast<MapPoint*>(nullptr) : static_cast<Front*>(h)->points[mp];
}
void ref_frame_set_outlier(void* fr, int idx, int outlier) { static_cast<Frame*>(fr)->mvbOutlier[idx] = outlier != 0; }

// out[i] = index of the map point attached to keypoint i of the frame, -1 for none
void ref_frame_get_map_points(void* h, void* fr, int* out) {
  Front* f = static_cast<Front*>(h);
  Frame& F = *static_cast<Frame*>(fr);
  for (int i = 0; i < F.N; i++) {
    out[i] = -1;
    if (!F.mvpMapPoints[i]) continue;
    for (size_t k = 0; k < f->points.size(); k++)
      if (f->points[k] == F.mvpMapPoints[i]) { out[i] = (int)k; break; }
  }
}

// R/src/ORBmatcher.cpp:63-163 on the points mps[0..nmp)
int ref_search_by_projection_mp(void* h, void* fr, const int* mps, int nmp, float th, float nnratio, int checkOri) {
  Front* f = static_cast<Front*>(h);
  std::vector<MapPoint*> v(nmp);
  for (int i = 0; i < nmp; i++) v[i] = f->points[mps[i]];
  ORBmatcher matcher(nnratio, checkOri != 0);
  return matcher.SearchByProjection(*static_cast<Frame*>(fr), v, th);
}

// R/src/ORBmatcher.cpp:1564-1718
int ref_search_by_projection_frame(void* cur, void* last, float th, int bMono, float nnratio, int checkOri) {
  ORBmatcher matcher(nnratio, checkOri != 0);
  return matcher.SearchByProjection(*static_cast<Frame*>(cur), *static_cast<Frame*>(last), th, bMono != 0);
}

// The projection of the last frame's points into the current frame that SearchByProjection performs before its
// windowed search (:1574-1611, same expressions on the same Mat arithmetic), exported so that the test can hand
// identical queries to the oracle and to the CUDA path. Per last-frame keypoint i: valid[i] (has a non-outlier
// point that projects inside the image bounds with positive depth), u, v, invzc. Also *forward / *backward.
void ref_project_last_frame(void* curp, void* lastp, int bMono, uint8_t* valid, float* u, float* v, float* invz,
                            int* forward, int* backward) {
  Frame &CurrentFrame = *static_cast<Frame*>(curp), &LastFrame = *static_cast<Frame*>(lastp);
  const cv::Mat Rcw = CurrentFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tcw = CurrentFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat twc = -Rcw.t() * tcw;
  const cv::Mat Rlw = LastFrame.mTcw.rowRange(0, 3).colRange(0, 3);
  const cv::Mat tlw = LastFrame.mTcw.rowRange(0, 3).col(3);
  const cv::Mat tlc = Rlw * twc + tlw;
  *forward = tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  *backward = -tlc.at<float>(2) > CurrentFrame.mb && !bMono;
  for (int i = 0; i < LastFrame.N; i++) {
    valid[i] = 0; u[i] = v[i] = invz[i] = 0;
    MapPoint* pMP = LastFrame.mvpMapPoints[i];
    if (!pMP || LastFrame.mvbOutlier[i]) continue;
    cv::Mat x3Dw = pMP->GetWorldPos();
    cv::Mat x3Dc = Rcw * x3Dw + tcw;
    const float xc = x3Dc.at<float>(0);
    const float yc = x3Dc.at<float>(1);
    const float invzc = 1.0 / x3Dc.at<float>(2);
    if (invzc < 0) continue;
    const float uu = CurrentFrame.fx * xc * invzc + CurrentFrame.cx;
    const float vv = CurrentFrame.fy * yc * invzc + CurrentFrame.cy;
    if (uu < CurrentFrame.mnMinX || uu > CurrentFrame.mnMaxX) continue;
    if (vv < CurrentFrame.mnMinY || vv > CurrentFrame.mnMaxY) continue;
    valid[i] = 1; u[i] = uu; v[i] = vv; invz[i] = invzc;
  }
}

// ---- key frames (R/src/KeyFrame.cpp compiled where it lies) ------------------------------------------------
// KeyFrame(F, pMap, pKFDB) (R/src/KeyFrame.cpp:30-58): copies keypoints, descriptors, grid, feature vector and pose
// of the frame (which needs ref_frame_set_pose first). The key frame database is outside the harness (NULL).
void* ref_keyframe_create(void* h, void* fr) {
  Front* f = static_cast<Front*>(h);
  KeyFrame* kf = new KeyFrame(*static_cast<Frame*>(fr), &f->map, static_cast<KeyFrameDatabase*>(nullptr));
  f->keyFrames.push_back(kf);
  return kf;
}

// mFeatVec is what DBoW2's transform() leaves (outside the path: the vocabulary file is not shipped); the harness
// fills it through DBoW2::FeatureVector::addFeature (Thirdparty/DBoW2/DBoW2/FeatureVector.cpp:31-45): node[i] is the
// vocabulary node of feature i, features are added in index order like transform() does
void ref_frame_set_featvec(void* fr, const int* node, int n) {
  Frame& F = *static_cast<Frame*>(fr);
  F.mFeatVec.clear();
  for (int i = 0; i < n; i++) if (node[i] >= 0) F.mFeatVec.addFeature((DBoW2::NodeId)node[i], (unsigned)i);
}
void ref_keyframe_set_featvec(void* kfp, const int* node, int n) {
  KeyFrame& K = *static_cast<KeyFrame*>(kfp);
  K.mFeatVec.clear();
  for (int i = 0; i < n; i++) if (node[i] >= 0) K.mFeatVec.addFeature((DBoW2::NodeId)node[i], (unsigned)i);
}

// KeyFrame::AddMapPoint + MapPoint::AddObservation, as LocalMapping / Tracking do for a new association
void ref_keyframe_add_map_point(void* h, void* kfp, int idx, int mp) {
  MapPoint* p = static_cast<Front*>(h)->points[mp];
  KeyFrame* kf = static_cast<KeyFrame*>(kfp);
  kf->AddMapPoint(p, (size_t)idx);
  p->AddObservation(kf, (size_t)idx);
}
int ref_keyframe_address_rank(void* h, void* kfp) {   // position of the key frame in pointer order (std::map<KeyFrame*,...>)
  Front* f = static_cast<Front*>(h);
  int r = 0;
  for (KeyFrame* k : f->keyFrames) if (k < static_cast<KeyFrame*>(kfp)) r++;
  return r;
}

// R/src/ORBmatcher.cpp:220-365; out[iF] = index of the map point assigned to frame keypoint iF or -1
int ref_search_by_bow_kf_frame(void* h, void* kfp, void* fr, float nnratio, int checkOri, int* out) {
  Frame& F = *static_cast<Frame*>(fr);
  std::vector<MapPoint*> v;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int n = matcher.SearchByBoW(static_cast<KeyFrame*>(kfp), F, v);
  for (int i = 0; i < F.N; i++) out[i] = point_index(static_cast<Front*>(h), v[i]);
  return n;
}

// R/src/ORBmatcher.cpp:632-772; out[i1] = index of the matched map point of key frame 2 or -1
int ref_search_by_bow_kf_kf(void* h, void* kf1, void* kf2, float nnratio, int checkOri, int* out) {
  std::vector<MapPoint*> v;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int n = matcher.SearchByBoW(static_cast<KeyFrame*>(kf1), static_cast<KeyFrame*>(kf2), v);
  for (size_t i = 0; i < v.size(); i++) out[i] = point_index(static_cast<Front*>(h), v[i]);
  return n;
}

// The epipole of camera 1 in image 2 as SearchForTriangulation's prologue computes it (R/src/ORBmatcher.cpp:790-803,
// same Mat arithmetic; the final expression carries the contraction GCC emits there: fma(fx*C2x, invz, cx)), exported so
// that the test hands identical inputs to the oracle and to the CUDA path.
void ref_epipole(void* kf1p, void* kf2p, float* ex, float* ey) {
  KeyFrame *pKF1 = static_cast<KeyFrame*>(kf1p), *pKF2 = static_cast<KeyFrame*>(kf2p);
  cv::Mat Cw = pKF1->GetCameraCenter();
  cv::Mat R2w = pKF2->GetRotation();
  cv::Mat t2w = pKF2->GetTranslation();
  cv::Mat C2 = R2w * Cw + t2w;
  const float invz = 1.0f / C2.at<float>(2);
  *ex = __builtin_fmaf(pKF2->fx * C2.at<float>(0), invz, pKF2->cx);
  *ey = __builtin_fmaf(pKF2->fy * C2.at<float>(1), invz, pKF2->cy);
}

// R/src/ORBmatcher.cpp:785-983; pairs = (idx1, idx2) x return value
int ref_search_for_triangulation(void* kf1, void* kf2, const float* F12, int onlyStereo, float nnratio, int checkOri,
                                 int* pairs, int cap) {
  cv::Mat F(3, 3, CV_32F);
  for (int i = 0; i < 9; i++) F.at<float>(i / 3, i % 3) = F12[i];
  std::vector<std::pair<size_t, size_t> > v;
  ORBmatcher matcher(nnratio, checkOri != 0);
  const int n = matcher.SearchForTriangulation(static_cast<KeyFrame*>(kf1), static_cast<KeyFrame*>(kf2), F, v, onlyStereo != 0);
  for (size_t i = 0; i < v.size() && (int)i < cap; i++) { pairs[2 * i] = (int)v[i].first; pairs[2 * i + 1] = (int)v[i].second; }
  return n == (int)v.size() ? n : -1000000;
}

void ref_keyframe_tables(void* kfp, float* scaleFactors, float* levelSigma2) {
  KeyFrame& K = *static_cast<KeyFrame*>(kfp);
  for (size_t i = 0; i < K.mvScaleFactors.size(); i++) { scaleFactors[i] = K.mvScaleFactors[i]; levelSigma2[i] = K.mvLevelSigma2[i]; }
}

// ---- MapPoint::ComputeDistinctiveDescriptors (R/src/MapPoint.cpp:306-385) -----------------------------------
// rows: the descriptors the function gathers (:326-333), in the iteration order of the observations map -- the
// flattening a drop-in caller performs. Returns their count.
int ref_map_point_observed_rows(void* h, int mp, uint8_t* rows, int cap) {
  MapPoint* p = static_cast<Front*>(h)->points[mp];
  std::map<KeyFrame*, size_t> obs = p->GetObservations();
  int n = 0;
  for (std::map<KeyFrame*, size_t>::iterator it = obs.begin(); it != obs.end(); ++it) {
    if (it->first->isBad()) continue;
    if (n < cap) std::memcpy(rows + (size_t)n * 32, it->first->mDescriptors.ptr((int)it->second), 32);
    n++;
  }
  return n;
}
// runs the reference's function and returns the descriptor it chose
void ref_map_point_distinctive(void* h, int mp, uint8_t* out32) {
  MapPoint* p = static_cast<Front*>(h)->points[mp];
  p->ComputeDistinctiveDescriptors();
  cv::Mat d = p->GetDescriptor();
  std::memcpy(out32, d.ptr(0), 32);
}

// ---- Fuse / SearchByProjection(KeyFrame*, Scw, ...) ------------------------------------------------------------
// The projection prologue those functions run per map point before their windowed search (R/src/ORBmatcher.cpp
// :1019-1061 and :401-443): same Mat arithmetic, same MapPoint members (GetNormal, PredictScale, ...); the scalar
// expressions carry the contractions of the reference's object code (u = fma(x, fx, cx), ur = fma(-bf, invz, u)).
// Exported so that the test hands identical queries to the oracle and to the CUDA path.
static bool window_query(KeyFrame* pKF, const cv::Mat& Rcw, const cv::Mat& tcw, const cv::Mat& Ow, MapPoint* pMP, float th,
                         float* u, float* v, float* ur, float* radius, int* level) {
  cv::Mat p3Dw = pMP->GetWorldPos();
  cv::Mat p3Dc = Rcw * p3Dw + tcw;
  if (p3Dc.at<float>(2) < 0.0f) return false;
  const float invz = 1 / p3Dc.at<float>(2);
  const float x = p3Dc.at<float>(0) * invz;
  const float y = p3Dc.at<float>(1) * invz;
  *u = __builtin_fmaf(x, pKF->fx, pKF->cx);
  *v = __builtin_fmaf(y, pKF->fy, pKF->cy);
  if (!pKF->IsInImage(*u, *v)) return false;
  *ur = __builtin_fmaf(-pKF->mbf, invz, *u);
  const float maxDistance = pMP->GetMaxDistanceInvariance();
  const float minDistance = pMP->GetMinDistanceInvariance();
  cv::Mat PO = p3Dw - Ow;
  const float dist3D = cv::norm(PO);
  if (dist3D < minDistance || dist3D > maxDistance) return false;
  cv::Mat Pn = pMP->GetNormal();
  if (PO.dot(Pn) < 0.5 * dist3D) return false;
  *level = pMP->PredictScale(dist3D, pKF);
  *radius = th * pKF->mvScaleFactors[*level];
  return true;
}

// queries of Fuse(pKF, vpMapPoints, th) (:995-1160); valid[i] = 0 for points the loop skips (:1012-1061)
void ref_fuse_queries(void* h, void* kfp, const int* mps, int nmp, float th, uint8_t* valid, float* uv, float* ur, float* radius,
                      int* level, uint8_t* desc) {
  Front* f = static_cast<Front*>(h);
  KeyFrame* pKF = static_cast<KeyFrame*>(kfp);
  cv::Mat Rcw = pKF->GetRotation(), tcw = pKF->GetTranslation(), Ow = pKF->GetCameraCenter();
  for (int i = 0; i < nmp; i++) {
    MapPoint* pMP = f->points[mps[i]];
    valid[i] = 0; uv[2 * i] = uv[2 * i + 1] = ur[i] = radius[i] = 0; level[i] = 0;
    std::memcpy(desc + (size_t)i * 32, pMP->GetDescriptor().ptr(0), 32);
    if (pMP->isBad() || pMP->IsInKeyFrame(pKF)) continue;
    valid[i] = window_query(pKF, Rcw, tcw, Ow, pMP, th, &uv[2 * i], &uv[2 * i + 1], &ur[i], &radius[i], &level[i]);
  }
}
int ref_fuse(void* h, void* kfp, const int* mps, int nmp, float th) {
  Front* f = static_cast<Front*>(h);
  std::vector<MapPoint*> v(nmp);
  for (int i = 0; i < nmp; i++) v[i] = f->points[mps[i]];
  ORBmatcher matcher(0.6f, true);
  return matcher.Fuse(static_cast<KeyFrame*>(kfp), v, th);
}
void ref_keyframe_get_map_points(void* h, void* kfp, int* out) {
  KeyFrame* kf = static_cast<KeyFrame*>(kfp);
  for (int i = 0; i < kf->N; i++) out[i] = point_index(static_cast<Front*>(h), kf->GetMapPoint((size_t)i));
}
void ref_keyframe_inv_level_sigma2(void* kfp, float* out) {
  KeyFrame& K = *static_cast<KeyFrame*>(kfp);
  for (size_t i = 0; i < K.mvInvLevelSigma2.size(); i++) out[i] = K.mvInvLevelSigma2[i];
}
int ref_map_point_observations(void* h, int mp) { return static_cast<Front*>(h)->points[mp]->Observations(); }
int ref_map_point_is_bad(void* h, int mp) { return static_cast<Front*>(h)->points[mp]->isBad() ? 1 : 0; }

// SearchByProjection(pKF, Scw, vpPoints, vpMatched, th) (:370-497). matched: pKF->N entries, map point index or -1,
// read and updated like the reference's vpMatched.
static void scw_parts(const float* Scw16, cv::Mat* Scw, cv::Mat* Rcw, cv::Mat* tcw, cv::Mat* Ow) {
  *Scw = cv::Mat(4, 4, CV_32F);
  for (int i = 0; i < 16; i++) Scw->at<float>(i / 4, i % 4) = Scw16[i];
  cv::Mat sRcw = Scw->rowRange(0, 3).colRange(0, 3);
  const float scw = sqrt(sRcw.row(0).dot(sRcw.row(0)));
  *Rcw = sRcw / scw;
  *tcw = Scw->rowRange(0, 3).col(3) / scw;
  *Ow = -Rcw->t() * *tcw;
}
void ref_scw_queries(void* h, void* kfp, const float* Scw16, const int* mps, int nmp, const int* matched, int th, uint8_t* valid,
                     float* uv, float* radius, int* level, uint8_t* desc) {
  Front* f = static_cast<Front*>(h);
  KeyFrame* pKF = static_cast<KeyFrame*>(kfp);
  cv::Mat Scw, Rcw, tcw, Ow;
  scw_parts(Scw16, &Scw, &Rcw, &tcw, &Ow);
  std::set<MapPoint*> spAlreadyFound;
  for (int i = 0; i < pKF->N; i++) if (matched[i] >= 0) spAlreadyFound.insert(f->points[matched[i]]);
  for (int i = 0; i < nmp; i++) {
    MapPoint* pMP = f->points[mps[i]];
    float ur;
    valid[i] = 0; uv[2 * i] = uv[2 * i + 1] = radius[i] = 0; level[i] = 0;
    std::memcpy(desc + (size_t)i * 32, pMP->GetDescriptor().ptr(0), 32);
    if (pMP->isBad() || spAlreadyFound.count(pMP)) continue;
    valid[i] = window_query(pKF, Rcw, tcw, Ow, pMP, (float)th, &uv[2 * i], &uv[2 * i + 1], &ur, &radius[i], &level[i]);
  }
}
int ref_search_by_projection_kf_scw(void* h, void* kfp, const float* Scw16, const int* mps, int nmp, int* matched, int th) {
  Front* f = static_cast<Front*>(h);
  KeyFrame* pKF = static_cast<KeyFrame*>(kfp);
  cv::Mat Scw(4, 4, CV_32F);
  for (int i = 0; i < 16; i++) Scw.at<float>(i / 4, i % 4) = Scw16[i];
  std::vector<MapPoint*> pts(nmp), vm(pKF->N, static_cast<MapPoint*>(nullptr));
  for (int i = 0; i < nmp; i++) pts[i] = f->points[mps[i]];
  for (int i = 0; i < pKF->N; i++) if (matched[i] >= 0) vm[i] = f->points[matched[i]];
  ORBmatcher matcher(0.75f, true);
  const int n = matcher.SearchByProjection(pKF, Scw, pts, vm, th);
  for (int i = 0; i < pKF->N; i++) matched[i] = point_index(f, vm[i]);
  return n;
}

// SearchByProjection(CurrentFrame, pKF, sAlreadyFound, th, ORBdist) (:1723-1851), Tracking::Relocalization's second pass
int ref_search_by_projection_frame_kf(void* h, void* cur, void* kfp, const int* found, int nfound, float th, int orbDist,
                                      float nnratio, int checkOri) {
  Front* f = static_cast<Front*>(h);
  std::set<MapPoint*> sFound;
  for (int i = 0; i < nfound; i++) sFound.insert(f->points[found[i]]);
  ORBmatcher matcher(nnratio, checkOri != 0);
  return matcher.SearchByProjection(*static_cast<Frame*>(cur), static_cast<KeyFrame*>(kfp), sFound, th, orbDist);
}

// Fuse(pKF, Scw, vpPoints, th, vpReplacePoint) (:1164-1290); replace[i] = index of the key frame's point that point i
// is to replace, or -1
int ref_fuse_scw(void* h, void* kfp, const float* Scw16, const int* mps, int nmp, float th, int* replace) {
  Front* f = static_cast<Front*>(h);
  cv::Mat Scw(4, 4, CV_32F);
  for (int i = 0; i < 16; i++) Scw.at<float>(i / 4, i % 4) = Scw16[i];
  std::vector<MapPoint*> pts(nmp), rep(nmp, static_cast<MapPoint*>(nullptr));
  for (int i = 0; i < nmp; i++) pts[i] = f->points[mps[i]];
  ORBmatcher matcher(0.8f, true);
  const int n = matcher.Fuse(static_cast<KeyFrame*>(kfp), Scw, pts, th, rep);
  for (int i = 0; i < nmp; i++) replace[i] = point_index(f, rep[i]);
  return n;
}

// SearchBySim3 (:1305-1560); matches12: pKF1->N entries, map point index or -1, read and updated like vpMatches12
int ref_search_by_sim3(void* h, void* kf1, void* kf2, int* matches12, float s12, const float* R12, const float* t12, float th) {
  Front* f = static_cast<Front*>(h);
  KeyFrame *pKF1 = static_cast<KeyFrame*>(kf1), *pKF2 = static_cast<KeyFrame*>(kf2);
  cv::Mat R(3, 3, CV_32F), t(3, 1, CV_32F);
  for (int i = 0; i < 9; i++) R.at<float>(i / 3, i % 3) = R12[i];
  for (int i = 0; i < 3; i++) t.at<float>(i) = t12[i];
  std::vector<MapPoint*> vm(pKF1->N, static_cast<MapPoint*>(nullptr));
  for (int i = 0; i < pKF1->N; i++) if (matches12[i] >= 0) vm[i] = f->points[matches12[i]];
  ORBmatcher matcher(0.75f, true);
  const int n = matcher.SearchBySim3(pKF1, pKF2, vm, s12, R, t, th);
  for (int i = 0; i < pKF1->N; i++) matches12[i] = point_index(f, vm[i]);
  return n;
}

// Hamming distance through whichever ORBmatcher this library carries (static member, :1901-1917)
int ref_descriptor_distance(const uint8_t* a, const uint8_t* b) {
  cv::Mat ma(1, 32, CV_8UC1, const_cast<uint8_t*>(a), 32), mb(1, 32, CV_8UC1, const_cast<uint8_t*>(b), 32);
  return ORBmatcher::DescriptorDistance(ma, mb);
}

}  // extern "C"
