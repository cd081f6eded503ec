// C entry points around the C++ mirror class, used by tests/test_gpu_host_mirror.py to drive
// ORB_SLAM2::ORBextractor (host/ORBextractor.h) exactly as Frame.cpp does, without OpenCV installed
// (the test build compiles the mirror against the cv shim; a real deployment compiles it against OpenCV).
#include <cstdint>
#include <cstring>
#include <vector>

#include "ORBextractor.h"
#include "orbx_shard.h"

struct mirror_keypoint { float x, y, size, angle, response; int32_t octave, class_id; };

extern "C" {

void* mirror_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void mirror_destroy(void* h) { delete static_cast<ORB_SLAM2::ORBextractor*>(h); }

int mirror_extract(void* h, const uint8_t* img, int w, int hgt, int stride, mirror_keypoint* kps, uint8_t* desc, int cap) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  cv::Mat image(hgt, w, CV_8UC1, const_cast<uint8_t*>(img), (size_t)stride);
  std::vector<cv::KeyPoint> keys(3);   // must be cleared by the call
  cv::Mat descriptors;
  ex(image, cv::Mat(), keys, descriptors);
  const int n = (int)keys.size();
  if (n > cap) return -n;
  if (n > 0 && (descriptors.rows != n || descriptors.cols != 32)) return -1000000;
  if (n == 0 && !descriptors.empty()) return -1000001;
  std::memcpy(kps, keys.data(), (size_t)n * sizeof(mirror_keypoint));
  for (int i = 0; i < n; i++) std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
  return n;
}

void mirror_tables(void* h, float* scale, float* invScale, float* sigma2, float* invSigma2) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  const int n = ex.GetLevels();
  const std::vector<float> a = ex.GetScaleFactors(), b = ex.GetInverseScaleFactors(), c = ex.GetScaleSigmaSquares(),
                           d = ex.GetInverseScaleSigmaSquares();
  for (int i = 0; i < n; i++) { scale[i] = a[i]; invScale[i] = b[i]; sigma2[i] = c[i]; invSigma2[i] = d[i]; }
}

void mirror_level_size(void* h, int level, int* w, int* hgt) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  *w = ex.mvImagePyramid[level].cols;
  *hgt = ex.mvImagePyramid[level].rows;
}
void mirror_get_level(void* h, int level, uint8_t* dst, int dstride) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  const cv::Mat& m = ex.mvImagePyramid[level];
  for (int r = 0; r < m.rows; r++) std::memcpy(dst + (size_t)r * dstride, m.ptr(r), (size_t)m.cols);
}

}  // extern "C"

// ---- windowed searches (host/orbx_search.h) ---------------------------------------------------
#include "orbx_search.h"

extern "C" {

static orbx_host::FrameView view(const orbx_keypoint* k, const unsigned char* d, int n, const float* b, const float* ur) {
  orbx_host::FrameView v;
  v.keysUn = k; v.descriptors = d; v.N = n; v.minX = b[0]; v.maxX = b[1]; v.minY = b[2]; v.maxY = b[3]; v.uRight = ur;
  return v;
}

int mirror_search_for_initialization(orbx_matcher* m, const orbx_keypoint* k1, const unsigned char* d1, int n1,
                                     const orbx_keypoint* k2, const unsigned char* d2, int n2, const float* bounds,
                                     float* prevMatched, int* matches12, int window, float nnratio, int checkOri) {
  return orbx_host::SearchForInitialization(m, view(k1, d1, n1, bounds, 0), view(k2, d2, n2, bounds, 0), prevMatched,
                                            matches12, window, nnratio, checkOri != 0);
}

int mirror_search_by_projection_mp(orbx_matcher* m, const orbx_keypoint* k, const unsigned char* d, int n,
                                   const float* uright, unsigned char* hasMP, const float* bounds, const float* mpUV,
                                   const float* mpR, const int* mpLevel, const float* mpUR, const unsigned char* mpDesc,
                                   int nMP, const float* scaleFactors, float nnratio, int* mpMatch,
                                   const unsigned char* mpHasObs) {
  return orbx_host::SearchByProjection(m, view(k, d, n, bounds, uright), hasMP, mpUV, mpR, mpLevel, mpUR, mpDesc, nMP,
                                       scaleFactors, nnratio, mpMatch, mpHasObs);
}

int mirror_search_by_projection_frame(orbx_matcher* m, const orbx_keypoint* k, const unsigned char* d, int n,
                                      const float* uright, unsigned char* hasMP, const float* bounds, const float* qUV,
                                      const float* qR, const int* qMin, const int* qMax, const float* qUR,
                                      const float* qAngle, const unsigned char* qDesc, int nQ, int thDist, int checkOri,
                                      int* curMatch, const unsigned char* qHasObs) {
  return orbx_host::SearchByProjection(m, view(k, d, n, bounds, uright), hasMP, qUV, qR, qMin, qMax, qUR, qAngle, qDesc, nQ,
                                       thDist, checkOri != 0, curMatch, qHasObs);
}

static orbx_host::BowView bow_view(const orbx_keypoint* k, const unsigned char* d, int n, const int* nodes, const int* off,
                                   const int* idx, int nn, const unsigned char* flag, const float* ur) {
  orbx_host::BowView v;
  v.keysUn = k; v.descriptors = d; v.N = n; v.featVec.nodes = nodes; v.featVec.off = off; v.featVec.idx = idx; v.featVec.n = nn;
  v.flag = flag; v.uRight = ur;
  return v;
}

int mirror_search_by_bow_kf_frame(orbx_matcher* m, const orbx_keypoint* kKF, const unsigned char* dKF, int nKF,
                                  const int* nodesKF, const int* offKF, const int* idxKF, int nnKF, const unsigned char* usable,
                                  const orbx_keypoint* kF, const unsigned char* dF, int nF, const int* nodesF, const int* offF,
                                  const int* idxF, int nnF, float nnratio, int checkOri, int* matchF) {
  return orbx_host::SearchByBoW(m, bow_view(kKF, dKF, nKF, nodesKF, offKF, idxKF, nnKF, usable, 0),
                                bow_view(kF, dF, nF, nodesF, offF, idxF, nnF, 0, 0), nnratio, checkOri != 0, matchF);
}

int mirror_search_by_bow_kf_kf(orbx_matcher* m, const orbx_keypoint* k1, const unsigned char* d1, int n1, const int* nodes1,
                               const int* off1, const int* idx1, int nn1, const unsigned char* usable1, const orbx_keypoint* k2,
                               const unsigned char* d2, int n2, const int* nodes2, const int* off2, const int* idx2, int nn2,
                               const unsigned char* usable2, float nnratio, int checkOri, int* matches12) {
  return orbx_host::SearchByBoWKeyFrames(m, bow_view(k1, d1, n1, nodes1, off1, idx1, nn1, usable1, 0),
                                         bow_view(k2, d2, n2, nodes2, off2, idx2, nn2, usable2, 0), nnratio, checkOri != 0,
                                         matches12);
}

int mirror_search_for_triangulation(orbx_matcher* m, const orbx_keypoint* k1, const unsigned char* d1, int n1, const int* nodes1,
                                    const int* off1, const int* idx1, int nn1, const unsigned char* hasMP1, const float* ur1,
                                    const orbx_keypoint* k2, const unsigned char* d2, int n2, const int* nodes2, const int* off2,
                                    const int* idx2, int nn2, const unsigned char* hasMP2, const float* ur2, const float* F12,
                                    float ex, float ey, const float* scaleFactors2, const float* levelSigma2_2, int onlyStereo,
                                    int checkOri, int* pairs) {
  std::vector<std::pair<int, int> > v;
  const int n = orbx_host::SearchForTriangulation(m, bow_view(k1, d1, n1, nodes1, off1, idx1, nn1, hasMP1, ur1),
                                                  bow_view(k2, d2, n2, nodes2, off2, idx2, nn2, hasMP2, ur2), F12, ex, ey,
                                                  scaleFactors2, levelSigma2_2, onlyStereo != 0, checkOri != 0, &v);
  for (size_t i = 0; i < v.size(); i++) { pairs[2 * i] = v[i].first; pairs[2 * i + 1] = v[i].second; }
  return n == (int)v.size() ? n : -1000000;
}

void mirror_search_best_in_window(orbx_matcher* m, const orbx_keypoint* k, const unsigned char* d, int n, const float* uright,
                                  const float* bounds, const float* invLevelSigma2, const float* qUV, const float* qUR,
                                  const float* qRadius, const int* qLevel, const unsigned char* qDesc, int nQ,
                                  unsigned char* taken, int th, int* bestIdx, int* bestDist) {
  orbx_host::SearchBestInWindow(m, view(k, d, n, bounds, uright), invLevelSigma2, qUV, qUR, qRadius, qLevel, qDesc, nQ, taken, th,
                                bestIdx, bestDist);
}

}  // extern "C"

// ---- vocabulary transform (host/orbx_bow.h) -------------------------------------------------------
#include "orbx_bow.h"

extern "C" {

void* mirror_voc_load(orbx_matcher* m, const char* path) {
  orbx_host::Vocabulary* v = new orbx_host::Vocabulary();
  if (!v->loadFromTextFile(m, path)) { delete v; return 0; }
  return v;
}
void mirror_voc_destroy(void* h) { delete static_cast<orbx_host::Vocabulary*>(h); }
int mirror_voc_words(void* h) { return static_cast<orbx_host::Vocabulary*>(h)->words; }

// flattened like ref_voc_transform (oracle/ref_shim/ref_voc_capi.cpp)
int mirror_voc_transform(void* h, const unsigned char* desc, int n, int levelsup, unsigned* words, double* values,
                         unsigned* nodes, int* off, unsigned* idx, int* nNodes, int* nIdx) {
  std::map<unsigned, double> v;
  std::map<unsigned, std::vector<unsigned> > fv;
  static_cast<orbx_host::Vocabulary*>(h)->transform(desc, n, &v, &fv, levelsup);
  int k = 0;
  for (std::map<unsigned, double>::const_iterator it = v.begin(); it != v.end(); ++it, ++k) { words[k] = it->first; values[k] = it->second; }
  int a = 0, b = 0;
  off[0] = 0;
  for (std::map<unsigned, std::vector<unsigned> >::const_iterator it = fv.begin(); it != fv.end(); ++it, ++a) {
    nodes[a] = it->first;
    for (size_t j = 0; j < it->second.size(); j++) idx[b++] = it->second[j];
    off[a + 1] = b;
  }
  *nNodes = a; *nIdx = b;
  return k;
}

}  // extern "C"

// orbx_host::ShardGroup (host/orbx_shard.h) over `ndev` devices: sharded top-2 of nq queries against ndb rows, and a
// sequence of `count` frames extracted by the group. Outputs are the caller's; returns 0, or -1 with the message in
// err (the wrapper throws).
extern "C" int mirror_shard_run(int ndev, const uint8_t* db, long long ndb, const uint8_t* q, int nq, float ratio, int th,
                                int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted, const uint8_t* frames, int count,
                                int w, int hgt, int nfeatures, mirror_keypoint* kps, uint8_t* desc, int cap, int32_t* n,
                                char* err, int errCap) {
  try {
    std::vector<int> devices;
    for (int i = 0; i < ndev; i++) devices.push_back(i);
    orbx_host::ShardGroup group(devices);
    if (group.Ranks() != ndev || group.LocalRanks() != ndev) throw std::runtime_error("group size");
    group.SetDatabase(db, ndb);
    const orbx_host::ShardGroup::Matches m = group.MatchBf2(q, nq, ratio, th);
    std::memcpy(idx, &m.idx[0], (size_t)nq * 4);
    std::memcpy(d1, &m.d1[0], (size_t)nq * 4);
    std::memcpy(d2, &m.d2[0], (size_t)nq * 4);
    std::memcpy(accepted, &m.accepted[0], (size_t)nq);
    orbx_params p;
    p.nfeatures = nfeatures; p.scale_factor = 1.2f; p.nlevels = 8; p.ini_th_fast = 20; p.min_th_fast = 7; p.device = 0;
    p.max_width = w; p.max_height = hgt; p.max_batch = 8;
    std::vector<const uint8_t*> imgs(count);
    for (int f = 0; f < count; f++) imgs[f] = frames + (size_t)f * w * hgt;
    group.ExtractSequence(p, &imgs[0], count, w, hgt, (size_t)w, reinterpret_cast<orbx_keypoint*>(kps), desc, cap, n);
    return 0;
  } catch (const std::exception& e) {
    std::strncpy(err, e.what(), (size_t)errCap - 1);
    err[errCap - 1] = 0;
    return -1;
  }
}

