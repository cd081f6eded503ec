// TEST INFRASTRUCTURE (oracle/_ref) -- C entry points around the REFERENCE's own ORBextractor class
// (R/include/ORBextractor.h:45-123, R/src/ORBextractor.cpp compiled where it lies against the cv shim in
// this directory). Used only by tests/ to pin the oracle's restatement of the extractor control flow
// (pyramid chaining, FAST cell grid, DistributeOctTree, orientation, descriptors, float codegen)
// against the reference's code itself.
#include <sys/mman.h>

#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "ORBextractor.h"

// ---- allocator control -----------------------------------------------------------------------
// DistributeOctTree sorts (point count, ExtractorNode*) pairs (R/src/ORBextractor.cpp:633,736), so ties
// between equally full nodes are broken by HEAP ADDRESS and the reference's output depends on the
// allocator's history. Mode 1 serves every allocation made inside this library from a bump arena
// with strictly increasing addresses: pointer order == creation order, which is the canonical
// (size, creation sequence) order the oracle pins. Mode 0 is the system allocator, as shipped.
// The library is linked with -Bsymbolic so these operators only replace new/delete for code
// compiled into it (the reference TU and the std:: templates it instantiates).
namespace {
std::atomic<int> g_mode{0};
char* g_arena = nullptr;
std::atomic<size_t> g_off{0};
constexpr size_t kArenaBytes = size_t(8) << 30;  // virtual reservation, touched on demand

void* arena_alloc(size_t n) {
  if (!g_arena) return nullptr;
  const size_t need = (n + 63) & ~size_t(63);
  const size_t off = g_off.fetch_add(need);
  if (off + need > kArenaBytes) return nullptr;
  return g_arena + off;
}
inline bool in_arena(void* p) { return g_arena && (char*)p >= g_arena && (char*)p < g_arena + kArenaBytes; }
}  // namespace

void* operator new(size_t n) {
  if (g_mode.load(std::memory_order_relaxed) == 1) {
    if (void* p = arena_alloc(n)) return p;
  }
  if (void* p = std::malloc(n ? n : 1)) return p;
  throw std::bad_alloc();
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept { if (p && !in_arena(p)) std::free(p); }
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

struct ref_keypoint { float x, y, size, angle, response; int32_t octave, class_id; };

extern "C" {

// 0 = system allocator (heap-address-dependent ties, as the reference ships); 1 = monotonic arena
int ref_set_allocator_mode(int mode) {
  if (mode == 1 && !g_arena) {
    void* p = mmap(nullptr, kArenaBytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
    if (p == MAP_FAILED) return -1;
    g_arena = static_cast<char*>(p);
  }
  g_mode.store(mode);
  return 0;
}
// bytes handed out by the arena so far
uint64_t ref_arena_used(void) { return g_off.load(); }

void* ref_extractor_create(int nfeatures, float scaleFactor, int nlevels, int iniTh, int minTh) {
  return new ORB_SLAM2::ORBextractor(nfeatures, scaleFactor, nlevels, iniTh, minTh);
}
void ref_extractor_destroy(void* h) { delete static_cast<ORB_SLAM2::ORBextractor*>(h); }

// returns the keypoint count (or -count when cap is too small); desc = n x 32
int ref_extract(void* h, const uint8_t* img, int w, int hgt, int stride, ref_keypoint* kps, uint8_t* desc, int cap) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  cv::Mat image(hgt, w, CV_8UC1, const_cast<uint8_t*>(img), (size_t)stride);
  std::vector<cv::KeyPoint> keys;
  cv::Mat descriptors;
  ex(image, cv::Mat(), keys, descriptors);
  const int n = (int)keys.size();
  if (n > cap) return -n;
  for (int i = 0; i < n; i++) {
    kps[i].x = keys[i].pt.x; kps[i].y = keys[i].pt.y; kps[i].size = keys[i].size; kps[i].angle = keys[i].angle;
    kps[i].response = keys[i].response; kps[i].octave = keys[i].octave; kps[i].class_id = keys[i].class_id;
    std::memcpy(desc + (size_t)i * 32, descriptors.ptr(i), 32);
  }
  return n;
}

void ref_tables(void* h, float* scale, float* invScale, float* sigma2, float* invSigma2) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  const int n = ex.GetLevels();
  const std::vector<float> a = ex.GetScaleFactors(), b = ex.GetInverseScaleFactors(), c = ex.GetScaleSigmaSquares(),
                           d = ex.GetInverseScaleSigmaSquares();
  for (int i = 0; i < n; i++) { scale[i] = a[i]; invScale[i] = b[i]; sigma2[i] = c[i]; invSigma2[i] = d[i]; }
}

void ref_level_size(void* h, int level, int* w, int* hgt) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  *w = ex.mvImagePyramid[level].cols;
  *hgt = ex.mvImagePyramid[level].rows;
}
void ref_get_level(void* h, int level, uint8_t* dst, int dstride) {
  ORB_SLAM2::ORBextractor& ex = *static_cast<ORB_SLAM2::ORBextractor*>(h);
  const cv::Mat& m = ex.mvImagePyramid[level];
  for (int r = 0; r < m.rows; r++) std::memcpy(dst + (size_t)r * dstride, m.ptr(r), (size_t)m.cols);
}

}  // extern "C"
