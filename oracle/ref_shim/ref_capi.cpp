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
size_t g_arenaBytes = 0;
size_t g_off = 0;                                // bump pointer, under g_lock
constexpr size_t kChunk = size_t(2) << 20;       // physical memory goes back to the system chunk by chunk
constexpr size_t kHeader = 64;                   // block size, kept in front of the payload (blocks are 64-byte aligned)
int* g_live = nullptr;                           // live blocks per chunk
std::atomic_flag g_lock = ATOMIC_FLAG_INIT;

struct Lock {
  Lock() { while (g_lock.test_and_set(std::memory_order_acquire)) {} }
  ~Lock() { g_lock.clear(std::memory_order_release); }
};

// Addresses only ever grow (pointer order == creation order for the lifetime of the process); a chunk whose blocks
// have all been freed and that the bump pointer has left behind is handed back with MADV_DONTNEED, so a long test
// session keeps the canonical order without keeping the memory.
void* arena_alloc(size_t n) {
  if (!g_arena) return nullptr;
  const size_t need = ((n + 63) & ~size_t(63)) + kHeader;
  Lock lock;
  if (g_off + need > g_arenaBytes) return nullptr;
  const size_t off = g_off;
  g_off += need;
  for (size_t c = off / kChunk; c <= (off + need - 1) / kChunk; c++) g_live[c]++;
  *reinterpret_cast<size_t*>(g_arena + off) = need;
  return g_arena + off + kHeader;
}
inline bool in_arena(void* p) { return g_arena && (char*)p >= g_arena && (char*)p < g_arena + g_arenaBytes; }
void arena_free(void* p) {
  Lock lock;
  const size_t off = (size_t)((char*)p - g_arena) - kHeader;
  const size_t need = *reinterpret_cast<size_t*>(g_arena + off);
  for (size_t c = off / kChunk; c <= (off + need - 1) / kChunk; c++)
    if (--g_live[c] == 0 && (c + 1) * kChunk <= g_off) madvise(g_arena + c * kChunk, kChunk, MADV_DONTNEED);
}
}  // namespace

void* operator new(size_t n) {
  if (g_mode.load(std::memory_order_relaxed) == 1) {
    if (void* p = arena_alloc(n)) return p;
  }
  if (void* p = std::malloc(n ? n : 1)) return p;
  throw std::bad_alloc();
}
void* operator new[](size_t n) { return operator new(n); }
void operator delete(void* p) noexcept {
  if (!p) return;
  if (in_arena(p)) arena_free(p);
  else std::free(p);
}
void operator delete[](void* p) noexcept { operator delete(p); }
void operator delete(void* p, size_t) noexcept { operator delete(p); }
void operator delete[](void* p, size_t) noexcept { operator delete(p); }

struct ref_keypoint { float x, y, size, angle, response; int32_t octave, class_id; };

extern "C" {

// 0 = system allocator (heap-address-dependent ties, as the reference ships); 1 = monotonic arena
int ref_set_allocator_mode(int mode) {
  if (mode == 1 && !g_arena) {
    // virtual reservation, touched on demand; smaller if the system refuses the large one
    for (size_t bytes = size_t(256) << 30; bytes >= (size_t(8) << 30); bytes >>= 1) {
      void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS | MAP_NORESERVE, -1, 0);
      if (p == MAP_FAILED) continue;
      void* live = mmap(nullptr, (bytes / kChunk) * sizeof(int), PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
      if (live == MAP_FAILED) { munmap(p, bytes); continue; }
      g_live = static_cast<int*>(live);
      g_arenaBytes = bytes;
      g_arena = static_cast<char*>(p);
      break;
    }
    if (!g_arena) return -1;
  }
  g_mode.store(mode);
  return 0;
}
// bytes handed out by the arena so far
uint64_t ref_arena_used(void) { Lock lock; return g_off; }

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
