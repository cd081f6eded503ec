// Shared declarations for the liborbx.so translation units (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <mutex>
#include <stdint.h>

#include "../../include/orbx.h"

namespace orbx {

// Programmatic dependent launch between the consecutive kernels of the extraction pipeline: a kernel launched
// with launch_pdl() is staged on the device while its predecessor in the stream still runs, and blocks in
// pdl_wait() until that grid has completed and flushed. Every pipeline kernel calls pdl_wait() before it touches
// anything a predecessor wrote. The completion trigger is left implicit (CTA exit): the launch latency of kernel
// k+1 hides behind kernel k (11 launches per chunk; 64-frame launch sets went from 119 k to 124 k frames/s), while
// an early griddepcontrol.launch_dependents made the waiting CTAs of k+1 crowd out k's tail (512 frames: 3.58 ->
// 4.27 ms) and is not used.
#ifdef __CUDACC__
// cudaFuncAttributeMaxDynamicSharedMemorySize belongs to the (kernel, device) pair and is shared by every host thread
// and every handle of the process: two extractors of different sizes on two threads (the stereo Frame constructor
// runs one per eye) must never lower it under each other's launches. Raise it monotonically under a lock; `state`
// is one zero-initialised static per kernel.
struct SmemLimit {
  std::mutex mu;
  size_t bytes[64] = {};
};
inline cudaError_t raise_dynamic_smem(const void* kernel, size_t bytes, SmemLimit* state) {
  if (bytes <= 48 * 1024) return cudaSuccess;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  std::lock_guard<std::mutex> lock(state->mu);
  if (bytes <= state->bytes[dev]) return cudaSuccess;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
  if (e == cudaSuccess) state->bytes[dev] = bytes;
  return e;
}

__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Per host thread: launches of the fork / join form of the pipeline (small batches, captured into a CUDA graph) are
// plain launches; griddepcontrol.wait returns at once in a grid that was not launched programmatically.
inline thread_local bool tl_programmatic_launch = true;

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr; cfg.numAttrs = tl_programmatic_launch ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

constexpr int kEdge = 19;         // EDGE_THRESHOLD  (R/src/ORBextractor.cpp:76)
constexpr int kMinBorder = 16;    // EDGE_THRESHOLD-3 (:831)
constexpr int kHalfPatch = 15;    // HALF_PATCH_SIZE (:75)
constexpr int kPatch = 31;        // PATCH_SIZE      (:74)
constexpr int kMaxLevels = ORBX_MAX_LEVELS;
constexpr int kMaxCellDim = 64;   // wCell/hCell upper bound handled by the FAST kernel (< 60 by construction)
constexpr int kMaxNodes = 3400;   // per-level octree quota + slack: 66 B of smem per node, 224 KB of the 227 KB a CTA may take

// Geometry of one pyramid level, fixed per (frame shape, extractor parameters).
struct LevelGeom {
  int w, h;            // level size (:1203)
  int pitch;           // row pitch in bytes of our own planes (multiple of 16)
  int maxBX, maxBY;    // w-16, h-16 (:833-834); minBorder is 16
  int nCols, nRows;    // cell grid (:844-845)
  int wCell, hCell;    // (:846-847)
  int nfeat;           // mnFeaturesPerLevel[level]
  int nIni;            // octree roots (:577)
  float hX;            // root width (:580)
  int candCap;         // capacity of the candidate list of one (frame, level)
  int keptCap;         // capacity of the kept list of one (frame, level)
  int cellBase;        // first entry of this level in the FAST cell table (cells the reference skips are not listed)
  int tileBase;        // first flat tile id of this level in the blur grid
  int tilesX, tilesY;
  float scale;         // mvScaleFactor[level]
  float patchSize;     // (float)(int)(31*scale)  (:905,914)
  long long planeOff;  // byte offset of frame 0 of this level in the pyramid / blur arenas
  long long planeBytes;  // pitch*h
  long long candOff;   // element offset of (frame 0, level) in the candidate arena
  long long keptOff;   // element offset in the kept arena
};

// One FAST cell the reference scans (R/src/ORBextractor.cpp:851-870), worked out once per frame shape on the host
// instead of ~150 integer instructions per cell and launch in the kernel. Tile column 0 of the cell's patch is image
// column gx0 (a multiple of 4); the first scanned pixel sits at tile column a (4..7), the last at b - 1.
struct alignas(16) FastCellDesc {
  uint8_t level, ch, a, b;   // ch = patch rows (scan rows = ch - 6)
  uint16_t iniY, gx0;        // first patch row / tile column 0, in level coordinates
  int16_t outX0, outY0;      // candidate x = outX0 + 4 * (q + 1) + i, y = outY0 + row (relative to minBorder)
  uint32_t pad;
};

struct FrameSetGeom {
  int nlevels;
  int iniTh, minTh;
  int totalCells;      // entries of the FAST cell table
  int totalTiles;      // blur grid x-dimension
  long long pyrFrameStride;   // bytes between the same level of consecutive frames is planeBytes; arenas are [level][frame]
  int candPerFrame;    // sum of candCap
  int keptPerFrame;    // sum of keptCap
  LevelGeom lv[kMaxLevels];
};

// candidate packing: x | y<<12 | score<<24, coordinates relative to minBorder (what the octree sees)
__host__ __device__ inline uint32_t pack_cand(int x, int y, int score) {
  return (uint32_t)x | ((uint32_t)y << 12) | ((uint32_t)score << 24);
}
__host__ __device__ inline int cand_x(uint32_t c) { return (int)(c & 0xfff); }
__host__ __device__ inline int cand_y(uint32_t c) { return (int)((c >> 12) & 0xfff); }
__host__ __device__ inline int cand_score(uint32_t c) { return (int)(c >> 24); }

// Host-side error plumbing (thread-local message).
void set_error(const char* fmt, ...);
#define ORBX_CUDA_TRY(expr)                                                              \
  do {                                                                                   \
    cudaError_t _e = (expr);                                                             \
    if (_e != cudaSuccess) {                                                             \
      ::orbx::set_error("%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), __FILE__, __LINE__); \
      return ORBX_ERR_CUDA;                                                              \
    }                                                                                    \
  } while (0)

// ---- kernel launchers (defined in the .cu files) ------------------------------------------------
struct alignas(8) ResizeTap { short p0, p1, c0, c1; };  // two source indices and their 11-bit weights
// Four consecutive destination pixels of a row (resize4_kernel): the <= 8 source bytes they read start
// at byte 4*word + shift/8 of the source row; sel01/sel23 are PRMT selectors gathering
// (left0,right0,left1,right1) / (left2,right2,left3,right3) from those 8 bytes; c[j] = c0 | c1<<16.
struct ResizeGroup { int word, shift; uint32_t sel01, sel23; uint32_t c[4]; };
struct ResizeTables {      // per destination level, device pointers
  const ResizeTap* xt;     // [dw]
  const ResizeTap* yt;     // [dh]
  const ResizeGroup* xg;   // [(dw+3)/4], or nullptr when a group would span more than 8 source bytes
};

// level `l` (l>=1) of every frame from level l-1
void launch_resize(const uint8_t* src, long long srcFrameStride, int srcPitch, int sw, int sh,
                   uint8_t* dst, long long dstFrameStride, int dstPitch, int dw, int dh,
                   const ResizeTables& t, int nframes, cudaStream_t s);

struct PyramidPtrs {
  const uint8_t* lvl[kMaxLevels];   // frame-0 base of each level
  long long frameStride[kMaxLevels];
  int pitch[kMaxLevels];
};

// returns 0 or a cudaError_t
// levels [levelBegin, levelEnd) only (levelEnd < 0: all levels)
int launch_fast(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, const FastCellDesc* cells, uint32_t* cand, int* candCount,
                int* overflow, int nframes, cudaStream_t s, int levelBegin = 0, int levelEnd = -1);

// returns 0 or a cudaError_t
int launch_octree(const FrameSetGeom* dGeom, const FrameSetGeom& hGeom, const uint32_t* cand,
                  const int* candCount, uint16_t* nodeScratch, uint32_t* kept, int* keptCount,
                  int nframes, cudaStream_t s, int levelBegin = 0, int levelEnd = -1);

void launch_blur(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, uint8_t* blurArena, int nframes, cudaStream_t s);

void launch_describe(const PyramidPtrs& pyr, const uint8_t* blurArena, const FrameSetGeom& hGeom,
                     const uint32_t* kept, const int* keptCount, const int* candCount, orbx_keypoint* outKps,
                     uint8_t* outDesc, int cap, int* outN, int nframes, cudaStream_t s);

// stand-alone octree on one candidate set (orbx_distribute); returns 0 or a cudaError_t
int launch_octree_single(const uint32_t* cand, int n, int width, int height, int nfeat, int wCell, int hCell,
                         int nCols, uint16_t* nodeScratch, uint32_t* kept, int* keptCount, int keptCap,
                         cudaStream_t s);
size_t octree_smem_bytes(int cap);

void blur_tiles_for(int w, int h, int* tilesX, int* tilesY);
void launch_blur_single(const uint8_t* src, int srcPitch, uint8_t* dst, int dstPitch, int w, int h, cudaStream_t s);
void launch_atan2(const float* y, const float* x, float* out, long long n, cudaStream_t s);
void launch_sincosf(const float* a, float* s_out, float* c_out, long long n, cudaStream_t s);
void launch_sincosf_sweep(uint32_t lo, uint32_t hi, float* s_out, float* c_out, cudaStream_t s);

}  // namespace orbx
