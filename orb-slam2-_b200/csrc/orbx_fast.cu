// Per-cell FAST-9/16 detection with the iniThFAST / minThFAST fallback (SURVEY.md 8a rows E3, E4;
// R/src/ORBextractor.cpp:819-892 calling cv::FAST(cell, th, nonmaxSuppression=true)).
//
// Reformulation used here (proved equal to the reference by tests/test_fast_parity.py):
//  * cv::FAST's score of a corner is S(p) = max over the 16 arcs of 9 contiguous ring pixels of
//    min |v - ring| (sign-consistent) minus 1, independent of the detection threshold, and a pixel
//    is a corner at threshold t  <=>  S(p) >= t. So one score map per cell serves both thresholds.
//  * cv::FAST's 3x3 non-max suppression only sees scores inside the cell's own scan window
//    [3,w-3)x[3,h-3) (everything else counts as 0), i.e. it is cell-local: keep p at threshold t
//    <=> S(p) >= t and S(p) > S(q) for the 8 neighbours q inside the window.
//  * The scan windows of neighbouring cells tile the level without overlap (cells overlap by
//    exactly the 6-pixel FAST apron), so every pixel is scored once.
//  * The order cv::FAST emits corners in only matters downstream as a tie-break (octree keeps the
//    FIRST max-response point of a node). The order is a function of (cell row, cell col, y, x), which
//    the octree kernel recomputes from the coordinates, so candidates are appended unordered.
//
// One CTA per cell: the (wCell+6)x(hCell+6) patch is staged in shared memory once, scored, suppressed,
// and the threshold decision ("cell empty after NMS at iniThFAST?") is a block-wide vote.
#include "orbx_common.cuh"

namespace orbx {

constexpr int kTileStride = 72;                      // bytes per staged row (cells are < 66 wide)
constexpr int kTileRows = kMaxCellDim + 6;           // 70
constexpr int kFastThreads = 256;

__device__ __forceinline__ bool has_run9(uint32_t m) {
  const uint32_t mm = m | (m << 16);
  uint32_t x = mm & (mm >> 1);
  x &= x >> 2;
  x &= x >> 4;
  x &= mm >> 8;
  return (x & 0xffffu) != 0;
}

// Score of the pixel at c (row stride `st`) if it is a FAST-9/16 corner at threshold th, else 0.
__device__ __forceinline__ int fast_score_at(const uint8_t* c, int st, int th) {
  const int v = c[0];
  // quick reject: any arc of 9 contains one pixel of each opposite pair
  const int lo = v - th, hi = v + th;
  const int p0 = c[3 * st], p8 = c[-3 * st];
  if (!((p0 < lo) | (p0 > hi) | (p8 < lo) | (p8 > hi))) return 0;
  const int p4 = c[3], p12 = c[-3];
  if (!((p4 < lo) | (p4 > hi) | (p12 < lo) | (p12 > hi))) return 0;

  int p[16];
  p[0] = p0; p[8] = p8; p[4] = p4; p[12] = p12;
  p[1] = c[3 * st + 1];  p[2] = c[2 * st + 2];   p[3] = c[st + 3];
  p[5] = c[-st + 3];     p[6] = c[-2 * st + 2];  p[7] = c[-3 * st + 1];
  p[9] = c[-3 * st - 1]; p[10] = c[-2 * st - 2]; p[11] = c[-st - 3];
  p[13] = c[st - 3];     p[14] = c[2 * st - 2];  p[15] = c[3 * st - 1];

  uint32_t md = 0, mb = 0;
#pragma unroll
  for (int k = 0; k < 16; k++) {
    md |= (uint32_t)(p[k] < lo) << k;
    mb |= (uint32_t)(p[k] > hi) << k;
  }
  const bool dark = has_run9(md), bright = has_run9(mb);
  if (!(dark | bright)) return 0;

  int e[16];
#pragma unroll
  for (int k = 0; k < 16; k++) e[k] = dark ? (v - p[k]) : (p[k] - v);
  int m2[16], m4[16], m8[16];
#pragma unroll
  for (int k = 0; k < 16; k++) m2[k] = min(e[k], e[(k + 1) & 15]);
#pragma unroll
  for (int k = 0; k < 16; k++) m4[k] = min(m2[k], m2[(k + 2) & 15]);
#pragma unroll
  for (int k = 0; k < 16; k++) m8[k] = min(m4[k], m4[(k + 4) & 15]);
  int best = -255;
#pragma unroll
  for (int k = 0; k < 16; k++) best = max(best, min(m8[k], e[(k + 8) & 15]));
  return best - 1;
}

__global__ void __launch_bounds__(kFastThreads)
fast_cells_kernel(PyramidPtrs pyr, const FrameSetGeom* __restrict__ g, uint32_t* __restrict__ cand,
                  int* __restrict__ candCount, int* __restrict__ overflow) {
  __shared__ __align__(16) uint8_t tile[kTileRows * kTileStride];
  __shared__ uint8_t score[kMaxCellDim * kMaxCellDim];
  __shared__ uint8_t keep[kMaxCellDim * kMaxCellDim];
  __shared__ uint32_t outList[kMaxCellDim * kMaxCellDim / 4 + 64];
  __shared__ int outCount, outBase;

  const int nl = g->nlevels;
  int level = 0;
  const int cell = blockIdx.x;
  while (level + 1 < nl && cell >= g->lv[level + 1].cellBase) level++;
  const LevelGeom& L = g->lv[level];
  const int cidx = cell - L.cellBase;
  const int ci = cidx / L.nCols, cj = cidx - ci * L.nCols;
  const int f = blockIdx.y;

  // cell rectangle (R/src/ORBextractor.cpp:851-870)
  const int iniY = kMinBorder + ci * L.hCell;
  const int iniX = kMinBorder + cj * L.wCell;
  if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6) return;
  const int maxY = min(iniY + L.hCell + 6, L.maxBY);
  const int maxX = min(iniX + L.wCell + 6, L.maxBX);
  const int cw = maxX - iniX, ch = maxY - iniY;
  const int sw = cw - 6, sh = ch - 6;  // cv::FAST scans [3,cw-3)x[3,ch-3)
  if (sw <= 0 || sh <= 0) return;

  const int tid = threadIdx.x;
  const int pitch = pyr.pitch[level];
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level] + (long long)iniY * pitch + iniX;
  for (int i = tid; i < ch * cw; i += kFastThreads) {
    const int r = i / cw, c = i - r * cw;
    tile[r * kTileStride + c] = src[(long long)r * pitch + c];
  }
  if (tid == 0) outCount = 0;
  __syncthreads();

  const int minTh = g->minTh, iniTh = g->iniTh;
  const int npx = sw * sh;
  for (int i = tid; i < npx; i += kFastThreads) {
    const int y = i / sw, x = i - y * sw;
    score[i] = (uint8_t)fast_score_at(&tile[(y + 3) * kTileStride + x + 3], kTileStride, minTh);
  }
  __syncthreads();

  int anyIni = 0;
  for (int i = tid; i < npx; i += kFastThreads) {
    const int s = score[i];
    int k = 0;
    if (s > 0) {
      const int y = i / sw, x = i - y * sw;
      k = 1;
#pragma unroll
      for (int dy = -1; dy <= 1; dy++) {
#pragma unroll
        for (int dx = -1; dx <= 1; dx++) {
          if (dx == 0 && dy == 0) continue;
          const int yy = y + dy, xx = x + dx;
          const int q = (yy >= 0 && yy < sh && xx >= 0 && xx < sw) ? score[yy * sw + xx] : 0;
          k &= (s > q);
        }
      }
      anyIni |= (k && s >= iniTh);
    }
    keep[i] = (uint8_t)k;
  }
  const int useIni = __syncthreads_or(anyIni);  // also orders keep[] writes before the reads below
  const int th = useIni ? iniTh : minTh;

  for (int i = tid; i < npx; i += kFastThreads) {
    if (keep[i] && score[i] >= th) {
      const int y = i / sw, x = i - y * sw;
      const int slot = atomicAdd(&outCount, 1);
      // coordinates relative to minBorder: cell-local (x+3, y+3) + (j*wCell, i*hCell)  (:889-890)
      outList[slot] = pack_cand(x + 3 + cj * L.wCell, y + 3 + ci * L.hCell, score[i]);
    }
  }
  __syncthreads();
  const int n = outCount;
  if (n == 0) return;
  const long long slotBase = L.candOff + (long long)f * L.candCap;
  if (tid == 0) outBase = atomicAdd(&candCount[f * nl + level], n);
  __syncthreads();
  const int base = outBase;
  if (base + n > L.candCap) {
    if (tid == 0) atomicExch(overflow, 1);
  }
  for (int i = tid; i < n; i += kFastThreads)
    if (base + i < L.candCap) cand[slotBase + base + i] = outList[i];
}

void launch_fast(const PyramidPtrs& pyr, const FrameSetGeom* dGeom, const FrameSetGeom& hGeom,
                 uint32_t* cand, int* candCount, int* overflow, int nframes, cudaStream_t s) {
  dim3 grid(hGeom.totalCells, nframes);
  fast_cells_kernel<<<grid, kFastThreads, 0, s>>>(pyr, dGeom, cand, candCount, overflow);
}

}  // namespace orbx
