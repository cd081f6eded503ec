// Per-cell FAST-9/16 detection with the iniThFAST / minThFAST fallback (SURVEY.md 8a rows E3, E4;
// R/src/ORBextractor.cpp:819-892 calling cv::FAST(cell, th, nonmaxSuppression=true)).
//
// Reformulation (equal to the reference; tests/test_gpu_extractor.py compares candidate sets and
// scores per level against the oracle):
//  * With ring pixels p_k (k = 0..15) and centre v, let A = min over the 16 arcs of 9 contiguous
//    ring pixels of (max of the arc) and B = max over arcs of (min of the arc). Then
//    m = max(v - A, B - v) is the largest margin by which some arc is uniformly darker/brighter
//    than the centre, the pixel is a corner at threshold t  <=>  m > t, and cv::FAST's
//    cornerScore<16> is m - 1 -- independent of t. One score map per cell serves both thresholds.
//  * cv::FAST's 3x3 non-max suppression only sees scores inside the cell's own scan window
//    [3,w-3)x[3,h-3) (everything else counts as 0): it is cell-local.
//  * The scan windows of neighbouring cells tile the level without overlap (cells overlap by
//    exactly the 6-pixel FAST apron), so every pixel is scored once.
//  * The order cv::FAST emits corners in only matters downstream as a tie-break (the octree keeps
//    the FIRST max-response point of a node); it is a function of (cell row, cell col, y, x), which
//    the octree kernel recomputes from the coordinates, so candidates are appended unordered.
//
// Kernel shape (v2): one CTA of 128 threads per cell. The cell patch is staged word-aligned in shared
// memory; every thread scores 4 horizontally adjacent pixels per task with branch-free packed
// 16x2 min/max (VIMNMX3.U16x2): 16 byte->16x2 unpacks and 80 three-input min/max per PAIR of
// pixels, no divergence, no per-pixel early-outs -- the cost per pixel is content independent
// (the v1 kernel was bound by the ALU pipe at 18/32 active lanes; see profiles/).
#include "orbx_common.cuh"

namespace orbx {

constexpr int kFastThreads = 128;
constexpr int kTileW = 19;                  // 32-bit words per staged row (cell <= 66 px + 7 px alignment slack)
constexpr int kTileRows = kMaxCellDim + 2;  // cell height <= 60 + 6
constexpr int kMapRows = kMaxCellDim;       // scan height <= 60, + one zero row above and below
constexpr int kOutCap = 1024;               // >= (60/2)^2 survivors of a 3x3 strict NMS

// bytes I, I+1 of the 12-byte window (w0,w1,w2) as two 16-bit lanes
template <int I>
__device__ __forceinline__ uint32_t ex2(uint32_t w0, uint32_t w1, uint32_t w2) {
  constexpr int wi = I >> 2, bi = I & 3;
  const uint32_t lo = wi == 0 ? w0 : (wi == 1 ? w1 : w2);
  if constexpr (bi < 3) {
    return __byte_perm(lo, 0u, bi | (4 << 4) | ((bi + 1) << 8) | (4 << 12));
  } else {
    const uint32_t hi = wi == 0 ? w1 : w2;
    return __byte_perm(lo, hi, 0x4433) & 0x00ff00ffu;
  }
}

// Two adjacent pixels (window bytes BASE, BASE+1 of the centre row): max(m - th, 0) in each 16-bit
// lane, where m is the corner margin defined above. W[j] = rows y-3+j of the 12-byte window.
template <int BASE>
__device__ __forceinline__ uint32_t fast_margin2(const uint32_t (&W)[7][3], uint32_t negThBias) {
#define ORBX_RING(dx, dy) ex2<BASE + (dx)>(W[(dy) + 3][0], W[(dy) + 3][1], W[(dy) + 3][2])
  uint32_t P[16];
  P[0] = ORBX_RING(0, 3);    P[1] = ORBX_RING(1, 3);    P[2] = ORBX_RING(2, 2);    P[3] = ORBX_RING(3, 1);
  P[4] = ORBX_RING(3, 0);    P[5] = ORBX_RING(3, -1);   P[6] = ORBX_RING(2, -2);   P[7] = ORBX_RING(1, -3);
  P[8] = ORBX_RING(0, -3);   P[9] = ORBX_RING(-1, -3);  P[10] = ORBX_RING(-2, -2); P[11] = ORBX_RING(-3, -1);
  P[12] = ORBX_RING(-3, 0);  P[13] = ORBX_RING(-3, 1);  P[14] = ORBX_RING(-2, 2);  P[15] = ORBX_RING(-1, 3);
  const uint32_t v = ORBX_RING(0, 0);
#undef ORBX_RING
  uint32_t T[16];
  // A = min over arcs of the arc maximum
#pragma unroll
  for (int k = 0; k < 16; k++) T[k] = __vimax3_u16x2(P[k], P[(k + 1) & 15], P[(k + 2) & 15]);
  uint32_t A = 0xffffffffu;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const uint32_t a0 = __vimax3_u16x2(T[k], T[(k + 3) & 15], T[(k + 6) & 15]);
    const uint32_t a1 = __vimax3_u16x2(T[k + 1], T[(k + 4) & 15], T[(k + 7) & 15]);
    A = __vimin3_u16x2(A, a0, a1);
  }
  // B = max over arcs of the arc minimum
#pragma unroll
  for (int k = 0; k < 16; k++) T[k] = __vimin3_u16x2(P[k], P[(k + 1) & 15], P[(k + 2) & 15]);
  uint32_t B = 0u;
#pragma unroll
  for (int k = 0; k < 16; k += 2) {
    const uint32_t b0 = __vimin3_u16x2(T[k], T[(k + 3) & 15], T[(k + 6) & 15]);
    const uint32_t b1 = __vimin3_u16x2(T[k + 1], T[(k + 4) & 15], T[(k + 7) & 15]);
    B = __vimax3_u16x2(B, b0, b1);
  }
  // both lanes biased by +256 so the plain 32-bit subtractions cannot borrow across lanes
  const uint32_t dark = v + 0x01000100u - A;
  const uint32_t bright = B + 0x01000100u - v;
  const uint32_t m = __vmaxu2(dark, bright);                // m + 256
  return __viaddmax_s16x2_relu(m, negThBias, 0u);           // max(m - th, 0)
}

__global__ void __launch_bounds__(kFastThreads)
fast_cells_kernel(const __grid_constant__ PyramidPtrs pyr, const __grid_constant__ FrameSetGeom g,
                  uint32_t* __restrict__ cand, int* __restrict__ candCount, int* __restrict__ overflow) {
  __shared__ uint32_t tile[kTileRows * kTileW];
  __shared__ uint32_t smap[(kMapRows + 2) * kTileW];
  __shared__ uint32_t outList[kOutCap];
  __shared__ int outCount, passCount, passCursor, outBase;

  const int nl = g.nlevels;
  int level = 0;
  const int cell = blockIdx.x;
  while (level + 1 < nl && cell >= g.lv[level + 1].cellBase) level++;
  const LevelGeom& L = g.lv[level];
  const int cidx = cell - L.cellBase;
  const int ci = cidx / L.nCols, cj = cidx - ci * L.nCols;
  const int f = blockIdx.y;

  // cell rectangle (R/src/ORBextractor.cpp:851-870)
  const int iniY = kMinBorder + ci * L.hCell;
  const int iniX = kMinBorder + cj * L.wCell;
  if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6) return;
  const int maxY = min(iniY + L.hCell + 6, L.maxBY);
  const int maxX = min(iniX + L.wCell + 6, L.maxBX);
  const int ch = maxY - iniY;
  const int sw = maxX - iniX - 6, sh = ch - 6;  // cv::FAST scans [3,cw-3)x[3,ch-3)
  if (sw <= 0 || sh <= 0) return;

  // Tile column 0 is global column gx0 (a multiple of 4, <= iniX-1): the first scanned pixel sits at
  // tile column a in 4..7, so every 4-pixel group starts on a word boundary and its ring (+-3) stays
  // inside the 12-byte window of words q, q+1, q+2.
  const int gx0 = ((iniX + 3) & ~3) - 4;
  const int a = iniX + 3 - gx0;
  const int b = a + sw;
  const int nG = (b - 1) >> 2;  // groups q = 0..nG-1 cover tile words 1..nG
  const int nW = nG + 2;        // staged words per row (<= kTileW)

  const int tid = threadIdx.x;
  const int pitch = pyr.pitch[level];
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level] + (long long)iniY * pitch + gx0;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)pitch) & 3) == 0;
  {
    int r = tid / nW, c = tid - r * nW;
    const int dr = kFastThreads / nW, dc = kFastThreads - dr * nW;
    while (r < ch) {
      const uint8_t* p = src + (long long)r * pitch + 4 * c;
      uint32_t w;
      if (aligned) {
        w = *reinterpret_cast<const uint32_t*>(p);
      } else {
        w = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
      }
      tile[r * kTileW + c] = w;
      r += dr; c += dc;
      if (c >= nW) { c -= nW; r++; }
    }
  }
  for (int i = tid; i < (sh + 2) * kTileW; i += kFastThreads) smap[i] = 0u;
  if (tid == 0) { outCount = 0; passCount = 0; passCursor = 0; }
  __syncthreads();

  const int minTh = g.minTh, iniTh = g.iniTh;
  // per 16-bit lane: -(256 + minTh)
  const uint32_t negTh = (uint32_t)(-(256 + minTh)) & 0xffffu;
  const uint32_t negThBias = negTh | (negTh << 16);
  const int nTasks = sh * nG;
  const int dy = kFastThreads / nG, dq = kFastThreads - dy * nG;

  // ---- score map: max(m - minTh, 0) for every scanned pixel (0 = not a corner at minTh) ----
  {
    int y = tid / nG, q = tid - y * nG;
    for (int t = tid; t < nTasks; t += kFastThreads) {
      const uint32_t* tp = &tile[y * kTileW + q];
      uint32_t W[7][3];
#pragma unroll
      for (int j = 0; j < 7; j++) {
        W[j][0] = tp[j * kTileW];
        W[j][1] = tp[j * kTileW + 1];
        W[j][2] = tp[j * kTileW + 2];
      }
      const uint32_t u01 = fast_margin2<4>(W, negThBias);
      const uint32_t u23 = fast_margin2<6>(W, negThBias);
      uint32_t word = __byte_perm(u01, u23, 0x6420);
      const int col0 = 4 * (q + 1);
      const int skipLo = a - col0, skipHi = col0 + 4 - b;
      if (skipLo > 0) word &= 0xffffffffu << (8 * skipLo);
      if (skipHi > 0) word &= 0xffffffffu >> (8 * skipHi);
      smap[(y + 1) * kTileW + q + 1] = word;
      y += dy; q += dq;
      if (q >= nG) { q -= nG; y++; }
    }
  }
  __syncthreads();

  // ---- cell-local 3x3 strict non-max suppression; survivors go to a shared list ----
  int anyIni = 0;
  {
    const uint8_t* sb = reinterpret_cast<const uint8_t*>(smap);
    constexpr int S = kTileW * 4;
    int y = tid / nG, q = tid - y * nG;
    for (int t = tid; t < nTasks; t += kFastThreads) {
      const uint32_t word = smap[(y + 1) * kTileW + q + 1];
      if (word) {
#pragma unroll
        for (int j = 0; j < 4; j++) {
          const int s = (word >> (8 * j)) & 0xff;
          if (s) {
            const int col = 4 * (q + 1) + j;
            const uint8_t* c = sb + (y + 1) * S + col;
            const bool keep = s > c[-S - 1] && s > c[-S] && s > c[-S + 1] && s > c[-1] && s > c[1] &&
                              s > c[S - 1] && s > c[S] && s > c[S + 1];
            if (keep) {
              const int score = s + minTh - 1;  // cv::FAST's cornerScore
              const int slot = atomicAdd(&outCount, 1);
              // coordinates relative to minBorder: cell-local (x+3, y+3) + (j*wCell, i*hCell)  (:889-890)
              if (slot < kOutCap) outList[slot] = pack_cand(col - a + 3 + cj * L.wCell, y + 3 + ci * L.hCell, score);
              anyIni |= score >= iniTh;
            }
          }
        }
      }
      y += dy; q += dq;
      if (q >= nG) { q -= nG; y++; }
    }
  }
  // "cell empty after NMS at iniThFAST?" (:875-883) is a block-wide vote
  const int useIni = __syncthreads_or(anyIni);
  const int th = useIni ? iniTh : minTh;
  const int n = min(outCount, kOutCap);
  if (n == 0) return;

  int mine = 0;
  for (int i = tid; i < n; i += kFastThreads) mine += cand_score(outList[i]) >= th;
  if (mine) atomicAdd(&passCount, mine);
  __syncthreads();
  const int total = passCount;
  const long long slotBase = L.candOff + (long long)f * L.candCap;
  if (tid == 0) outBase = atomicAdd(&candCount[f * nl + level], total);
  __syncthreads();
  const int base = outBase;
  if (base + total > L.candCap) {
    if (tid == 0) atomicExch(overflow, 1);
  }
  for (int i = tid; i < n; i += kFastThreads) {
    const uint32_t c = outList[i];
    if (cand_score(c) >= th) {
      const int slot = base + atomicAdd(&passCursor, 1);
      if (slot < L.candCap) cand[slotBase + slot] = c;
    }
  }
}

void launch_fast(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, uint32_t* cand, int* candCount, int* overflow,
                 int nframes, cudaStream_t s) {
  dim3 grid(hGeom.totalCells, nframes);
  fast_cells_kernel<<<grid, kFastThreads, 0, s>>>(pyr, hGeom, cand, candCount, overflow);
}

}  // namespace orbx
