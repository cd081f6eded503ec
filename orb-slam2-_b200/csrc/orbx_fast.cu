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
// Kernel shape (v3): one WARP per cell, shared memory sized on the host for the largest cell of the shape. The cell patch is staged word-aligned in shared memory; every lane scores
// 4 horizontally adjacent pixels per task with branch-free packed 16x2 min/max (VIMNMX3.U16x2):
// the patch is staged as 16-bit lanes, so 6 of the 16 ring offsets are plain words and 10 cost one PRMT;
// 72 packed min/max per PAIR of pixels, no divergence, no per-pixel early-outs -- the cost per pixel
// of the scoring pass is content independent. (v1 was one thread per pixel with
// early-outs: ALU-pipe bound at 18/32 active lanes. v2 was one 128-thread CTA per cell: 80 % lane use
// because sh*nG tasks rarely fill whole rounds of 128; with 32-lane rounds the tail is ~5 %.)
#include <type_traits>

#include "orbx_common.cuh"

namespace orbx {

// One warp per CTA: the warps of a CTA finish their cells at different times and a CTA's slot is only handed on when
// its last warp is done; with single-warp CTAs (32 resident per SM at 64 registers) the SM stays full. Measured on
// C1, 512 frames: 8 / 4 / 2 / 1 warps per CTA = 1.154 / 1.133 / 1.133 / 1.120 ms (register caps below 64 spill and lose).
#ifndef ORBX_FAST_WARPS
#define ORBX_FAST_WARPS 1
#endif
#ifndef ORBX_FAST_MINB
#define ORBX_FAST_MINB 32
#endif
constexpr int kFastWarps = ORBX_FAST_WARPS;   // cells per CTA
constexpr int kFastThreads = kFastWarps * 32;

// cInv16[n] = 65536/n + 1: (i * cInv16[n]) >> 16 == i / n for n <= 64, i <= 1023 (and for i <= 4096 when n <= 26);
// the kernel has no integer division on its path.
struct Inv16Table { uint32_t v[65]; };  // v[1] = 65537 needs 17 bits
constexpr Inv16Table make_inv16() {
  Inv16Table t{};
  t.v[0] = 0;
  for (int n = 1; n <= 64; n++) t.v[n] = (uint32_t)(65536 / n + 1);
  return t;
}
__constant__ Inv16Table cInv16 = make_inv16();
__device__ __forceinline__ int div_small(int i, int n) {  // see cInv16
  return (int)(((uint32_t)i * cInv16.v[n]) >> 16);
}

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

// Pixels C, C+1 of a row staged as 16-bit lanes (W[k] = pixels 2k, 2k+1): free when C is even, one
// PRMT when it straddles two words.
template <int C>
__device__ __forceinline__ uint32_t pair16(const uint32_t (&W)[6]) {
  if constexpr ((C & 1) == 0) return W[C >> 1];
  else return __byte_perm(W[C >> 1], W[(C >> 1) + 1], 0x5432);
}

// Two adjacent pixels (window columns BASE, BASE+1 of the centre row, BASE even): max(m - th, 0) in each
// 16-bit lane, where m is the corner margin defined above. W[j] = rows y-3+j of the 12-pixel window.
// 6 of the 16 ring offsets have an even dx and come straight out of the staged words.
template <int BASE>
__device__ __forceinline__ uint32_t fast_margin2(const uint32_t (&W)[7][6], uint32_t negThBias) {
#define ORBX_RING(dx, dy) pair16<BASE + (dx)>(W[(dy) + 3])
  uint32_t P[16];
  P[0] = ORBX_RING(0, 3);    P[1] = ORBX_RING(1, 3);    P[2] = ORBX_RING(2, 2);    P[3] = ORBX_RING(3, 1);
  P[4] = ORBX_RING(3, 0);    P[5] = ORBX_RING(3, -1);   P[6] = ORBX_RING(2, -2);   P[7] = ORBX_RING(1, -3);
  P[8] = ORBX_RING(0, -3);   P[9] = ORBX_RING(-1, -3);  P[10] = ORBX_RING(-2, -2); P[11] = ORBX_RING(-3, -1);
  P[12] = ORBX_RING(-3, 0);  P[13] = ORBX_RING(-3, 1);  P[14] = ORBX_RING(-2, 2);  P[15] = ORBX_RING(-1, 3);
  const uint32_t v = ORBX_RING(0, 0);
#undef ORBX_RING
  // A = min over the 16 arcs of the arc maximum. Arcs i and i+1 (i even) share p[i+1..i+8] =: M, and
  // min(max(M, p[i]), max(M, p[i+9])) = max(M, min(p[i], p[i+9])): 8 pair terms instead of 16 arcs.
  // U[j] = p[2j+1..2j+2], V[j] = p[2j+1..2j+4], M of arc pair m = V[m] and V[m+2]: 36 VIMNMX per side.
  uint32_t U[8], V[8];
#pragma unroll
  for (int j = 0; j < 8; j++) U[j] = __vmaxu2(P[2 * j + 1], P[(2 * j + 2) & 15]);
#pragma unroll
  for (int j = 0; j < 8; j++) V[j] = __vmaxu2(U[j], U[(j + 1) & 7]);
  uint32_t A = 0xffffffffu;
#pragma unroll
  for (int m = 0; m < 8; m += 2) {
    const uint32_t x0 = __vimax3_u16x2(V[m], V[(m + 2) & 7], __vminu2(P[2 * m], P[(2 * m + 9) & 15]));
    const uint32_t x1 = __vimax3_u16x2(V[m + 1], V[(m + 3) & 7], __vminu2(P[2 * m + 2], P[(2 * m + 11) & 15]));
    A = __vimin3_u16x2(A, x0, x1);
  }
  // B = max over arcs of the arc minimum, the same with min and max exchanged
#pragma unroll
  for (int j = 0; j < 8; j++) U[j] = __vminu2(P[2 * j + 1], P[(2 * j + 2) & 15]);
#pragma unroll
  for (int j = 0; j < 8; j++) V[j] = __vminu2(U[j], U[(j + 1) & 7]);
  uint32_t B = 0u;
#pragma unroll
  for (int m = 0; m < 8; m += 2) {
    const uint32_t x0 = __vimin3_u16x2(V[m], V[(m + 2) & 7], __vmaxu2(P[2 * m], P[(2 * m + 9) & 15]));
    const uint32_t x1 = __vimin3_u16x2(V[m + 1], V[(m + 3) & 7], __vmaxu2(P[2 * m + 2], P[(2 * m + 11) & 15]));
    B = __vimax3_u16x2(B, x0, x1);
  }
  // both lanes biased by +256 so the plain 32-bit subtractions cannot borrow across lanes
  const uint32_t dark = v + 0x01000100u - A;
  const uint32_t bright = B + 0x01000100u - v;
  const uint32_t m = __vmaxu2(dark, bright);                // m + 256
  return __viaddmax_s16x2_relu(m, negThBias, 0u);           // max(m - th, 0)
}

// Upper bound of the margin of two adjacent pixels from the four compass ring pixels (n = (0,3), s = (0,-3),
// e = (3,0), w = (-3,0)) alone: max(bound - th, 0) per 16-bit lane. See pass 1 of the kernel. The minimum over
// the four adjacent compass pairs of the pair maximum is, by distributivity, max(min(n,s), min(e,w)): 3 packed
// operations per side instead of 6.
__device__ __forceinline__ uint32_t fast_compass_bound(uint32_t n, uint32_t s, uint32_t e, uint32_t w, uint32_t v,
                                                       uint32_t negThBias) {
  const uint32_t Ac = __vmaxu2(__vminu2(n, s), __vminu2(e, w));
  const uint32_t Bc = __vminu2(__vmaxu2(n, s), __vmaxu2(e, w));
  const uint32_t dark = v + 0x01000100u - Ac;
  const uint32_t bright = Bc + 0x01000100u - v;
  return __viaddmax_s16x2_relu(__vmaxu2(dark, bright), negThBias, 0u);
}

// words of shared memory one cell needs: patch staged as 16-bit lanes (ch x 2nW) + score map ((sh+2) x nW)
// + the queue of tasks with a non-zero score word (16 bits per task); even, so that every warp's patch
// stays 8-byte aligned
__host__ __device__ inline int fast_cell_words(int wCell, int hCell) {
  const int nW = ((7 + wCell - 1) >> 2) + 2;  // worst alignment: a = 7
  const int rounds = (hCell * (nW - 2) + 31) / 32;  // one class-B ballot word per pass-1 round
  const int words = (hCell + 6) * 2 * nW + (hCell + 2) * nW + (hCell * (nW - 2) + 1) / 2 + 1 + rounds;
  return (words + 1) & ~1;
}

__global__ void __launch_bounds__(kFastThreads, ORBX_FAST_MINB)
fast_cells_kernel(const __grid_constant__ PyramidPtrs pyr, const __grid_constant__ FrameSetGeom g,
                  const FastCellDesc* __restrict__ cells, uint32_t* __restrict__ cand, int* __restrict__ candCount,
                  int* __restrict__ overflow, int wordsPerWarp, int cellBegin, int cellEnd) {
  extern __shared__ __align__(16) uint32_t smemFast[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int cell = cellBegin + blockIdx.x * kFastWarps + warp;
  if (cell >= cellEnd) return;  // warps are independent: no block-wide barrier below
  // the cell's rectangle comes from the table of the frame shape (host-built, the same for every frame)
  const uint4 cd = __ldg(reinterpret_cast<const uint4*>(cells) + cell);
  const int nl = g.nlevels;
  const int level = cd.x & 0xff, ch = (cd.x >> 8) & 0xff, a = (cd.x >> 16) & 0xff, b = cd.x >> 24;
  const int iniY = cd.y & 0xffff, gx0 = cd.y >> 16;
  const int outX0 = (int)(short)(cd.z & 0xffff), outY0 = (int)(short)(cd.z >> 16);
  const LevelGeom& L = g.lv[level];
  const int f = blockIdx.y;
  const int sh = ch - 6;        // cv::FAST scans [3,cw-3)x[3,ch-3)
  const int nG = (b - 1) >> 2;  // groups q = 0..nG-1 cover tile words 1..nG
  const int nW = nG + 2;        // staged words per row
  pdl_wait();

  uint32_t* tile = smemFast + warp * wordsPerWarp;  // ch rows x 2nW words: one 16-bit lane per pixel
  uint32_t* smap = tile + ch * 2 * nW;              // (sh+2) rows x nW words, rows 0 and sh+1 stay zero
  uint16_t* queue = reinterpret_cast<uint16_t*>(smap + (sh + 2) * nW);  // y<<8 | q of the tasks worth a second look

  const int pitch = pyr.pitch[level];
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level] + (long long)iniY * pitch + gx0;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)pitch) & 3) == 0;
  {
    // Word idx of the patch (row idx / nW, word idx % nW) goes to tile[2 idx]; lane handles idx = lane, lane+32, ...
    // Row and column advance incrementally (32 = dr * nW + dc per step), so a word costs a pointer add, the load,
    // two PRMT and one 8-byte store. Eight loads are in flight per lane before the first one is used: with one
    // load per loop trip the warp sat out a full DRAM latency 13 times per cell.
    const int total = ch * nW;
    const int dr = div_small(32, nW), dc = 32 - dr * nW;
    int r = div_small(lane, nW);
    int c = lane - r * nW;
    const uint8_t* p = src + (long long)r * pitch + 4 * c;
    const long long stepNoWrap = (long long)dr * pitch + 4 * dc;
    const long long stepWrap = stepNoWrap + pitch - 4 * nW;
    constexpr int kInFlight = 8;
    auto stage = [&](auto alignedTag) {
      constexpr bool kAligned = decltype(alignedTag)::value;
      for (int base = lane; base < total; base += 32 * kInFlight) {
        uint32_t w[kInFlight];
#pragma unroll
        for (int k = 0; k < kInFlight; k++) {
          w[k] = 0u;
          if (base + 32 * k < total) {
            if constexpr (kAligned) {
              w[k] = *reinterpret_cast<const uint32_t*>(p);
            } else {
              w[k] = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
            }
          }
          c += dc;
          const bool wrap = c >= nW;
          p += wrap ? stepWrap : stepNoWrap;
          if (wrap) c -= nW;
        }
#pragma unroll
        for (int k = 0; k < kInFlight; k++) {
          // bytes -> 16-bit lanes once here instead of once per ring offset in the scoring loop
          if (base + 32 * k < total)
            *reinterpret_cast<uint2*>(&tile[2 * (base + 32 * k)]) =
                make_uint2(__byte_perm(w[k], 0u, 0x4140), __byte_perm(w[k], 0u, 0x4342));
        }
      }
    };
    if (aligned) stage(std::true_type{});
    else stage(std::false_type{});
  }
  for (int i = lane; i < (sh + 2) * nW; i += 32) smap[i] = 0u;
  __syncwarp();

  const int minTh = g.minTh, iniTh = g.iniTh;
  // per 16-bit lane: -(256 + minTh)
  const uint32_t negTh = (uint32_t)(-(256 + minTh)) & 0xffffu;
  const uint32_t negThBias = negTh | (negTh << 16);
  // Scores are kept as max(m - minTh, 0): a pixel whose margin exceeds iniTh stores at least iniByte.
  const int iniByte = max(iniTh - minTh + 1, 1);
  const uint32_t negIni = ((uint32_t)(-(iniByte - 1)) & 0xffffu) * 0x00010001u;
  const int nTasks = sh * nG;
  const int dy = div_small(32, nG), dq = 32 - dy * nG;
  const int y0 = div_small(lane, nG), q0 = lane - y0 * nG;
  const uint32_t below = (1u << lane) - 1u;

  // ---- pass 1, every task: compass bound. Any 9-arc of the ring contains two adjacent compass pixels
  //      (p0, p4, p8, p12), so  A >= Ac = min over adjacent compass pairs of their max  and
  //      B <= Bc = max over those pairs of their min, hence m <= max(v - Ac, Bc - v): a task whose four
  //      bounds are all <= minTh has an all-zero score word and is dropped after 12 instead of 92 packed
  //      operations per pixel pair. The others fall in two classes: list A (some bound above iniTh) is queued as
  //      y<<8 | q from the front of the queue; class B (bounds in (minTh, iniTh]) is only ever scored if the cell
  //      turns out empty at iniTh (phase 1 below) and is kept as ballot words until then. Rounds are warp-uniform
  //      (lanes past the last task idle inside the round) so that the ballots are full-warp ones. ----
  int nA = 0;
  bool anyB = false;
  uint32_t* bmask = reinterpret_cast<uint32_t*>(queue + ((nTasks + 1) & ~1));  // class-B ballots, one word per round
  {
    int y = y0, q = q0;
    for (int t0 = 0; t0 < nTasks; t0 += 32) {
      uint32_t any = 0u, hi = 0u;
      if (t0 + lane < nTasks) {
        const uint2* tp = reinterpret_cast<const uint2*>(&tile[(y * nW + q) * 2]);  // window = tile columns 4q..4q+11
        const uint2 top = tp[1], bot = tp[6 * nW + 1];                             // columns 4..7 of rows y-3, y+3
        const uint2 c01 = tp[3 * nW], c23 = tp[3 * nW + 1], c45 = tp[3 * nW + 2];
        const uint32_t C[6] = {c01.x, c01.y, c23.x, c23.y, c45.x, c45.y};
        const uint32_t b01 = fast_compass_bound(top.x, bot.x, pair16<7>(C), pair16<1>(C), C[2], negThBias);
        const uint32_t b23 = fast_compass_bound(top.y, bot.y, pair16<9>(C), pair16<3>(C), C[3], negThBias);
        any = b01 | b23;
        hi = __viaddmax_s16x2_relu(__vmaxu2(b01, b23), negIni, 0u);
      }
      // Class A is compacted right away; class B is only remembered as the round's ballot word: its list is needed
      // by a few per cent of the cells (those empty at iniTh) and is rebuilt from the ballots there.
      const uint32_t maskA = __ballot_sync(0xffffffffu, hi != 0u);
      const uint32_t maskB = __ballot_sync(0xffffffffu, any != 0u) & ~maskA;
      if (hi != 0u) queue[nA + __popc(maskA & below)] = (uint16_t)((y << 8) | q);
      if (lane == 0) bmask[t0 >> 5] = maskB;
      nA += __popc(maskA);
      anyB = anyB || maskB != 0u;
      y += dy; q += dq;
      if (q >= nG) { q -= nG; y++; }
    }
  }
  __syncwarp();
  if (nA == 0 && !anyB) return;
  int nB = 0;

  // Phase 0 scores list A and suppresses at iniTh. A pixel above iniTh can only lose to a neighbour whose margin
  // is also above iniTh, and such a neighbour's task is in A: the phase-0 result is exactly cv::FAST(cell, iniTh).
  // If it is empty (:875-883), phase 1 scores list B as well and suppresses everything at minTh.
  uint32_t* kept = tile;  // survivors' score words; only written once the patch is dead (see below)
  int nQ = 0;             // tasks whose score word holds a corner at minTh: queue[0 .. nQ)
  int nS = 0;             // tasks that keep a corner after suppression: queue[0 .. nS), kept[0 .. nS)
#pragma unroll 1
  for (int phase = 0; phase < 2; phase++) {
    if (phase == 1) {
      // list B from the ballots, at the back of the queue (task t of round r is r*32 + lane = y*nG + q)
      int back = nTasks - 1;
      for (int t0 = 0; t0 < nTasks; t0 += 32) {
        const uint32_t mb = bmask[t0 >> 5];
        if ((mb >> lane) & 1u) {
          const int t = t0 + lane, y = div_small(t, nG), q = t - y * nG;
          queue[back - __popc(mb & below)] = (uint16_t)((y << 8) | q);
        }
        back -= __popc(mb);
      }
      nB = nTasks - 1 - back;
      __syncwarp();
    }
    const int srcBase = phase == 0 ? 0 : nTasks - nB;
    const int srcN = phase == 0 ? nA : nB;
    // ---- pass 2, listed tasks: score map max(m - minTh, 0) (0 = not a corner at minTh). Only a task whose
    //      word holds a corner can keep one after suppression: those are appended to queue[0 .. nQ). Entry j of
    //      a round is read before the round's ballot and nQ never overtakes the read position (phase 0: nQ <=
    //      entries read; phase 1: nQ <= nA + entries of B read <= nTasks - nB + entries read). ----
    for (int j0 = 0; j0 < srcN; j0 += 32) {
      uint32_t word = 0u;
      int e = 0;
      if (j0 + lane < srcN) {
        e = queue[srcBase + j0 + lane];
        const int y = e >> 8, q = e & 0xff;
        const uint2* tp = reinterpret_cast<const uint2*>(&tile[(y * nW + q) * 2]);
        uint32_t W[7][6];
#pragma unroll
        for (int j = 0; j < 7; j++) {
          const uint2 w01 = tp[j * nW], w23 = tp[j * nW + 1], w45 = tp[j * nW + 2];
          W[j][0] = w01.x; W[j][1] = w01.y; W[j][2] = w23.x; W[j][3] = w23.y; W[j][4] = w45.x; W[j][5] = w45.y;
        }
        const uint32_t u01 = fast_margin2<4>(W, negThBias);
        const uint32_t u23 = fast_margin2<6>(W, negThBias);
        word = __byte_perm(u01, u23, 0x6420);
        const int col0 = 4 * (q + 1);
        const int skipLo = a - col0, skipHi = col0 + 4 - b;
        if (skipLo > 0) word &= 0xffffffffu << (8 * skipLo);
        if (skipHi > 0) word &= 0xffffffffu >> (8 * skipHi);
        smap[(y + 1) * nW + q + 1] = word;
      }
      const uint32_t liveMask = __ballot_sync(0xffffffffu, word != 0u);
      if (word != 0u) queue[nQ + __popc(liveMask & (below))] = (uint16_t)e;
      nQ += __popc(liveMask);
    }
    __syncwarp();

    // ---- cell-local 3x3 strict non-max suppression, packed: the 8-neighbour maximum of two adjacent pixels is
    //      4 VIMNMX on 16x2 lanes. Scores below this phase's threshold are cleared and the tasks that keep a
    //      corner are compacted in place once more, their score words going to kept[], which reuses the patch
    //      rows: a word is only stored when a survivor exists, i.e. when no further phase needs the patch (and
    //      then nothing was compacted either: queue[0 .. nQ) is still whole for phase 1). ----
    const uint32_t negKeep = phase == 0 ? negIni : 0u;
    for (int j0 = 0; j0 < nQ; j0 += 32) {
      uint32_t word = 0u;
      int e = 0;
      if (j0 + lane < nQ) {
        e = queue[j0 + lane];
        const int y = e >> 8, q = e & 0xff;
        const uint32_t* mp = &smap[y * nW + q];  // map rows y, y+1, y+2 = image rows y-1, y, y+1
        uint32_t E[3][5];                        // column pairs (3,4) (4,5) (5,6) (6,7) (7,8) of the 12-byte window
#pragma unroll
        for (int r = 0; r < 3; r++) {
          const uint32_t w0 = mp[r * nW], w1 = mp[r * nW + 1], w2 = mp[r * nW + 2];
          E[r][0] = ex2<3>(w0, w1, w2);
          E[r][1] = ex2<4>(w0, w1, w2);
          E[r][2] = ex2<5>(w0, w1, w2);
          E[r][3] = ex2<6>(w0, w1, w2);
          E[r][4] = ex2<7>(w0, w1, w2);
        }
        uint32_t k2[2];
#pragma unroll
        for (int p = 0; p < 2; p++) {  // pixel pairs (0,1) and (2,3): left/centre/right pairs E[.][2p .. 2p+2]
          const uint32_t up = __vimax3_u16x2(E[0][2 * p], E[0][2 * p + 1], E[0][2 * p + 2]);
          const uint32_t dn = __vimax3_u16x2(E[2][2 * p], E[2][2 * p + 1], E[2][2 * p + 2]);
          const uint32_t M = __vimax3_u16x2(up, dn, __vmaxu2(E[1][2 * p], E[1][2 * p + 2]));
          const uint32_t sc = E[1][2 * p + 1];
          const uint32_t d = __viaddmax_s16x2_relu(sc + 0x01000100u - M, 0xff00ff00u, 0u);  // max(s - M, 0)
          const uint32_t won = __vminu2(sc, d << 8);                                         // s if s > M else 0
          // keep s only if s >= threshold byte: max(s - (th-1), 0) != 0 per lane
          const uint32_t over = __viaddmax_s16x2_relu(won, negKeep, 0u);
          k2[p] = __vminu2(won, over * 0xffu);  // over <= 255 per lane: over*255 >= won iff over != 0
        }
        word = __byte_perm(k2[0], k2[1], 0x6420);
      }
      const uint32_t keepMask = __ballot_sync(0xffffffffu, word != 0u);
      if (word != 0u) {
        const int slot = nS + __popc(keepMask & (below));
        queue[slot] = (uint16_t)e;
        kept[slot] = word;
      }
      nS += __popc(keepMask);
    }
    __syncwarp();
    if (nS > 0 || iniByte == 1) break;  // iniByte == 1: phase 0 already was the suppression at minTh
  }
  if (nS == 0) return;

  int mine = 0;
  for (int j = lane; j < nS; j += 32) {
    const uint32_t word = kept[j];
#pragma unroll
    for (int i = 0; i < 4; i++) mine += ((word >> (8 * i)) & 0xff) != 0;
  }
  // exclusive prefix of `mine` over the warp + total
  int incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += v;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  int base = 0;
  if (lane == 0) {
    base = atomicAdd(&candCount[f * nl + level], total);
    if (base + total > L.candCap) atomicExch(overflow, 1);
  }
  base = __shfl_sync(0xffffffffu, base, 0);
  int slot = base + incl - mine;
  const long long slotBase = L.candOff + (long long)f * L.candCap;
  for (int j = lane; j < nS; j += 32) {
    const uint32_t word = kept[j];
    const int e = queue[j], y = e >> 8, q = e & 0xff;
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const int sc = (word >> (8 * i)) & 0xff;
      if (sc) {
        // cv::FAST's cornerScore = sc + minTh - 1; coordinates relative to minBorder:
        // cell-local (x+3, y+3) + (j*wCell, i*hCell)  (:889-890)
        if (slot < L.candCap)
          cand[slotBase + slot] = pack_cand(outX0 + 4 * (q + 1) + i, outY0 + y, sc + minTh - 1);
        slot++;
      }
    }
  }
}

int launch_fast(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, const FastCellDesc* cells, uint32_t* cand, int* candCount,
                int* overflow, int nframes, cudaStream_t s, int levelBegin, int levelEnd) {
  if (levelEnd < 0) levelEnd = hGeom.nlevels;
  int words = 0;
  for (int l = 0; l < hGeom.nlevels; l++) words = max(words, fast_cell_words(hGeom.lv[l].wCell, hGeom.lv[l].hCell));
  const size_t smem = (size_t)words * 4 * kFastWarps;
  static SmemLimit limit;
  cudaError_t e = raise_dynamic_smem(reinterpret_cast<const void*>(fast_cells_kernel), smem, &limit);
  if (e != cudaSuccess) return (int)e;
  // cells of levels [levelBegin, levelEnd): the table entries of a level are contiguous
  const int cellBegin = hGeom.lv[levelBegin].cellBase;
  const int cellEnd = levelEnd < hGeom.nlevels ? hGeom.lv[levelEnd].cellBase : hGeom.totalCells;
  if (cellEnd <= cellBegin) return 0;
  dim3 grid((cellEnd - cellBegin + kFastWarps - 1) / kFastWarps, nframes);
  launch_pdl(fast_cells_kernel, grid, dim3(kFastThreads), smem, s, pyr, hGeom, cells, cand, candCount, overflow, words, cellBegin,
             cellEnd);
  return 0;
}

}  // namespace orbx
