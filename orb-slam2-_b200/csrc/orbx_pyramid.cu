// Pyramid construction and descriptor pre-blur.
//
//  resize : level l of every frame from level l-1, cv::resize INTER_LINEAR u8 semantics
//           (R/src/ORBextractor.cpp:1212). The 11-bit fixed-point coefficient tables are computed
//           once per shape on the host (orbx_geom.cu) exactly as OpenCV does; the kernels are pure
//           integer arithmetic.
//           resize4_kernel: one thread per 4 destination pixels x 8 destination rows (horizontally
//           interpolated source rows are reused between them). Per source row it loads the three
//           aligned words that hold the <= 8 source bytes of the group, realigns them with a funnel
//           shift, gathers (left,right) byte pairs with two PRMTs whose selectors come from the
//           table, and does the horizontal interpolation with IDP.2A (16-bit weights x 8-bit
//           pixels); one 32-bit store per group.
//           resize_kernel: the general per-pixel form, used when a group spans more than 8 source
//           bytes (scale factors >= 1.75) or the source rows are not 4-byte aligned.
//  blur   : cv::GaussianBlur 7x7 sigma 2 BORDER_REFLECT_101 on a border-less copy of each level
//           (R/src/ORBextractor.cpp:1166-1167): 8.8 fixed-point kernel {18,34,48,56,48,34,18},
//           exact 16-bit horizontal pass, 16.16 vertical pass, one rounding.
//           Eight lanes per 32x32 pixel strip (four strips per warp): each lane owns 4 adjacent
//           columns, the horizontal pass is 10 IDP.4A per row on the lane's word and its neighbours'
//           words (two width-8 shuffles), the vertical pass runs over a 7-row register window, loads
//           run 4 rows ahead. No shared memory, every source row is read once per strip (+6 halo
//           rows), one 32-bit store per lane and row; image borders are rebuilt with PRMT.
//
// Both are HBM-streaming stencils: each source byte is fetched from DRAM once and each destination
// byte written once.
#include "orbx_common.cuh"

namespace orbx {

// ------------------------------------------------------------------------------------------------
// resize
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
resize_kernel(const uint8_t* __restrict__ src, long long srcFrameStride, int srcPitch,
              uint8_t* __restrict__ dst, long long dstFrameStride, int dstPitch, int dw, int dh,
              const ResizeTap* __restrict__ xt, const ResizeTap* __restrict__ yt) {
  pdl_wait();
  const int dx = blockIdx.x * 32 + threadIdx.x;
  const int dy = blockIdx.y * 8 + threadIdx.y;
  if (dx >= dw || dy >= dh) return;
  const uint8_t* S = src + (long long)blockIdx.z * srcFrameStride;
  const ResizeTap tx = xt[dx];
  const ResizeTap ty = yt[dy];
  const uint8_t* r0 = S + (long long)ty.p0 * srcPitch;
  const uint8_t* r1 = S + (long long)ty.p1 * srcPitch;
  const int h0 = (int)r0[tx.p0] * tx.c0 + (int)r0[tx.p1] * tx.c1;
  const int h1 = (int)r1[tx.p0] * tx.c0 + (int)r1[tx.p1] * tx.c1;
  const int v = ((((int)ty.c0 * (h0 >> 4)) >> 16) + (((int)ty.c1 * (h1 >> 4)) >> 16) + 2) >> 2;
  dst[(long long)blockIdx.z * dstFrameStride + (long long)dy * dstPitch + dx] = (uint8_t)min(max(v, 0), 255);
}

// One horizontally interpolated source row for the group's four pixels from its three aligned words, already >> 4
// (the vertical pass only ever uses that form, resize.cpp VResizeLinear).
__device__ __forceinline__ void resize_hrow4(uint32_t w0, uint32_t w1, uint32_t w2, const ResizeGroup& G, uint32_t (&h)[4]) {
  const uint32_t lo = __funnelshift_r(w0, w1, G.shift);
  const uint32_t hi = __funnelshift_r(w1, w2, G.shift);
  const uint32_t b01 = __byte_perm(lo, hi, G.sel01);  // (L0, R0, L1, R1)
  const uint32_t b23 = __byte_perm(lo, hi, G.sel23);
  h[0] = __dp2a_lo(G.c[0], b01, 0u) >> 4;
  h[1] = __dp2a_hi(G.c[1], b01, 0u) >> 4;
  h[2] = __dp2a_lo(G.c[2], b23, 0u) >> 4;
  h[3] = __dp2a_hi(G.c[3], b23, 0u) >> 4;
}

#ifndef ORBX_RESIZE_ROWS
#define ORBX_RESIZE_ROWS 8
#endif
#ifndef ORBX_RESIZE_BANDS
#define ORBX_RESIZE_BANDS 4
#endif
#ifndef ORBX_RESIZE_PREFETCH
#define ORBX_RESIZE_PREFETCH 0
#endif
constexpr int kResizeRows = ORBX_RESIZE_ROWS;    // destination rows per thread (source rows are reused between them)
constexpr int kResizeBands = ORBX_RESIZE_BANDS;  // threadIdx.y
#ifndef ORBX_RESIZE_ROWS_SMALL
#define ORBX_RESIZE_ROWS_SMALL 2
#endif
constexpr int kResizeRowsSmall = ORBX_RESIZE_ROWS_SMALL;
#ifndef ORBX_RESIZE_FILL
#define ORBX_RESIZE_FILL 1536
#endif
constexpr long long kResizeFillThreads = 148LL * ORBX_RESIZE_FILL;  // fewer threads than this at kResizeRows cannot fill the GPU

// Requires non-negative weights (build_resize_groups checks both axes; always true when shrinking): then
// (c * (h >> 4)) >> 16 == umulhi(c << 16, h >> 4), two IMAD.HI per pixel instead of multiply + shift twice, and
// the result needs no clamp: c0 + c1 == 2048 and h >> 4 <= 32640 bound it by (1020 + 2) >> 2 = 255.
// A thread walks down consecutive source rows, so the words of the row after the last one used are requested one
// destination row ahead (the kernel was waiting on exactly these loads 59 % of the time).
#ifndef ORBX_RESIZE_MINB
#define ORBX_RESIZE_MINB 12
#endif
// ROWS = destination rows per thread. A thread loads its source rows one after the other (each use waits for its
// load), so a launch that cannot fill the GPU anyway -- one frame, or the small levels of a small batch -- is bound by
// that chain: kResizeRowsSmall rows per thread make four times the threads with a quarter of the chain, at the
// price of re-interpolating more shared source rows (1.5 instead of 1.33 per destination row).
template <int ROWS>
__global__ void __launch_bounds__(32 * kResizeBands, ORBX_RESIZE_MINB)
resize4_kernel(const uint8_t* __restrict__ src, long long srcFrameStride, int srcPitch, int sh,
               uint8_t* __restrict__ dst, long long dstFrameStride, int dstPitch, int dw, int dh,
               const ResizeGroup* __restrict__ xg, const ResizeTap* __restrict__ yt) {
  pdl_wait();
  const int gx = blockIdx.x * 32 + threadIdx.x;  // group of 4 destination pixels
  const int dy0 = (blockIdx.y * kResizeBands + threadIdx.y) * ROWS;
  if (4 * gx >= dw || dy0 >= dh) return;
  const uint8_t* S = src + (long long)blockIdx.z * srcFrameStride;
  uint8_t* D = dst + (long long)blockIdx.z * dstFrameStride + 4 * gx + (long long)dy0 * dstPitch;
  const ResizeGroup G = xg[gx];
  const int maxWord = (srcPitch >> 2) - 1;
  // byte offsets of the three aligned words of a row that hold the group's <= 8 source bytes. (Keeping three
  // 64-bit column pointers instead saves instructions but costs 16 registers, and this kernel lives on occupancy:
  // it spends most of its time waiting for the row loads.)
  const uint32_t o0 = 4u * (uint32_t)G.word, o1 = 4u * (uint32_t)min(G.word + 1, maxWord),
                 o2 = 4u * (uint32_t)min(G.word + 2, maxWord);
  int nxtRow = -1;
  uint32_t n0 = 0, n1 = 0, n2 = 0;  // words of source row nxtRow, in flight (ORBX_RESIZE_PREFETCH)
  auto request = [&](int r, uint32_t& a, uint32_t& b, uint32_t& c) {
    const uint32_t row = (uint32_t)r * (uint32_t)srcPitch;
    a = *reinterpret_cast<const uint32_t*>(S + (row + o0));
    b = *reinterpret_cast<const uint32_t*>(S + (row + o1));
    c = *reinterpret_cast<const uint32_t*>(S + (row + o2));
  };
  auto take = [&](int r, uint32_t (&h)[4]) {
    uint32_t a, b, c;
#if ORBX_RESIZE_PREFETCH
    if (r == nxtRow) { a = n0; b = n1; c = n2; }
    else request(r, a, b, c);
    nxtRow = min(r + 1, sh - 1);
    request(nxtRow, n0, n1, n2);
#else
    request(r, a, b, c);
#endif
    resize_hrow4(a, b, c, G, h);
  };
  // horizontally interpolated source rows ia / ib are kept between destination rows (dy is uniform
  // across the warp, so the reuse branches do not diverge)
  int ia = -1, ib = -1;
  uint32_t ha[4] = {0, 0, 0, 0}, hb[4] = {0, 0, 0, 0};
  const int dyEnd = min(dy0 + ROWS, dh);
  uint2 tyNext = *reinterpret_cast<const uint2*>(&yt[dy0]);
  for (int dy = dy0; dy < dyEnd; dy++) {
    const uint2 ty = tyNext;  // (p0 | p1 << 16, c0 | c1 << 16), all >= 0
    tyNext = *reinterpret_cast<const uint2*>(&yt[min(dy + 1, dh - 1)]);
    const int p0 = (int)(ty.x & 0xffffu), p1 = (int)(ty.x >> 16);
    const uint32_t c0s = ty.y << 16, c1s = ty.y & 0xffff0000u;
    if (p0 != ia) {
      if (p0 == ib) {
#pragma unroll
        for (int j = 0; j < 4; j++) ha[j] = hb[j];
      } else {
        take(p0, ha);
      }
      ia = p0;
    }
    if (p1 != ib) {
      if (p1 == ia) {
#pragma unroll
        for (int j = 0; j < 4; j++) hb[j] = ha[j];
      } else {
        take(p1, hb);
      }
      ib = p1;
    }
    uint32_t v[4];
#pragma unroll
    for (int j = 0; j < 4; j++) v[j] = (__umulhi(c0s, ha[j]) + __umulhi(c1s, hb[j]) + 2u) >> 2;
    // destination planes are ours: pitch is a multiple of 16, so the whole word is inside the row
    *reinterpret_cast<uint32_t*>(D) = __byte_perm(__byte_perm(v[0], v[1], 0x0040), __byte_perm(v[2], v[3], 0x0040), 0x5410);
    D += dstPitch;
  }
}

void launch_resize(const uint8_t* src, long long srcFrameStride, int srcPitch, int sw, int sh,
                   uint8_t* dst, long long dstFrameStride, int dstPitch, int dw, int dh,
                   const ResizeTables& t, int nframes, cudaStream_t s) {
  (void)sw;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)srcPitch | (uintptr_t)srcFrameStride |
                         reinterpret_cast<uintptr_t>(dst) | (uintptr_t)dstPitch | (uintptr_t)dstFrameStride) & 3) == 0;
  if (t.xg && aligned) {
    const long long threadsAtFull = (long long)((dw + 3) / 4) * ((dh + kResizeRows - 1) / kResizeRows) * nframes;
    if (threadsAtFull >= kResizeFillThreads) {
      const int rowsPerCta = kResizeRows * kResizeBands;
      dim3 grid(((dw + 3) / 4 + 31) / 32, (dh + rowsPerCta - 1) / rowsPerCta, nframes);
      launch_pdl(resize4_kernel<kResizeRows>, grid, dim3(32, kResizeBands), 0, s, src, srcFrameStride, srcPitch, sh, dst,
                 dstFrameStride, dstPitch, dw, dh, t.xg, t.yt);
    } else {
      const int rowsPerCta = kResizeRowsSmall * kResizeBands;
      dim3 grid(((dw + 3) / 4 + 31) / 32, (dh + rowsPerCta - 1) / rowsPerCta, nframes);
      launch_pdl(resize4_kernel<kResizeRowsSmall>, grid, dim3(32, kResizeBands), 0, s, src, srcFrameStride, srcPitch, sh, dst,
                 dstFrameStride, dstPitch, dw, dh, t.xg, t.yt);
    }
  } else {
    dim3 grid((dw + 31) / 32, (dh + 7) / 8, nframes);
    launch_pdl(resize_kernel, grid, dim3(32, 8), 0, s, src, srcFrameStride, srcPitch, dst, dstFrameStride, dstPitch, dw, dh,
               t.xt, t.yt);
  }
}

// ------------------------------------------------------------------------------------------------
// blur
// ------------------------------------------------------------------------------------------------
#ifndef ORBX_BLUR_H
#define ORBX_BLUR_H 30
#endif
constexpr int kBlurStripW = 32, kBlurStripH = ORBX_BLUR_H;  // one strip = 8 lanes x 4 px, a multiple of 6 output rows
constexpr int kBlurWarps = 4;                      // 4 strips per warp, 16 per CTA
constexpr int kBlurPrefetch = 6;                   // rows in flight per lane

__device__ __forceinline__ int reflect101(int p, int len) {
  // len >= 4 on every level we accept, so one reflection per side is enough for a 3-pixel halo
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

// horizontal taps of pixel j over the 12-byte window (left word, own word, right word), byte i of the
// window = column x-4+i; kernel k = {18,34,48,56,48,34,18} centred on column x+j
__device__ __forceinline__ void blur_hrow(uint32_t l, uint32_t c, uint32_t r, int (&H)[4]) {
  constexpr uint32_t L0 = 0x30221200u, C0 = 0x12223038u;                    // j=0: bytes 1..7
  constexpr uint32_t L1 = 0x22120000u, C1 = 0x22303830u, R1 = 0x00000012u;  // j=1: bytes 2..8
  constexpr uint32_t L2 = 0x12000000u, C2 = 0x30383022u, R2 = 0x00001222u;  // j=2: bytes 3..9
  constexpr uint32_t C3 = 0x38302212u, R3 = 0x00122230u;                    // j=3: bytes 4..10
  H[0] = (int)__dp4a(l, L0, __dp4a(c, C0, 0u));
  H[1] = (int)__dp4a(l, L1, __dp4a(c, C1, __dp4a(r, R1, 0u)));
  H[2] = (int)__dp4a(l, L2, __dp4a(c, C2, __dp4a(r, R2, 0u)));
  H[3] = (int)__dp4a(c, C3, __dp4a(r, R3, 0u));
}

// Vertical pass on packed rows: P[k] holds (H[k-1], H[k]) as two 16-bit lanes (H <= 255*256), so the seven taps of
// the output row centred on r-3 are four IDP.2A (16-bit x 8-bit) over P[r-5], P[r-3], P[r-1], P[r], the rounding
// constant riding in the accumulator; the four results (< 2^24, wanted bits 16..23) are packed with three PRMT.
__device__ __forceinline__ uint32_t blur_vcol(const uint32_t (&P)[6][4], int r) {
  constexpr uint32_t K01 = 18u | (34u << 8), K23 = 48u | (56u << 8), K45 = 48u | (34u << 8), K6 = 18u << 8;
  uint32_t v[4];
#pragma unroll
  for (int j = 0; j < 4; j++)
    v[j] = __dp2a_lo(P[(r + 1) % 6][j], K01,
                     __dp2a_lo(P[(r + 3) % 6][j], K23, __dp2a_lo(P[(r + 5) % 6][j], K45, __dp2a_lo(P[r % 6][j], K6, 32768u))));
  return __byte_perm(__byte_perm(v[0], v[1], 0x0062), __byte_perm(v[2], v[3], 0x0062), 0x5410);
}

// Fast path: source rows 4-byte aligned, w >= 12. Eight lanes own a 32-px wide strip; every lane does
// one aligned 32-bit load per row (plus one halo load on the strip's first and last lane); image
// borders (BORDER_REFLECT_101) are rebuilt from the loaded words with PRMT selectors that are fixed
// per strip. `valid` is false for the lanes of a quarter-warp without a strip: they run the
// shuffles with everybody else and store nothing. kInterior: rows y0-3 .. y0+kBlurStripH+2 all exist, so the row
// pointer just advances by the pitch and no store needs a row test.
// Rows go by in groups of six (the period of the packed window): the first group only fills the window, the
// others are one rolled loop around six unrolled rows. (Unrolling all 38 rows of a 32-row strip twice -- interior
// and border form -- made 74 KB of code and a quarter of the stall samples were instruction-cache misses.)
template <bool kInterior>
__device__ __forceinline__ void blur_strip8(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst,
                                            int dstPitch, int w, int h, int x0, int y0, bool valid) {
  const int sub = threadIdx.x & 7;
  const int x = x0 + 4 * sub;
  const int lastX = (w - 1) & ~3;           // column of the last (possibly partial) word of a row
  const int xl = min(x, lastX);
  const bool first = sub == 0, last = sub == 7;
  const bool isLastWord = x == lastX;
  const int e = w - 1 - lastX;              // index of the last valid byte in the last word, 0..3
  // selC: (left, raw) -> the last word with columns >= w reflected; selR: (left, raw) -> the word after it
  uint32_t selC = 0, selR = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    selC |= (uint32_t)(i <= e ? 4 + i : 4 + 2 * e - i) << (4 * i);
    selR |= (uint32_t)max(2 * e - i, 0) << (4 * i);
  }
  const uint32_t selOwn = isLastWord ? selC : 0x7654u;                       // applied to (l, raw)
  const uint32_t selHalo = (x + 4 == lastX) ? selC : 0x7654u;                // applied to (own, rawRight)
  // the halo word sits at -4 (first lane) / +4 (last lane) from the lane's own word: the first lane of a strip has
  // x <= lastX, and a last lane only loads its halo when x + 4 <= lastX, so the clamp in xl never applies to a
  // lane that loads one
  const bool halo = valid && ((first && x > 0) || (last && x + 4 <= lastX));
  const int haloOff = first ? -4 : 4;
  const bool storeOk = valid && x <= lastX;

  const uint8_t* colBase = src + xl;
  const uint8_t* fp = colBase + (long long)(y0 - 3) * srcPitch;  // kInterior only
  uint8_t* dp = dst + (long long)y0 * dstPitch + x;
  uint32_t qc[kBlurPrefetch], qh[kBlurPrefetch];
  auto fetch = [&](int r, int slot) {
    const uint8_t* row;
    if constexpr (kInterior) {
      row = fp;
      fp += srcPitch;
    } else {
      row = colBase + (long long)reflect101(min(y0 + r - 3, h + 2), h) * srcPitch;
    }
    qc[slot] = valid ? *reinterpret_cast<const uint32_t*>(row) : 0u;
    qh[slot] = halo ? *reinterpret_cast<const uint32_t*>(row + haloOff) : 0u;
  };
  uint32_t P[6][4];
  int Hprev[4] = {0, 0, 0, 0};
  // row r = 6 * group + k: horizontal pass into the packed window (k = r % 6 at compile time)
  auto hpass = [&](int k) {
    const uint32_t raw = qc[k], hw = qh[k];
    uint32_t l = __shfl_up_sync(0xffffffffu, raw, 1, 8);
    if (first) l = hw;                                    // x == 0 is fixed below
    const uint32_t c = __byte_perm(l, raw, selOwn);
    uint32_t rr = __shfl_down_sync(0xffffffffu, c, 1, 8);
    if (last) rr = __byte_perm(c, hw, selHalo);
    if (isLastWord) rr = __byte_perm(l, raw, selR);
    if (x == 0) l = __byte_perm(c, rr, 0x1234);           // columns -4..-1 = columns 4,3,2,1
    int H[4];
    blur_hrow(l, c, rr, H);
#pragma unroll
    for (int j = 0; j < 4; j++) {
      P[k][j] = __byte_perm((uint32_t)Hprev[j], (uint32_t)H[j], 0x5410);
      Hprev[j] = H[j];
    }
  };
  static_assert(kBlurPrefetch == 6 && kBlurStripH % 6 == 0, "rows are processed in groups of six");
#pragma unroll
  for (int k = 0; k < 6; k++) fetch(k, k);
#pragma unroll
  for (int k = 0; k < 6; k++) {  // group 0 fills the window
    hpass(k);
    fetch(6 + k, k);
  }
#pragma unroll 1
  for (int g = 1; g <= kBlurStripH / 6; g++) {
    const bool more = g < kBlurStripH / 6;
#pragma unroll
    for (int k = 0; k < 6; k++) {
      hpass(k);
      if (more) fetch(6 * g + 6 + k, k);
      const uint32_t out = blur_vcol(P, k);
      // destination planes are ours (pitch multiple of 16): the word never leaves the row
      if (storeOk && (kInterior || y0 + 6 * g + k - 6 < h)) *reinterpret_cast<uint32_t*>(dp) = out;
      dp += dstPitch;
    }
  }
}

// interior strips take the cheaper row addressing when the whole warp agrees (its four strips are neighbours)
__device__ __forceinline__ void blur_strip_dispatch(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst,
                                                    int dstPitch, int w, int h, int x0, int y0, bool valid) {
  const bool interior = !valid || (y0 >= 3 && y0 + kBlurStripH + 3 <= h);
  if (__all_sync(0xffffffffu, interior)) blur_strip8<true>(src, srcPitch, dst, dstPitch, w, h, x0, y0, valid);
  else blur_strip8<false>(src, srcPitch, dst, dstPitch, w, h, x0, y0, valid);
}

__global__ void __launch_bounds__(kBlurWarps * 32)
blur_kernel(const __grid_constant__ PyramidPtrs pyr, const __grid_constant__ FrameSetGeom g,
            uint8_t* __restrict__ blurArena) {
  pdl_wait();
  const int task = (blockIdx.x * kBlurWarps + (threadIdx.x >> 5)) * 4 + ((threadIdx.x & 31) >> 3);
  const bool valid = task < g.totalTiles;
  int level = 0;
  const int nl = g.nlevels;
  while (level + 1 < nl && task >= g.lv[level + 1].tileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = valid ? task - L.tileBase : 0;
  const int ty = t / L.tilesX, tx = t - ty * L.tilesX;
  const int f = blockIdx.y;
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level];
  uint8_t* dst = blurArena + L.planeOff + (long long)f * L.planeBytes;
  blur_strip_dispatch(src, pyr.pitch[level], dst, L.pitch, L.w, L.h, tx * kBlurStripW, ty * kBlurStripH, valid);
}

__global__ void __launch_bounds__(kBlurWarps * 32)
blur_single_kernel(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst, int dstPitch, int w,
                   int h, int tilesX, int tiles) {
  const int task = (blockIdx.x * kBlurWarps + (threadIdx.x >> 5)) * 4 + ((threadIdx.x & 31) >> 3);
  const bool valid = task < tiles;
  const int t = valid ? task : 0;
  const int ty = t / tilesX, tx = t - ty * tilesX;
  blur_strip_dispatch(src, srcPitch, dst, dstPitch, w, h, tx * kBlurStripW, ty * kBlurStripH, valid);
}

// General form (any alignment, any size >= 4x4): one thread per pixel, used by orbx_blur7_u8 for images
// narrower than 12 px; the pipeline never needs it (levels are >= 62 px wide and staged aligned).
__global__ void __launch_bounds__(256)
blur_generic_kernel(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst, int dstPitch, int w,
                    int h) {
  const int x = blockIdx.x * 32 + (threadIdx.x & 31), y = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (x >= w || y >= h) return;
  const int k[7] = {18, 34, 48, 56, 48, 34, 18};
  uint32_t v = 0;
  for (int j = 0; j < 7; j++) {
    const uint8_t* row = src + (long long)reflect101(y + j - 3, h) * srcPitch;
    uint32_t hsum = 0;
    for (int i = 0; i < 7; i++) hsum += (uint32_t)k[i] * row[reflect101(x + i - 3, w)];
    v += (uint32_t)k[j] * hsum;
  }
  dst[(long long)y * dstPitch + x] = (uint8_t)((v + 32768u) >> 16);
}

void blur_tiles_for(int w, int h, int* tilesX, int* tilesY) {
  *tilesX = (w + kBlurStripW - 1) / kBlurStripW;
  *tilesY = (h + kBlurStripH - 1) / kBlurStripH;
}

void launch_blur(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, uint8_t* blurArena, int nframes, cudaStream_t s) {
  const int perCta = kBlurWarps * 4;
  dim3 grid((hGeom.totalTiles + perCta - 1) / perCta, nframes);
  launch_pdl(blur_kernel, grid, dim3(kBlurWarps * 32), 0, s, pyr, hGeom, blurArena);
}

void launch_blur_single(const uint8_t* src, int srcPitch, uint8_t* dst, int dstPitch, int w, int h, cudaStream_t s) {
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)srcPitch) & 3) == 0;
  if (aligned && w >= 12) {
    int tx, ty;
    blur_tiles_for(w, h, &tx, &ty);
    const int perCta = kBlurWarps * 4;
    blur_single_kernel<<<(tx * ty + perCta - 1) / perCta, kBlurWarps * 32, 0, s>>>(src, srcPitch, dst, dstPitch, w, h,
                                                                                   tx, tx * ty);
  } else {
    dim3 grid((w + 31) / 32, (h + 7) / 8);
    blur_generic_kernel<<<grid, 256, 0, s>>>(src, srcPitch, dst, dstPitch, w, h);
  }
}

}  // namespace orbx
