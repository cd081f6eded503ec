// Pyramid construction and descriptor pre-blur.
//
//  resize : level l of every frame from level l-1, cv::resize INTER_LINEAR u8 semantics
//           (R/src/ORBextractor.cpp:1212). The 11-bit fixed-point coefficient tables are computed
//           once per shape on the host (orbx_geom.cu) exactly as OpenCV does; the kernels are pure
//           integer arithmetic.
//           resize4_kernel: one thread per 4 destination pixels. Per source row it loads the three
//           aligned words that hold the <= 8 source bytes of the group, realigns them with a funnel
//           shift, gathers (left,right) byte pairs with two PRMTs whose selectors come from the
//           table, and does the horizontal interpolation with IDP.2A (16-bit weights x 8-bit
//           pixels); one 32-bit store per group.
//           resize_kernel: the general per-pixel form, used when a group spans more than 8 source
//           bytes (scale factors >= 1.75) or the source rows are not 4-byte aligned.
//  blur   : cv::GaussianBlur 7x7 sigma 2 BORDER_REFLECT_101 on a border-less copy of each level
//           (R/src/ORBextractor.cpp:1166-1167): 8.8 fixed-point kernel {18,34,48,56,48,34,18},
//           exact 16-bit horizontal pass, 16.16 vertical pass, one rounding.
//           One warp per 128x32 pixel strip: each lane owns 4 adjacent columns, the horizontal pass
//           is 10 IDP.4A per row on the lane's word and its neighbours' words (two shuffles), the
//           vertical pass runs over a 7-row register window. No shared memory, every source row is
//           read once per strip (+6 halo rows), one 32-bit store per lane and row.
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

__device__ __forceinline__ void resize_row4(const uint32_t* __restrict__ row, const ResizeGroup& G, int maxWord,
                                            uint32_t (&h)[4]) {
  const uint32_t w0 = row[G.word];
  const uint32_t w1 = row[min(G.word + 1, maxWord)];
  const uint32_t w2 = row[min(G.word + 2, maxWord)];
  const uint32_t lo = __funnelshift_r(w0, w1, G.shift);
  const uint32_t hi = __funnelshift_r(w1, w2, G.shift);
  const uint32_t b01 = __byte_perm(lo, hi, G.sel01);  // (L0, R0, L1, R1)
  const uint32_t b23 = __byte_perm(lo, hi, G.sel23);
  h[0] = __dp2a_lo(G.c[0], b01, 0u);
  h[1] = __dp2a_hi(G.c[1], b01, 0u);
  h[2] = __dp2a_lo(G.c[2], b23, 0u);
  h[3] = __dp2a_hi(G.c[3], b23, 0u);
}

__global__ void __launch_bounds__(256)
resize4_kernel(const uint8_t* __restrict__ src, long long srcFrameStride, int srcPitch,
               uint8_t* __restrict__ dst, long long dstFrameStride, int dstPitch, int dw, int dh,
               const ResizeGroup* __restrict__ xg, const ResizeTap* __restrict__ yt) {
  const int gx = blockIdx.x * 32 + threadIdx.x;  // group of 4 destination pixels
  const int dy = blockIdx.y * 8 + threadIdx.y;
  if (4 * gx >= dw || dy >= dh) return;
  const uint8_t* S = src + (long long)blockIdx.z * srcFrameStride;
  const ResizeGroup G = xg[gx];
  const ResizeTap ty = yt[dy];
  const int maxWord = (srcPitch >> 2) - 1;
  uint32_t h0[4], h1[4];
  resize_row4(reinterpret_cast<const uint32_t*>(S + (long long)ty.p0 * srcPitch), G, maxWord, h0);
  resize_row4(reinterpret_cast<const uint32_t*>(S + (long long)ty.p1 * srcPitch), G, maxWord, h1);
  uint32_t out = 0;
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const int v = ((((int)ty.c0 * (int)(h0[j] >> 4)) >> 16) + (((int)ty.c1 * (int)(h1[j] >> 4)) >> 16) + 2) >> 2;
    out |= (uint32_t)min(max(v, 0), 255) << (8 * j);
  }
  // destination planes are ours: pitch is a multiple of 16, so the whole word is inside the row
  *reinterpret_cast<uint32_t*>(dst + (long long)blockIdx.z * dstFrameStride + (long long)dy * dstPitch + 4 * gx) = out;
}

void launch_resize(const uint8_t* src, long long srcFrameStride, int srcPitch, int sw, int sh,
                   uint8_t* dst, long long dstFrameStride, int dstPitch, int dw, int dh,
                   const ResizeTables& t, int nframes, cudaStream_t s) {
  (void)sw; (void)sh;
  dim3 block(32, 8);
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)srcPitch | (uintptr_t)srcFrameStride |
                         reinterpret_cast<uintptr_t>(dst) | (uintptr_t)dstPitch | (uintptr_t)dstFrameStride) & 3) == 0;
  if (t.xg && aligned) {
    dim3 grid(((dw + 3) / 4 + 31) / 32, (dh + 7) / 8, nframes);
    resize4_kernel<<<grid, block, 0, s>>>(src, srcFrameStride, srcPitch, dst, dstFrameStride, dstPitch, dw, dh,
                                          t.xg, t.yt);
  } else {
    dim3 grid((dw + 31) / 32, (dh + 7) / 8, nframes);
    resize_kernel<<<grid, block, 0, s>>>(src, srcFrameStride, srcPitch, dst, dstFrameStride, dstPitch, dw, dh,
                                         t.xt, t.yt);
  }
}

// ------------------------------------------------------------------------------------------------
// blur
// ------------------------------------------------------------------------------------------------
constexpr int kBlurStripW = 128, kBlurStripH = 32;
constexpr int kBlurWarps = 4;

__device__ __forceinline__ int reflect101(int p, int len) {
  // len >= 4 on every level we accept, so one reflection per side is enough for a 3-pixel halo
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

// pixels x..x+3 of a row as one word, BORDER_REFLECT_101 outside [0,w)
__device__ __forceinline__ uint32_t blur_load4(const uint8_t* __restrict__ row, int x, int w, bool aligned) {
  if (aligned && x >= 0 && x + 3 < w) return *reinterpret_cast<const uint32_t*>(row + x);
  uint32_t v = 0;
#pragma unroll
  for (int i = 0; i < 4; i++) {
    const int c = reflect101(min(x + i, w + 2), w);
    v |= (uint32_t)row[c] << (8 * i);
  }
  return v;
}

__device__ __forceinline__ uint32_t dp4a_u(uint32_t a, uint32_t b, uint32_t c) { return __dp4a(a, b, c); }

// One warp: output columns x0 + 4*lane .. +3, rows y0 .. y0+kBlurStripH-1.
__device__ __forceinline__ void blur_strip(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst,
                                           int dstPitch, int w, int h, int x0, int y0) {
  const int lane = threadIdx.x & 31;
  const int x = x0 + 4 * lane;
  const bool aligned = ((reinterpret_cast<uintptr_t>(src) | (uintptr_t)srcPitch) & 3) == 0;
  // horizontal taps of pixel j over the 12-byte window (left word, own word, right word), byte i of
  // the window = column x-4+i; kernel k = {18,34,48,56,48,34,18} centred on column x+j
  constexpr uint32_t L0 = 0x30221200u, C0 = 0x12223038u;                    // j=0: bytes 1..7
  constexpr uint32_t L1 = 0x22120000u, C1 = 0x22303830u, R1 = 0x00000012u;  // j=1: bytes 2..8
  constexpr uint32_t L2 = 0x12000000u, C2 = 0x30383022u, R2 = 0x00001222u;  // j=2: bytes 3..9
  constexpr uint32_t C3 = 0x38302212u, R3 = 0x00122230u;                    // j=3: bytes 4..10
  int H[7][4];
  const bool active = x < w;  // lanes past the right edge still feed their left neighbour
  const bool feeds = x < w + 4;
#pragma unroll
  for (int r = 0; r < kBlurStripH + 6; r++) {
    const int yy = reflect101(min(y0 + r - 3, h + 2), h);
    const uint8_t* row = src + (long long)yy * srcPitch;
    uint32_t c = 0;
    if (feeds) c = blur_load4(row, x, w, aligned);
    uint32_t l = __shfl_up_sync(0xffffffffu, c, 1);
    uint32_t rr = __shfl_down_sync(0xffffffffu, c, 1);
    if (lane == 0) l = blur_load4(row, x - 4, w, aligned);
    if (lane == 31 && active) rr = blur_load4(row, x + 4, w, aligned);
    int* Hr = H[r % 7];
    Hr[0] = (int)dp4a_u(l, L0, dp4a_u(c, C0, 0u));
    Hr[1] = (int)dp4a_u(l, L1, dp4a_u(c, C1, dp4a_u(rr, R1, 0u)));
    Hr[2] = (int)dp4a_u(l, L2, dp4a_u(c, C2, dp4a_u(rr, R2, 0u)));
    Hr[3] = (int)dp4a_u(c, C3, dp4a_u(rr, R3, 0u));
    if (r >= 6) {
      const int y = y0 + r - 6;
      // rows r-6..r of the window are taps 0..6
      uint32_t out = 0;
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const uint32_t v = 18u * (uint32_t)(H[(r - 6) % 7][j] + H[r % 7][j]) +
                           34u * (uint32_t)(H[(r - 5) % 7][j] + H[(r - 1) % 7][j]) +
                           48u * (uint32_t)(H[(r - 4) % 7][j] + H[(r - 2) % 7][j]) +
                           56u * (uint32_t)H[(r - 3) % 7][j];
        out |= ((v + 32768u) >> 16) << (8 * j);
      }
      // destination planes are ours (pitch multiple of 16): the word never leaves the row
      if (active && y < h) *reinterpret_cast<uint32_t*>(dst + (long long)y * dstPitch + x) = out;
    }
  }
}

__global__ void __launch_bounds__(kBlurWarps * 32)
blur_kernel(const __grid_constant__ PyramidPtrs pyr, const __grid_constant__ FrameSetGeom g,
            uint8_t* __restrict__ blurArena) {
  const int task = blockIdx.x * kBlurWarps + (threadIdx.x >> 5);
  if (task >= g.totalTiles) return;
  int level = 0;
  const int nl = g.nlevels;
  while (level + 1 < nl && task >= g.lv[level + 1].tileBase) level++;
  const LevelGeom& L = g.lv[level];
  const int t = task - L.tileBase;
  const int ty = t / L.tilesX, tx = t - ty * L.tilesX;
  const int f = blockIdx.y;
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level];
  uint8_t* dst = blurArena + L.planeOff + (long long)f * L.planeBytes;
  blur_strip(src, pyr.pitch[level], dst, L.pitch, L.w, L.h, tx * kBlurStripW, ty * kBlurStripH);
}

__global__ void __launch_bounds__(kBlurWarps * 32)
blur_single_kernel(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst, int dstPitch, int w,
                   int h, int tilesX, int tiles) {
  const int task = blockIdx.x * kBlurWarps + (threadIdx.x >> 5);
  if (task >= tiles) return;
  const int ty = task / tilesX, tx = task - ty * tilesX;
  blur_strip(src, srcPitch, dst, dstPitch, w, h, tx * kBlurStripW, ty * kBlurStripH);
}

void blur_tiles_for(int w, int h, int* tilesX, int* tilesY) {
  *tilesX = (w + kBlurStripW - 1) / kBlurStripW;
  *tilesY = (h + kBlurStripH - 1) / kBlurStripH;
}

void launch_blur(const PyramidPtrs& pyr, const FrameSetGeom& hGeom, uint8_t* blurArena, int nframes, cudaStream_t s) {
  dim3 grid((hGeom.totalTiles + kBlurWarps - 1) / kBlurWarps, nframes);
  blur_kernel<<<grid, kBlurWarps * 32, 0, s>>>(pyr, hGeom, blurArena);
}

void launch_blur_single(const uint8_t* src, int srcPitch, uint8_t* dst, int dstPitch, int w, int h, cudaStream_t s) {
  int tx, ty;
  blur_tiles_for(w, h, &tx, &ty);
  blur_single_kernel<<<(tx * ty + kBlurWarps - 1) / kBlurWarps, kBlurWarps * 32, 0, s>>>(src, srcPitch, dst, dstPitch,
                                                                                         w, h, tx, tx * ty);
}

}  // namespace orbx
