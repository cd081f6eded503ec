// Pyramid construction and descriptor pre-blur.
//
//  resize_kernel : level l of every frame from level l-1, cv::resize INTER_LINEAR u8 semantics
//                  (R/src/ORBextractor.cpp:1212). The 11-bit fixed-point coefficient tables are
//                  computed once per shape on the host (orbx_api.cu: build_resize_tables) exactly as
//                  OpenCV does; the kernel is pure integer arithmetic.
//  blur_kernel   : cv::GaussianBlur 7x7 sigma 2 BORDER_REFLECT_101 on a border-less copy of each level
//                  (R/src/ORBextractor.cpp:1166-1167): 8.8 fixed-point kernel {18,34,48,56,48,34,18},
//                  exact 16-bit horizontal pass, 16.16 vertical pass, one rounding.
//
// Both are HBM-streaming stencils: each source byte is read once into shared memory / L1 and each
// destination byte written once.
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

void launch_resize(const uint8_t* src, long long srcFrameStride, int srcPitch, int sw, int sh,
                   uint8_t* dst, long long dstFrameStride, int dstPitch, int dw, int dh,
                   const ResizeTables& t, int nframes, cudaStream_t s) {
  (void)sw; (void)sh;
  dim3 block(32, 8);
  dim3 grid((dw + 31) / 32, (dh + 7) / 8, nframes);
  resize_kernel<<<grid, block, 0, s>>>(src, srcFrameStride, srcPitch, dst, dstFrameStride, dstPitch, dw, dh,
                                       t.xt, t.yt);
}

// ------------------------------------------------------------------------------------------------
// blur
// ------------------------------------------------------------------------------------------------
constexpr int kBlurTW = 64, kBlurTH = 32;

__device__ __forceinline__ int reflect101(int p, int len) {
  // len >= 4 on every level we accept, so one reflection per side is enough for a 3-pixel halo
  if (p < 0) p = -p;
  if (p >= len) p = 2 * (len - 1) - p;
  return p;
}

__device__ __forceinline__ void blur_tile(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst,
                                          int dstPitch, int w, int h, int tx0, int ty0) {
  __shared__ uint8_t in[kBlurTH + 6][kBlurTW + 8];
  __shared__ uint16_t hb[kBlurTH + 6][kBlurTW];
  const int tid = threadIdx.x;
  for (int i = tid; i < (kBlurTH + 6) * (kBlurTW + 6); i += blockDim.x) {
    const int r = i / (kBlurTW + 6), c = i - r * (kBlurTW + 6);
    const int y = reflect101(min(ty0 + r - 3, h + 2), h), x = reflect101(min(tx0 + c - 3, w + 2), w);
    in[r][c] = src[(long long)y * srcPitch + x];
  }
  __syncthreads();
  for (int i = tid; i < (kBlurTH + 6) * kBlurTW; i += blockDim.x) {
    const int r = i / kBlurTW, c = i - r * kBlurTW;
    const uint8_t* p = &in[r][c];
    hb[r][c] = (uint16_t)(18 * (p[0] + p[6]) + 34 * (p[1] + p[5]) + 48 * (p[2] + p[4]) + 56 * p[3]);
  }
  __syncthreads();
  for (int i = tid; i < kBlurTH * kBlurTW; i += blockDim.x) {
    const int r = i / kBlurTW, c = i - r * kBlurTW;
    const int x = tx0 + c, y = ty0 + r;
    if (x < w && y < h) {
      const uint32_t v = 18u * (hb[r][c] + hb[r + 6][c]) + 34u * (hb[r + 1][c] + hb[r + 5][c]) +
                         48u * (hb[r + 2][c] + hb[r + 4][c]) + 56u * hb[r + 3][c];
      dst[(long long)y * dstPitch + x] = (uint8_t)((v + 32768u) >> 16);
    }
  }
}

__global__ void __launch_bounds__(256)
blur_kernel(PyramidPtrs pyr, const FrameSetGeom* __restrict__ g, uint8_t* __restrict__ blurArena) {
  int level = 0;
  const int tile = blockIdx.x;
  const int nl = g->nlevels;
  while (level + 1 < nl && tile >= g->lv[level + 1].tileBase) level++;
  const LevelGeom& L = g->lv[level];
  const int t = tile - L.tileBase;
  const int ty = t / L.tilesX, tx = t - ty * L.tilesX;
  const int f = blockIdx.y;
  const uint8_t* src = pyr.lvl[level] + (long long)f * pyr.frameStride[level];
  uint8_t* dst = blurArena + L.planeOff + (long long)f * L.planeBytes;
  blur_tile(src, pyr.pitch[level], dst, L.pitch, L.w, L.h, tx * kBlurTW, ty * kBlurTH);
}

__global__ void __launch_bounds__(256)
blur_single_kernel(const uint8_t* __restrict__ src, int srcPitch, uint8_t* __restrict__ dst, int dstPitch, int w,
                   int h) {
  blur_tile(src, srcPitch, dst, dstPitch, w, h, blockIdx.x * kBlurTW, blockIdx.y * kBlurTH);
}

void blur_tiles_for(int w, int h, int* tilesX, int* tilesY) {
  *tilesX = (w + kBlurTW - 1) / kBlurTW;
  *tilesY = (h + kBlurTH - 1) / kBlurTH;
}

void launch_blur(const PyramidPtrs& pyr, const FrameSetGeom* dGeom, const FrameSetGeom& hGeom,
                 uint8_t* blurArena, int nframes, cudaStream_t s) {
  dim3 grid(hGeom.totalTiles, nframes);
  blur_kernel<<<grid, 256, 0, s>>>(pyr, dGeom, blurArena);
}

void launch_blur_single(const uint8_t* src, int srcPitch, uint8_t* dst, int dstPitch, int w, int h, cudaStream_t s) {
  dim3 grid((w + kBlurTW - 1) / kBlurTW, (h + kBlurTH - 1) / kBlurTH);
  blur_single_kernel<<<grid, 256, 0, s>>>(src, srcPitch, dst, dstPitch, w, h);
}

}  // namespace orbx
