// Windowed Hamming search on the Frame grid: the data-parallel half of
// ORBmatcher::SearchForInitialization / SearchByProjection (R/src/ORBmatcher.cpp:499-617, :63-163,
// :1564-1718). For every query (u, v, r, level range, descriptor) the device produces, in the
// reference's order, the candidate list Frame::GetFeaturesInArea would return (R/src/Frame.cpp:387-440)
// together with DescriptorDistance(query, candidate). The order-dependent resolve (already-matched
// exclusion, best/second-best with levels, rotation histogram) stays on the host, verbatim from the
// reference (orb-slam2-_b200/host/orbx_search.cc).
//
//  grid_build_kernel : AssignFeaturesToGrid (:244-260) + PosInGrid (:442-452) as a CSR over the
//                      64x48 cells, cell id = ix*48+iy, items ascending by keypoint index (= push order).
//                      Because GetFeaturesInArea walks ix outer / iy inner, the cells a query visits for
//                      one ix are ONE contiguous item range of the CSR.
//  window_kernel     : one warp per query; pass 0 counts, pass 1 writes (idx, distance) at the query's
//                      offset with ballot-ordered compaction.
#include "orbx_common.cuh"

namespace orbx {

constexpr int kGridCols = 64, kGridRows = 48, kGridCells = kGridCols * kGridRows;  // R/include/Frame.h:37-38

struct GridParams {
  float minX, minY, invW, invH;  // mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv
};

__global__ void __launch_bounds__(1024)
grid_build_kernel(const orbx_keypoint* __restrict__ kps, int n, GridParams gp, int* __restrict__ cellStart /*[cells+1]*/,
                  int* __restrict__ items /*[n]*/) {
  __shared__ int cnt[kGridCells];
  __shared__ int warpSums[32];
  const int tid = threadIdx.x, T = blockDim.x;
  for (int i = tid; i < kGridCells; i += T) cnt[i] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += T) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
    if (px >= 0 && px < kGridCols && py >= 0 && py < kGridRows) atomicAdd(&cnt[px * kGridRows + py], 1);
  }
  __syncthreads();
  // exclusive scan of cnt (3 cells per thread)
  constexpr int per = kGridCells / 1024;  // 3
  int local[per], s = 0;
#pragma unroll
  for (int k = 0; k < per; k++) { local[k] = cnt[tid * per + k]; s += local[k]; }
  const int lane = tid & 31, warp = tid >> 5;
  int v = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
  if (lane == 31) warpSums[warp] = v;
  __syncthreads();
  if (warp == 0) {
    int w = warpSums[lane];
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
    warpSums[lane] = w;
  }
  __syncthreads();
  int run = v - s + (warp ? warpSums[warp - 1] : 0);
#pragma unroll
  for (int k = 0; k < per; k++) { cellStart[tid * per + k] = run; cnt[tid * per + k] = run; run += local[k]; }
  if (tid == T - 1) cellStart[kGridCells] = run;
  __syncthreads();
  // fill (unordered), then sort every cell's few items ascending = the reference's push_back order
  for (int i = tid; i < n; i += T) {
    const int px = (int)roundf(__fmul_rn(__fsub_rn(kps[i].x, gp.minX), gp.invW));
    const int py = (int)roundf(__fmul_rn(__fsub_rn(kps[i].y, gp.minY), gp.invH));
    if (px >= 0 && px < kGridCols && py >= 0 && py < kGridRows) items[atomicAdd(&cnt[px * kGridRows + py], 1)] = i;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < per; k++) {
    const int c = tid * per + k;
    const int b = cellStart[c], e = cnt[c];
    for (int i = b + 1; i < e; i++) {
      const int key = items[i];
      int j = i - 1;
      while (j >= b && items[j] > key) { items[j + 1] = items[j]; j--; }
      items[j + 1] = key;
    }
  }
}

__device__ __forceinline__ int hamming_row(const uint4& a0, const uint4& a1, const uint4* __restrict__ row) {
  const uint4 b0 = row[0], b1 = row[1];
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// FILL = false: count[q] = number of candidates. FILL = true: write them at off[q].
template <bool FILL>
__global__ void __launch_bounds__(256)
window_kernel(const orbx_keypoint* __restrict__ kps, const uint4* __restrict__ desc, GridParams gp,
              const int* __restrict__ cellStart, const int* __restrict__ items, const float* __restrict__ qxy,
              const float* __restrict__ qr, const int* __restrict__ qMinLevel, const int* __restrict__ qMaxLevel,
              const uint4* __restrict__ qdesc, int nq, int* __restrict__ count, const int* __restrict__ off,
              int* __restrict__ candIdx, int* __restrict__ candDist) {
  const int q = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (q >= nq) return;
  const float x = qxy[2 * q], y = qxy[2 * q + 1], r = qr[q];
  const int minLevel = qMinLevel[q], maxLevel = qMaxLevel[q];
  // R/src/Frame.cpp:392-407
  const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
  const int nMaxCellX = min(kGridCols - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, gp.minX), r), gp.invW)));
  const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
  const int nMaxCellY = min(kGridRows - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, gp.minY), r), gp.invH)));
  int total = 0;
  if (nMinCellX < kGridCols && nMaxCellX >= 0 && nMinCellY < kGridRows && nMaxCellY >= 0) {
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    uint4 a0, a1;
    int base = 0;
    if (FILL) { a0 = qdesc[2 * q]; a1 = qdesc[2 * q + 1]; base = off[q]; }
    for (int ix = nMinCellX; ix <= nMaxCellX; ix++) {
      const int b = cellStart[ix * kGridRows + nMinCellY], e = cellStart[ix * kGridRows + nMaxCellY + 1];
      for (int i0 = b; i0 < e; i0 += 32) {
        const int i = i0 + lane;
        bool pass = false;
        int idx = 0;
        if (i < e) {
          idx = items[i];
          const orbx_keypoint kp = kps[idx];
          pass = true;
          if (bCheckLevels) {
            if (kp.octave < minLevel) pass = false;
            if (maxLevel >= 0 && kp.octave > maxLevel) pass = false;
          }
          const float distx = __fsub_rn(kp.x, x), disty = __fsub_rn(kp.y, y);
          pass = pass && fabsf(distx) < r && fabsf(disty) < r;
        }
        const unsigned m = __ballot_sync(0xffffffffu, pass);
        if (FILL && pass) {
          const int slot = base + total + __popc(m & ((1u << lane) - 1));
          candIdx[slot] = idx;
          candDist[slot] = hamming_row(a0, a1, desc + 2 * (long long)idx);
        }
        total += __popc(m);
      }
    }
  }
  if (!FILL && lane == 0) count[q] = total;
}

// exclusive scan of count[0..n) into off[0..n], off[n] = total (single CTA)
__global__ void __launch_bounds__(1024)
scan_kernel(const int* __restrict__ count, int n, int* __restrict__ off) {
  __shared__ int warpSums[32];
  __shared__ int carry;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int b = 0; b < n; b += 1024) {
    const int i = b + tid;
    const int c = i < n ? count[i] : 0;
    int v = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, v, o); if (lane >= o) v += t; }
    if (lane == 31) warpSums[warp] = v;
    __syncthreads();
    if (warp == 0) {
      int w = warpSums[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += t; }
      warpSums[lane] = w;
    }
    __syncthreads();
    const int excl = carry + v - c + (warp ? warpSums[warp - 1] : 0);
    if (i < n) off[i] = excl;
    __syncthreads();
    if (tid == 1023) carry = excl + c;
    __syncthreads();
  }
  if (tid == 0) off[n] = carry;
}

void launch_grid_build(const orbx_keypoint* kps, int n, float minX, float maxX, float minY, float maxY, int* cellStart,
                       int* items, cudaStream_t s) {
  GridParams gp;
  gp.minX = minX; gp.minY = minY;
  gp.invW = (float)kGridCols / (maxX - minX);  // R/src/Frame.cpp:168-169
  gp.invH = (float)kGridRows / (maxY - minY);
  grid_build_kernel<<<1, 1024, 0, s>>>(kps, n, gp, cellStart, items);
}

void launch_window(bool fill, const orbx_keypoint* kps, const uint8_t* desc, float minX, float maxX, float minY, float maxY,
                   const int* cellStart, const int* items, const float* qxy, const float* qr, const int* qMinLevel,
                   const int* qMaxLevel, const uint8_t* qdesc, int nq, int* count, const int* off, int* candIdx,
                   int* candDist, cudaStream_t s) {
  GridParams gp;
  gp.minX = minX; gp.minY = minY;
  gp.invW = (float)kGridCols / (maxX - minX);
  gp.invH = (float)kGridRows / (maxY - minY);
  const int blocks = (nq + 7) / 8;
  if (fill)
    window_kernel<true><<<blocks, 256, 0, s>>>(kps, reinterpret_cast<const uint4*>(desc), gp, cellStart, items, qxy, qr,
                                               qMinLevel, qMaxLevel, reinterpret_cast<const uint4*>(qdesc), nq, count, off,
                                               candIdx, candDist);
  else
    window_kernel<false><<<blocks, 256, 0, s>>>(kps, reinterpret_cast<const uint4*>(desc), gp, cellStart, items, qxy, qr,
                                                qMinLevel, qMaxLevel, reinterpret_cast<const uint4*>(qdesc), nq, count, off,
                                                candIdx, candDist);
}

void launch_scan(const int* count, int n, int* off, cudaStream_t s) { scan_kernel<<<1, 1024, 0, s>>>(count, n, off); }

int grid_cells() { return kGridCells; }

}  // namespace orbx
