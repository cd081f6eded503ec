// Frame::ComputeStereoMatches on the device (R/src/Frame.cpp:551-770), for one pair or a
// batch of pairs (grid.y = pair).
//
//  stereo_rows_kernel   : the row table vRowIndices (:558-589) as a CSR over image rows: right keypoint iR is listed
//                         in rows floor(yR - 2*scale[octR]) .. ceil(yR + 2*scale[octR]). One CTA per pair.
//  stereo_match_kernel  : one warp per left keypoint.
//     1. candidates = the list of row (int)yL (:604-641), kept iff |octave difference| <= 1 and
//        uL - maxD <= uR <= uL. The reference scans the list in ascending iR and keeps the FIRST minimum of
//        DescriptorDistance below TH_HIGH: that is min over (dist << 16 | iR), so the list order is free.
//     2. SAD refinement (:662-713): 11x11 patch of the left level against 11 horizontal offsets of the right
//        level, each patch minus its own centre pixel; all values are small integers, exact in float.
//     3. parabola fit, disparity, depth (:719-750) by lane 0 with explicit round-to-nearest float ops.
//  stereo_median_kernel : the global filter (:752-769): drop matches whose SAD is >= 1.5*1.4*median.
#include <climits>

#include "orbx_common.cuh"
#include "orbx_stereo.cuh"

namespace orbx {

constexpr int kThHigh = 100, kThLow = 50;  // R/src/ORBmatcher.cpp:37-38
constexpr int kMedianSmem = 4096;           // SADs ranked in shared memory by the median filter

__device__ __forceinline__ int stereo_count(const int* n, int frame, int fixed, int cap) {
  return n ? min(n[frame], cap) : fixed;
}

__global__ void __launch_bounds__(1024)
stereo_rows_kernel(const __grid_constant__ StereoLevels lv, const __grid_constant__ StereoArrays a) {
  extern __shared__ int sRows[];  // rows0 counters, then cursors
  __shared__ int warpSums[32];
  __shared__ int carry;
  const int p = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int frameR = lv.arrRight0 + p * lv.arrStep;
  const int nr = stereo_count(a.nR, frameR, lv.fixedNR, lv.cap);
  const orbx_keypoint* kpR = a.kpR + (long long)frameR * lv.cap;
  int* rowStart = a.rowStart + (long long)p * (lv.rows0 + 1);
  int* items = a.rowItems + (long long)p * a.itemCap;
  const int rows = lv.rows0;
  for (int i = tid; i < rows; i += 1024) sRows[i] = 0;
  if (tid == 0) carry = 0;
  __syncthreads();
  for (int iR = tid; iR < nr; iR += 1024) {
    const orbx_keypoint r = kpR[iR];
    const float rad = __fmul_rn(2.0f, lv.scale[r.octave]);
    const int maxr = min((int)ceilf(__fadd_rn(r.y, rad)), rows - 1), minr = max((int)floorf(__fsub_rn(r.y, rad)), 0);
    for (int y = minr; y <= maxr; y++) atomicAdd(&sRows[y], 1);
  }
  __syncthreads();
  // exclusive scan of the row counts, 1024 rows per trip
  for (int b = 0; b < rows; b += 1024) {
    const int i = b + tid;
    const int c = i < rows ? sRows[i] : 0;
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
    if (i < rows) { rowStart[i] = excl; sRows[i] = excl; }
    __syncthreads();
    if (tid == 1023) carry = excl + c;
    __syncthreads();
  }
  if (tid == 0) rowStart[rows] = carry;
  for (int iR = tid; iR < nr; iR += 1024) {
    const orbx_keypoint r = kpR[iR];
    const float rad = __fmul_rn(2.0f, lv.scale[r.octave]);
    const int maxr = min((int)ceilf(__fadd_rn(r.y, rad)), rows - 1), minr = max((int)floorf(__fsub_rn(r.y, rad)), 0);
    for (int y = minr; y <= maxr; y++) {
      const int pos = atomicAdd(&sRows[y], 1);
      if (pos < a.itemCap) items[pos] = iR;  // itemCap is sized on the host so that this always holds
    }
  }
}

__global__ void __launch_bounds__(256)
stereo_match_kernel(const __grid_constant__ StereoLevels lv, const __grid_constant__ StereoArrays a, float mbf, float mb) {
  const int p = blockIdx.y;
  const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int frameL = lv.arrLeft0 + p * lv.arrStep, frameR = lv.arrRight0 + p * lv.arrStep;
  const int nl = stereo_count(a.nL, frameL, lv.fixedNL, lv.cap);
  if (iL >= nl) return;
  const orbx_keypoint* kpL = a.kpL + (long long)frameL * lv.cap;
  const orbx_keypoint* kpR = a.kpR + (long long)frameR * lv.cap;
  const uint4* descL = a.descL + 2LL * frameL * lv.cap;
  const uint4* descR = a.descR + 2LL * frameR * lv.cap;
  float* uRight = a.uRight + (long long)p * lv.cap;
  float* depth = a.depth + (long long)p * lv.cap;
  int* sad = a.sad + (long long)p * lv.cap;
  const int* rowStart = a.rowStart + (long long)p * (lv.rows0 + 1);
  const int* items = a.rowItems + (long long)p * a.itemCap;

  if (lane == 0) { uRight[iL] = -1.0f; depth[iL] = -1.0f; sad[iL] = -1; }
  const orbx_keypoint k = kpL[iL];
  const int levelL = k.octave;
  const float vL = k.y, uL = k.x;
  const int row = (int)vL;  // vRowIndices[vL]
  const float maxD = __fdiv_rn(mbf, mb);
  const float minU = __fsub_rn(uL, maxD), maxU = uL;  // minD = 0
  if (maxU < 0 || row < 0 || row >= lv.rows0) return;
  const uint4 a0 = descL[2 * iL], a1 = descL[2 * iL + 1];
  unsigned best = 0xffffffffu;
  const int jEnd = min(rowStart[row + 1], a.itemCap);
  for (int j = rowStart[row] + lane; j < jEnd; j += 32) {
    const int iR = items[j];
    const orbx_keypoint r = kpR[iR];
    if (r.octave < levelL - 1 || r.octave > levelL + 1) continue;
    if (r.x >= minU && r.x <= maxU) {
      const uint4 b0 = descR[2 * iR], b1 = descR[2 * iR + 1];
      const int d = __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
                    __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
      best = min(best, ((unsigned)d << 16) | (unsigned)iR);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  const int bestDist = (int)(best >> 16);
  // dist < bestDist(=TH_HIGH) to be taken at all, then bestDist < thOrbDist (:625, :662)
  if (best == 0xffffffffu || bestDist >= kThHigh || bestDist >= (kThHigh + kThLow) / 2) return;
  const int bestIdxR = (int)(best & 0xffffu);

  const float uR0 = kpR[bestIdxR].x;
  const float sf = lv.invScale[levelL];
  const float scaleduL = roundf(__fmul_rn(k.x, sf));
  const float scaledvL = roundf(__fmul_rn(k.y, sf));
  const float scaleduR0 = roundf(__fmul_rn(uR0, sf));
  constexpr int w = 5, L = 5;
  const float iniu = scaleduR0 + (float)(L - w), endu = scaleduR0 + (float)(L + w + 1);
  if (iniu < 0 || endu >= (float)lv.width[levelL]) return;
  const int y0 = (int)(scaledvL - w), xL0 = (int)(scaleduL - w);
  const uint8_t* IL = lv.left[levelL] + (long long)(lv.pyrLeft0 + p * lv.pyrStep) * lv.strideL[levelL];
  const uint8_t* IR = lv.right[levelL] + (long long)(lv.pyrRight0 + p * lv.pyrStep) * lv.strideR[levelL];
  const int pL = lv.pitchL[levelL], pR = lv.pitchR[levelL];
  const int cL = IL[(long long)(y0 + w) * pL + xL0 + w];
  // left patch values of this lane's pixels (121 pixels, 4 per lane)
  int lp[4], py[4], px[4];
#pragma unroll
  for (int t = 0; t < 4; t++) {
    const int q = lane + 32 * t;
    py[t] = q / 11; px[t] = q - py[t] * 11;
    lp[t] = q < 121 ? (int)IL[(long long)(y0 + py[t]) * pL + xL0 + px[t]] - cL : 0;
  }
  int bestSad = INT_MAX, bestincR = 0;
  int d1 = 0, d2 = 0, d3 = 0, prev = 0;  // SADs around the running best
  bool wantNext = false;
  for (int incR = -L; incR <= L; incR++) {
    const int xR0 = (int)(scaleduR0 + (float)incR - (float)w);
    const int cR = IR[(long long)(y0 + w) * pR + xR0 + w];
    int acc = 0;
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const int q = lane + 32 * t;
      if (q < 121) acc += abs(lp[t] - ((int)IR[(long long)(y0 + py[t]) * pR + xR0 + px[t]] - cR));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (wantNext) { d3 = acc; wantNext = false; }
    if (acc < bestSad) {  // strict: the first minimum wins (:699-703)
      bestSad = acc; bestincR = incR; d1 = prev; d2 = acc; wantNext = true;
    }
    prev = acc;
  }
  if (bestincR == -L || bestincR == L) return;
  if (lane == 0) {
    const float dist1 = (float)d1, dist2 = (float)d2, dist3 = (float)d3;
    const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
    if (deltaR < -1 || deltaR > 1) return;  // also false for NaN (0/0), like the reference's comparison
    float bestuR = __fmul_rn(lv.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= 0 && disparity < maxD) {
      if (disparity <= 0) { disparity = 0.01f; bestuR = __double2float_rn(__dsub_rn((double)uL, 0.01)); }  // :742-743: uL-0.01 is a double expression
      depth[iL] = __fdiv_rn(mbf, disparity);
      uRight[iL] = bestuR;
      sad[iL] = bestSad;
    }
  }
}

// (:752-769) median of the SADs in (sad, iL) order, threshold 1.5f*1.4f*median, drop sad >= threshold. One CTA per pair.
__global__ void __launch_bounds__(1024)
stereo_median_kernel(const __grid_constant__ StereoLevels lv, const __grid_constant__ StereoArrays a) {
  __shared__ int sCount, sMedian;
  const int p = blockIdx.x;
  const int tid = threadIdx.x, T = blockDim.x;
  const int nl = stereo_count(a.nL, lv.arrLeft0 + p * lv.arrStep, lv.fixedNL, lv.cap);
  float* uRight = a.uRight + (long long)p * lv.cap;
  float* depth = a.depth + (long long)p * lv.cap;
  const int* sad = a.sad + (long long)p * lv.cap;
  if (tid == 0) { sCount = 0; sMedian = -1; }
  __syncthreads();
  int c = 0;
  for (int i = tid; i < nl; i += T) c += sad[i] >= 0;
  if (c) atomicAdd(&sCount, c);
  __syncthreads();
  const int size = sCount;
  if (size == 0) { if (tid == 0) a.nKept[p] = 0; return; }
  const int k = size / 2;  // vDistIdx[size/2] after the sort
  // the median is the value v with #(t < v) <= k < #(t <= v). Usually a few hundred of the nl keypoints carry a
  // SAD: they are compacted into shared memory first and ranked among themselves.
  __shared__ int sVals[kMedianSmem];
  __shared__ int sFill;
  if (tid == 0) sFill = 0;
  __syncthreads();
  if (size <= kMedianSmem) {
    for (int i = tid; i < nl; i += T) {
      const int s = sad[i];
      if (s >= 0) sVals[atomicAdd(&sFill, 1)] = s;
    }
    __syncthreads();
    for (int i = tid; i < size; i += T) {
      const int s = sVals[i];
      int lo = 0, hi = 0;
      for (int j = 0; j < size; j++) {
        const int t = sVals[j];
        lo += t < s;
        hi += t <= s;
      }
      if (lo <= k && k < hi) sMedian = s;  // every writer writes the same value
    }
  } else {
    for (int i = tid; i < nl; i += T) {
      const int s = sad[i];
      if (s < 0) continue;
      int lo = 0, hi = 0;
      for (int j = 0; j < nl; j++) {
        const int t = sad[j];
        if (t < 0) continue;
        lo += t < s;
        hi += t <= s;
      }
      if (lo <= k && k < hi) sMedian = s;
    }
  }
  __syncthreads();
  const float thDist = __fmul_rn(1.5f * 1.4f, (float)sMedian);
  if (tid == 0) sCount = 0;
  __syncthreads();
  int kept = 0;
  for (int i = tid; i < nl; i += T) {
    const int s = sad[i];
    if (s < 0) continue;
    if ((float)s < thDist) kept++;
    else { uRight[i] = -1.0f; depth[i] = -1.0f; }
  }
  if (kept) atomicAdd(&sCount, kept);
  __syncthreads();
  if (tid == 0) a.nKept[p] = sCount;
}

// rows of the row table one right keypoint can occupy: 2*ceil(2*maxScale) + 3
int stereo_items_per_keypoint(float maxScale) { return 2 * (int)ceilf(2.0f * maxScale) + 3; }

// `maxNL` bounds the left keypoint count of any pair (grid sizing)
int launch_stereo(const StereoLevels& lv, const StereoArrays& a, int pairs, int maxNL, float mbf, float mb, cudaStream_t s) {
  if (pairs <= 0) return 0;
  const size_t smem = (size_t)lv.rows0 * sizeof(int);
  static SmemLimit limit;
  cudaError_t e = raise_dynamic_smem(reinterpret_cast<const void*>(stereo_rows_kernel), smem, &limit);
  if (e != cudaSuccess) return (int)e;
  stereo_rows_kernel<<<pairs, 1024, smem, s>>>(lv, a);
  if (maxNL > 0) stereo_match_kernel<<<dim3((maxNL + 7) / 8, pairs), 256, 0, s>>>(lv, a, mbf, mb);
  stereo_median_kernel<<<pairs, 1024, 0, s>>>(lv, a);
  return 0;
}

}  // namespace orbx
