// Frame::ComputeStereoMatches on the device (R/src/Frame.cpp:551-770).
//
// Every left keypoint is independent until the final median filter, so one warp owns one left keypoint:
//  1. row-band candidates (:568-589, :604-641): right keypoint iR is a candidate iff
//     floor(yR - 2*scale[octR]) <= (int)yL <= ceil(yR + 2*scale[octR]), |octave difference| <= 1 and
//     uL - maxD <= uR <= uL. The reference scans the row table in ascending iR and keeps the FIRST minimum
//     of DescriptorDistance below TH_HIGH: that is min over (dist << 16 | iR), order free.
//  2. SAD refinement (:662-713): 11x11 patch of the left level against 11 horizontal offsets of the right
//     level, each patch minus its own centre pixel; all values are small integers, exact in float.
//  3. parabola fit, disparity, depth (:719-750) by lane 0 with explicit round-to-nearest float ops.
// stereo_median_kernel applies the global filter (:756-769): drop matches whose SAD is >= 1.5*1.4*median.
#include <climits>

#include "orbx_common.cuh"

namespace orbx {

constexpr int kThHigh = 100, kThLow = 50;  // R/src/ORBmatcher.cpp:37-38

struct StereoLevels {
  const uint8_t* left[kMaxLevels];
  const uint8_t* right[kMaxLevels];
  int pitchL[kMaxLevels], pitchR[kMaxLevels];
  int width[kMaxLevels];
  float scale[kMaxLevels], invScale[kMaxLevels];
  int rows0;  // mvImagePyramid[0].rows
};

__global__ void __launch_bounds__(256)
stereo_match_kernel(const __grid_constant__ StereoLevels lv, const orbx_keypoint* __restrict__ kpL,
                    const uint4* __restrict__ descL, int nl, const orbx_keypoint* __restrict__ kpR,
                    const uint4* __restrict__ descR, int nr, float mbf, float mb, float* __restrict__ uRight,
                    float* __restrict__ depth, int* __restrict__ sad) {
  const int iL = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (iL >= nl) return;
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
  for (int iR = lane; iR < nr; iR += 32) {
    const orbx_keypoint r = kpR[iR];
    const float rad = __fmul_rn(2.0f, lv.scale[r.octave]);
    const int maxr = (int)ceilf(__fadd_rn(r.y, rad)), minr = (int)floorf(__fsub_rn(r.y, rad));
    if (row < minr || row > maxr) continue;
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
  const uint8_t* IL = lv.left[levelL];
  const uint8_t* IR = lv.right[levelL];
  const int pL = lv.pitchL[levelL], pR = lv.pitchR[levelL];
  const int cL = IL[(long long)(y0 + w) * pL + xL0 + w];
  // left patch values of this lane's pixels (121 pixels, 4 per lane)
  int lp[4], py[4], px[4];
#pragma unroll
  for (int t = 0; t < 4; t++) {
    const int p = lane + 32 * t;
    py[t] = p / 11; px[t] = p - py[t] * 11;
    lp[t] = p < 121 ? (int)IL[(long long)(y0 + py[t]) * pL + xL0 + px[t]] - cL : 0;
  }
  int bestSad = INT_MAX, bestincR = 0;
  int d1 = 0, d2 = 0, d3 = 0, prev = 0, prevprev = 0;  // SADs around the running best
  bool wantNext = false;
  for (int incR = -L; incR <= L; incR++) {
    const int xR0 = (int)(scaleduR0 + (float)incR - (float)w);
    const int cR = IR[(long long)(y0 + w) * pR + xR0 + w];
    int acc = 0;
#pragma unroll
    for (int t = 0; t < 4; t++) {
      const int p = lane + 32 * t;
      if (p < 121) acc += abs(lp[t] - ((int)IR[(long long)(y0 + py[t]) * pR + xR0 + px[t]] - cR));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (wantNext) { d3 = acc; wantNext = false; }
    if (acc < bestSad) {  // strict: the first minimum wins (:699-703)
      bestSad = acc; bestincR = incR; d1 = prev; d2 = acc; wantNext = true;
    }
    prevprev = prev; prev = acc;
  }
  (void)prevprev;
  if (bestincR == -L || bestincR == L) return;
  if (lane == 0) {
    const float dist1 = (float)d1, dist2 = (float)d2, dist3 = (float)d3;
    const float den = __fmul_rn(2.0f, __fsub_rn(__fadd_rn(dist1, dist3), __fmul_rn(2.0f, dist2)));
    const float deltaR = __fdiv_rn(__fsub_rn(dist1, dist3), den);
    if (deltaR < -1 || deltaR > 1) return;  // also false for NaN (0/0), like the reference's comparison
    float bestuR = __fmul_rn(lv.scale[levelL], __fadd_rn(__fadd_rn(scaleduR0, (float)bestincR), deltaR));
    float disparity = __fsub_rn(uL, bestuR);
    if (disparity >= 0 && disparity < maxD) {
      if (disparity <= 0) { disparity = 0.01f; bestuR = __fsub_rn(uL, 0.01f); }
      depth[iL] = __fdiv_rn(mbf, disparity);
      uRight[iL] = bestuR;
      sad[iL] = bestSad;
    }
  }
}

// (:752-769) median of the SADs in (sad, iL) order, threshold 1.5f*1.4f*median, drop sad >= threshold.
__global__ void __launch_bounds__(1024)
stereo_median_kernel(int nl, float* __restrict__ uRight, float* __restrict__ depth, const int* __restrict__ sad,
                     int* __restrict__ nKept) {
  __shared__ int sCount, sMedian;
  const int tid = threadIdx.x, T = blockDim.x;
  if (tid == 0) { sCount = 0; sMedian = -1; }
  __syncthreads();
  int c = 0;
  for (int i = tid; i < nl; i += T) c += sad[i] >= 0;
  if (c) atomicAdd(&sCount, c);
  __syncthreads();
  const int size = sCount;
  if (size == 0) { if (tid == 0) *nKept = 0; return; }
  const int k = size / 2;  // vDistIdx[size/2] after the sort
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
    if (lo <= k && k < hi) sMedian = s;  // every writer writes the same value
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
  if (tid == 0) *nKept = sCount;
}

void launch_stereo(const StereoLevels& lv, const orbx_keypoint* kpL, const uint8_t* descL, int nl,
                   const orbx_keypoint* kpR, const uint8_t* descR, int nr, float mbf, float mb, float* uRight,
                   float* depth, int* sad, int* nKept, cudaStream_t s) {
  if (nl > 0)
    stereo_match_kernel<<<(nl + 7) / 8, 256, 0, s>>>(lv, kpL, reinterpret_cast<const uint4*>(descL), nl, kpR,
                                                     reinterpret_cast<const uint4*>(descR), nr, mbf, mb, uRight, depth, sad);
  stereo_median_kernel<<<1, 1024, 0, s>>>(nl, uRight, depth, sad, nKept);
}

}  // namespace orbx
