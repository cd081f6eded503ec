// Node-restricted Hamming blocks and distinctive-descriptor selection.
//
// block_distance_kernel: the data-parallel half of ORBmatcher::SearchByBoW (R/src/ORBmatcher.cpp:220-365,
// :632-772) and SearchForTriangulation (:785-983). Those loops walk the vocabulary nodes two feature vectors
// share and, inside one node, compare every feature of the first list with every feature of the second. The
// device fills the distance matrix of every shared node; the order-dependent bookkeeping (already-matched
// flags, running best / second best, rotation histogram) stays on the host (host/orbx_search.cc).
// One warp per (node, first-list entry): the 32-byte row of that entry sits in registers, lanes stride over the
// node's second list, 8 POPC per pair, distances written as uint16 (0..256) row-major inside the node's matrix.
//
// distinctive_kernel: MapPoint::ComputeDistinctiveDescriptors (R/src/MapPoint.cpp:306-385) for many map points
// in one launch. Per point: all-pairs distances among its N observed descriptors, per row the median
// vDists[0.5*(N-1)] of the sorted row (self distance 0 included), keep the first row with the least median.
// One CTA per point, one warp per row at a time; the median is read off a 257-bin histogram of the row built
// with shared-memory atomics and scanned with ballots, so no N x N matrix is stored and N is unbounded.
#include "orbx_common.cuh"

namespace orbx {

__device__ __forceinline__ int hamming_rows(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

constexpr int kBlockWarps = 8;

__global__ void __launch_bounds__(kBlockWarps * 32)
block_distance_kernel(const uint4* __restrict__ descA, const uint4* __restrict__ descB, const int* __restrict__ aOff,
                      const int* __restrict__ aIdx, const int* __restrict__ bOff, const int* __restrict__ bIdx,
                      const int* __restrict__ entryBlock, const long long* __restrict__ dOff, int nEntries,
                      unsigned short* __restrict__ dist) {
  const int lane = threadIdx.x & 31;
  const int e = blockIdx.x * kBlockWarps + (threadIdx.x >> 5);
  if (e >= nEntries) return;
  const int k = entryBlock[e];
  const int b0 = bOff[k], nb = bOff[k + 1] - b0;
  const int ia = e - aOff[k];
  const int ra = aIdx[e];
  const uint4 a0 = descA[2 * ra], a1 = descA[2 * ra + 1];
  unsigned short* out = dist + dOff[k] + (long long)ia * nb;
  for (int j = lane; j < nb; j += 32) {
    const int rb = bIdx[b0 + j];
    out[j] = (unsigned short)hamming_rows(a0, a1, descB[2 * rb], descB[2 * rb + 1]);
  }
}

void launch_block_distance(const uint8_t* descA, const uint8_t* descB, const int* aOff, const int* aIdx, const int* bOff,
                           const int* bIdx, const int* entryBlock, const long long* dOff, int nEntries,
                           unsigned short* dist, cudaStream_t s) {
  if (nEntries <= 0) return;
  const int grid = (nEntries + kBlockWarps - 1) / kBlockWarps;
  block_distance_kernel<<<grid, kBlockWarps * 32, 0, s>>>(reinterpret_cast<const uint4*>(descA),
                                                          reinterpret_cast<const uint4*>(descB), aOff, aIdx, bOff, bIdx,
                                                          entryBlock, dOff, nEntries, dist);
}

constexpr int kDistinctWarps = 4;
constexpr int kHistBins = 288;  // 257 used, padded to 9 x 32

__global__ void __launch_bounds__(kDistinctWarps * 32)
distinctive_kernel(const uint4* __restrict__ desc, const int* __restrict__ obsOff, int nPoints, int* __restrict__ best,
                   int* __restrict__ bestMedian) {
  __shared__ unsigned hist[kDistinctWarps][kHistBins];
  __shared__ unsigned bestKey[kDistinctWarps];
  const int p = blockIdx.x;
  if (p >= nPoints) return;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int r0 = obsOff[p], N = obsOff[p + 1] - r0;
  if (N <= 0) {  // no observation: the reference returns before touching mDescriptor (:320-321, :335-336)
    if (threadIdx.x == 0) { best[p] = -1; bestMedian[p] = -1; }
    return;
  }
  const int rank = (int)(0.5 * (N - 1));  // vDists[0.5*(N-1)] (:366): the double is truncated by the index conversion
  unsigned key = 0xFFFFFFFFu;             // median << 20 | row: the first row with the least median wins (:369)
  for (int i = warp; i < N; i += kDistinctWarps) {
    for (int b = lane; b < kHistBins; b += 32) hist[warp][b] = 0;
    __syncwarp();
    const uint4 a0 = desc[2 * (r0 + i)], a1 = desc[2 * (r0 + i) + 1];
    for (int j = lane; j < N; j += 32) {
      const int d = hamming_rows(a0, a1, desc[2 * (r0 + j)], desc[2 * (r0 + j) + 1]);
      atomicAdd(&hist[warp][d], 1u);
    }
    __syncwarp();
    // the value at sorted position `rank`: first bin whose inclusive prefix count exceeds rank
    int before = 0, median = 256;
    for (int b0 = 0; b0 < kHistBins; b0 += 32) {
      const int c = (int)hist[warp][b0 + lane];
      int incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (lane >= o) incl += t;
      }
      const unsigned hit = __ballot_sync(0xffffffffu, before + incl > rank);
      if (hit) { median = b0 + __ffs(hit) - 1; break; }
      before += __shfl_sync(0xffffffffu, incl, 31);
    }
    const unsigned k = ((unsigned)median << 20) | (unsigned)i;
    key = min(key, k);
    __syncwarp();
  }
  if (lane == 0) bestKey[warp] = key;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned k = bestKey[0];
    for (int w = 1; w < kDistinctWarps; w++) k = min(k, bestKey[w]);
    best[p] = (int)(k & 0xFFFFFu);
    bestMedian[p] = (int)(k >> 20);
  }
}

void launch_distinctive(const uint8_t* desc, const int* obsOff, int nPoints, int* best, int* bestMedian, cudaStream_t s) {
  if (nPoints <= 0) return;
  distinctive_kernel<<<nPoints, kDistinctWarps * 32, 0, s>>>(reinterpret_cast<const uint4*>(desc), obsOff, nPoints, best,
                                                             bestMedian);
}

}  // namespace orbx

// ---- vocabulary descent ---------------------------------------------------------------------
// DBoW2::TemplatedVocabulary::transform(feature, word_id, weight, nid, levelsup)
// (R/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1242-1296), what Frame::ComputeBoW / KeyFrame::ComputeBoW run once
// per descriptor (R/src/Frame.cpp:262-274, R/src/KeyFrame.cpp:66-75): from the root, move to the child with the
// least FORB::distance (first minimum wins, :1268-1278) until a leaf; remember the node passed at level L - levelsup.
// One warp per descriptor: its row in registers, lane j scores child j (k <= 20 in the text format), the arg-min is
// a warp reduction of (distance << 8 | child position). The tree (32 bytes per node, ~35 MB for ORBvoc's 1.08 M
// nodes) stays resident in HBM / L2; a descent reads L x k rows.
namespace orbx {

__global__ void __launch_bounds__(kBlockWarps * 32)
voc_descend_kernel(const uint4* __restrict__ feat, int n, const uint4* __restrict__ nodeDesc,
                   const int* __restrict__ childOff, const int* __restrict__ childIdx, int nidLevel,
                   int* __restrict__ leafNode, int* __restrict__ levelNode) {
  const int lane = threadIdx.x & 31;
  const int f = blockIdx.x * kBlockWarps + (threadIdx.x >> 5);
  if (f >= n) return;
  const uint4 a0 = feat[2 * f], a1 = feat[2 * f + 1];
  int node = 0, level = 0;
  int nid = nidLevel <= 0 ? 0 : -1;   // -1: a leaf above the requested level (the reference leaves *nid unset there)
  for (;;) {
    const int c0 = childOff[node], nc = childOff[node + 1] - c0;
    if (nc == 0) break;
    ++level;
    unsigned best = 0xFFFFFFFFu;
    for (int j = lane; j < nc; j += 32) {
      const int c = childIdx[c0 + j];
      const unsigned d = (unsigned)hamming_rows(a0, a1, nodeDesc[2 * c], nodeDesc[2 * c + 1]);
      best = min(best, (d << 8) | (unsigned)j);   // nc <= 256 children per node
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    node = childIdx[c0 + (int)(best & 0xFFu)];
    if (level == nidLevel) nid = node;
  }
  if (lane == 0) { leafNode[f] = node; levelNode[f] = nid; }
}

void launch_voc_descend(const uint8_t* feat, int n, const uint8_t* nodeDesc, const int* childOff, const int* childIdx,
                        int nidLevel, int* leafNode, int* levelNode, cudaStream_t s) {
  if (n <= 0) return;
  voc_descend_kernel<<<(n + kBlockWarps - 1) / kBlockWarps, kBlockWarps * 32, 0, s>>>(
      reinterpret_cast<const uint4*>(feat), n, reinterpret_cast<const uint4*>(nodeDesc), childOff, childIdx, nidLevel,
      leafNode, levelNode);
}

}  // namespace orbx
