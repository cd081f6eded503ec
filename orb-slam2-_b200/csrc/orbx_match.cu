// Hamming matching kernels (R/src/ORBmatcher.cpp DescriptorDistance
// :1901-1917 and the running best / second-best loops :266-300, :526-556).
//
// Brute-force top-2: bound by the integer pipe (8 XOR + 8 POPC per pair), not by HBM: a 32-byte
// database row is read once per CTA from global memory and reused by 1024 queries from shared
// memory (broadcast LDS), queries live in registers. No tensor cores: Hamming distance is not a
// dense contraction and sm_100a has no native 1-bit MMA.
//
// The sequential rule "if(d<best){best2=best;best=d;idx=i} else if(d<best2) best2=d" over rows in
// index order is a (distance, index)-ordered top-2 of a multiset, which is associative. Each thread
// keeps the two smallest 32-bit keys  key = distance<<23 | rowInChunk  (unsigned min/max only):
// the smallest key is the first minimum, the second smallest key carries the second-smallest
// distance counting duplicates. Chunks and database shards merge with the same rule.
#include <climits>

#include "orbx_common.cuh"

namespace orbx {

constexpr int kBfThreads = 256;
constexpr int kBfQPT = 4;                       // queries per thread
constexpr int kBfQueriesPerCta = kBfThreads * kBfQPT;
constexpr int kBfTileRows = 512;                // database rows staged per step (16 KB)
constexpr unsigned kKeyNone = 0xFFFFFFFFu;

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// grid: (query groups, chunks). partial arrays are [chunk][nq].
__global__ void __launch_bounds__(kBfThreads)
bf2_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb, int rowsPerChunk,
           long long idxBase, int* __restrict__ pIdx, int* __restrict__ pD1, int* __restrict__ pD2) {
  __shared__ uint4 tile[kBfTileRows * 2];
  const int tid = threadIdx.x;
  const int qBase = blockIdx.x * kBfQueriesPerCta;
  const long long rowBegin = (long long)blockIdx.y * rowsPerChunk;
  const long long rowEnd = min(rowBegin + rowsPerChunk, ndb);

  uint4 qa[kBfQPT], qb[kBfQPT];
  unsigned k1[kBfQPT], k2[kBfQPT];
#pragma unroll
  for (int k = 0; k < kBfQPT; k++) {
    const int qi = min(qBase + tid + k * kBfThreads, nq - 1);
    qa[k] = q[2 * (long long)qi];
    qb[k] = q[2 * (long long)qi + 1];
    k1[k] = kKeyNone;
    k2[k] = kKeyNone;
  }

  for (long long t0 = rowBegin; t0 < rowEnd; t0 += kBfTileRows) {
    const int rows = (int)min((long long)kBfTileRows, rowEnd - t0);
    __syncthreads();
    for (int i = tid; i < rows * 2; i += kBfThreads) tile[i] = db[2 * t0 + i];
    __syncthreads();
    unsigned rowKey = (unsigned)(t0 - rowBegin);
#pragma unroll 2
    for (int r = 0; r < rows; r++, rowKey++) {
      const uint4 d0 = tile[2 * r], d1 = tile[2 * r + 1];
#pragma unroll
      for (int k = 0; k < kBfQPT; k++) {
        const unsigned key = ((unsigned)hamming256(qa[k], qb[k], d0, d1) << 23) | rowKey;
        k2[k] = min(k2[k], max(k1[k], key));
        k1[k] = min(k1[k], key);
      }
    }
  }
#pragma unroll
  for (int k = 0; k < kBfQPT; k++) {
    const int qi = qBase + tid + k * kBfThreads;
    if (qi < nq) {
      const long long o = (long long)blockIdx.y * nq + qi;
      pIdx[o] = k1[k] == kKeyNone ? -1 : (int)(idxBase + rowBegin + (k1[k] & 0x7FFFFFu));
      pD1[o] = k1[k] == kKeyNone ? INT_MAX : (int)(k1[k] >> 23);
      pD2[o] = k2[k] == kKeyNone ? INT_MAX : (int)(k2[k] >> 23);
    }
  }
}

// Merge nshards partial results ([shard][nq], shards in ascending index order); optional ratio test.
__global__ void __launch_bounds__(256)
merge_kernel(const int* __restrict__ pIdx, const int* __restrict__ pD1, const int* __restrict__ pD2, int nshards,
             int nq, float ratio, int th, int* __restrict__ oIdx, int* __restrict__ oD1, int* __restrict__ oD2,
             uint8_t* __restrict__ accepted) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  int best = INT_MAX, best2 = INT_MAX, idx = -1;
  for (int s = 0; s < nshards; s++) {
    const long long o = (long long)s * nq + qi;
    const int a1 = pD1[o], a2 = pD2[o];
    if (a1 < best) {
      best2 = min(best, a2);
      best = a1;
      idx = pIdx[o];
    } else {
      best2 = min(best2, a1);
    }
  }
  oIdx[qi] = idx;
  oD1[qi] = best;
  oD2[qi] = best2;
  if (accepted)  // R/src/ORBmatcher.cpp:554-556: bestDist<=TH && bestDist < (float)bestDist2*mfNNratio
    accepted[qi] = (best <= th && (float)best < __fmul_rn((float)best2, ratio)) ? 1 : 0;
}

__global__ void distance_kernel(const uint4* __restrict__ a, const uint4* __restrict__ b, int* __restrict__ d,
                                long long n) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) d[i] = hamming256(a[2 * i], a[2 * i + 1], b[2 * i], b[2 * i + 1]);
}

int bf2_plan_chunks(int nq, long long ndb, int smCount, int* rowsPerChunk) {
  const int qGroups = (nq + kBfQueriesPerCta - 1) / kBfQueriesPerCta;
  long long chunks = (8LL * smCount + qGroups - 1) / qGroups;  // ~8 CTAs per SM in flight overall
  const long long maxChunks = (ndb + kBfTileRows - 1) / kBfTileRows;
  chunks = max(1LL, min(chunks, maxChunks));
  long long rpc = (ndb + chunks - 1) / chunks;
  rpc = (rpc + kBfTileRows - 1) / kBfTileRows * kBfTileRows;
  rpc = min(rpc, (long long)(1 << 23));
  *rowsPerChunk = (int)max(rpc, (long long)kBfTileRows);
  return (int)((ndb + *rowsPerChunk - 1) / *rowsPerChunk);
}

void launch_bf2(const uint8_t* q, int nq, const uint8_t* db, long long ndb, int rowsPerChunk, int chunks,
                long long idxBase, int* pIdx, int* pD1, int* pD2, cudaStream_t s) {
  dim3 grid((nq + kBfQueriesPerCta - 1) / kBfQueriesPerCta, chunks);
  bf2_kernel<<<grid, kBfThreads, 0, s>>>(reinterpret_cast<const uint4*>(q), nq, reinterpret_cast<const uint4*>(db), ndb,
                                         rowsPerChunk, idxBase, pIdx, pD1, pD2);
}

void launch_merge(const int* pIdx, const int* pD1, const int* pD2, int nshards, int nq, float ratio, int th,
                  int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s) {
  merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(pIdx, pD1, pD2, nshards, nq, ratio, th, oIdx, oD1, oD2, accepted);
}

void launch_distance(const uint8_t* a, const uint8_t* b, int* d, long long n, cudaStream_t s) {
  if (n > 0)
    distance_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint4*>(a),
                                                                 reinterpret_cast<const uint4*>(b), d, n);
}

}  // namespace orbx
