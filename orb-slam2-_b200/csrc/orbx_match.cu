// Hamming matching kernels (R/src/ORBmatcher.cpp DescriptorDistance
// :1901-1917 and the running best / second-best loops :266-300, :526-556).
//
// Brute-force top-2: bound by the integer pipe (8 XOR + 8 LOP3 + 4 POPC per pair after carry-save compression), not by HBM: a 32-byte
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

#ifndef ORBX_BF_THREADS
#define ORBX_BF_THREADS 256
#endif
constexpr int kBfThreads = ORBX_BF_THREADS;
#ifndef ORBX_BF_QPT
#define ORBX_BF_QPT 4
#endif
constexpr int kBfQPT = ORBX_BF_QPT;             // queries per thread
constexpr int kBfQueriesPerCta = kBfThreads * kBfQPT;
constexpr int kBfTileRows = 512;                // database rows staged per step (16 KB)
constexpr unsigned kKeyNone = 0xFFFFFFFFu;
#ifndef ORBX_BF_GROUP
#define ORBX_BF_GROUP 4
#endif
constexpr int kBfGroup = ORBX_BF_GROUP;         // rows per guarded group (even)
#ifndef ORBX_BF_GUARD_MIN_ROWS
#define ORBX_BF_GUARD_MIN_ROWS 8192
#endif
constexpr int kBfGuardMinRows = ORBX_BF_GUARD_MIN_ROWS;  // chunks shorter than this use the unguarded kernel

__device__ __forceinline__ int hamming256(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1) {
  return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
         __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}

// ---- brute-force inner step -------------------------------------------------------------------------
// POPC issues at 16 lanes/clk/SM on sm_100a (tools/pipe_probe.cu: 0.124 warp-instr/clk/SMSP), a quarter of
// the LOP3/VIMNMX rate, so 8 POPC per pair made the first version of this kernel POPC-bound (ncu r1_l: ALU
// pipe 45 %, 15.2 of 16 POPC lanes/clk). The 8 XOR words are first compressed with carry-save adders
// (a full adder over three words is two LOP3: parity 0x96 and majority 0xE8):
//   (x0,x1,x2)->(s1,c1)  (x3,x4,x5)->(s2,c2)  (s1,s2,x6)->(s3,c3)  (c1,c2,c3)->(s4,c4)
//   distance = popc(s3) + popc(x7) + 2 popc(s4) + 4 popc(c4)
// 4 POPC + 8 LOP3 + 8 XOR per pair. The weighted sum and the key packing (distance << 23 | row) are one
// IMAD each on the otherwise idle FMA pipe; the multipliers come in as kernel arguments so that ptxas
// keeps them as IMADs instead of strength-reducing them to ALU shifts.
struct KeyMul { unsigned m23, m24, m25; };

__device__ __forceinline__ unsigned lop3_xor(unsigned a, unsigned b, unsigned c) {
  unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0x96;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ unsigned lop3_maj(unsigned a, unsigned b, unsigned c) {
  unsigned r; asm("lop3.b32 %0, %1, %2, %3, 0xE8;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}
__device__ __forceinline__ unsigned imad(unsigned a, unsigned b, unsigned c) {
  unsigned r; asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c)); return r;
}

__device__ __forceinline__ unsigned hamming_key(const uint4& a0, const uint4& a1, const uint4& b0, const uint4& b1,
                                                unsigned rowKey, const KeyMul& km) {
  const unsigned x0 = a0.x ^ b0.x, x1 = a0.y ^ b0.y, x2 = a0.z ^ b0.z, x3 = a0.w ^ b0.w;
  const unsigned x4 = a1.x ^ b1.x, x5 = a1.y ^ b1.y, x6 = a1.z ^ b1.z, x7 = a1.w ^ b1.w;
  const unsigned s1 = lop3_xor(x0, x1, x2), c1 = lop3_maj(x0, x1, x2);
  const unsigned s2 = lop3_xor(x3, x4, x5), c2 = lop3_maj(x3, x4, x5);
  const unsigned s3 = lop3_xor(s1, s2, x6), c3 = lop3_maj(s1, s2, x6);
  const unsigned s4 = lop3_xor(c1, c2, c3), c4 = lop3_maj(c1, c2, c3);
  unsigned key = imad(__popc(s3), km.m23, rowKey);
  key = imad(__popc(x7), km.m23, key);
  key = imad(__popc(s4), km.m24, key);
  return imad(__popc(c4), km.m25, key);
}

// grid: (query groups, chunks). partial arrays are [chunk][nq]. GUARD: the grouped fast path below; it pays once a
// chunk's second-best keys have settled, i.e. for long chunks (launch_bf2 picks).
template <bool GUARD>
__global__ void __launch_bounds__(kBfThreads)
bf2_kernel(const uint4* __restrict__ q, int nq, const uint4* __restrict__ db, long long ndb, int rowsPerChunk,
           long long idxBase, int* __restrict__ pIdx, int* __restrict__ pD1, int* __restrict__ pD2, const KeyMul km) {
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
    int r = 0;
    // Groups of kBfGroup rows: the fast path only tracks the smallest key of the group per query (half a VIMNMX3
    // per pair) and compares it with the query's second-best key once per group. A key >= k2 (>= k1) changes
    // neither k1 nor k2, so a group without a key below k2 is skipped exactly; after n rows a group holds an
    // improving row with probability ~2*kBfGroup/n, so the exact update (5 min/max per row pair) runs on a few
    // per cent of the groups only (ALU-pipe operations per pair: 8 XOR + 8 LOP3 + 0.5 VIMNMX3 + 0.25 ISETP instead of
    // + 2.5 VIMNMX). Measured on config 5 (10k x 10M, 1 GPU): unguarded 801; exact update for all four queries of a thread
    // when any improves, groups of 16 / 8 / 4 rows: 777 / 818 / 841; exact update per improving query (this code),
    // groups of 16 / 8 / 4 / 2: 863 / 874 / 868 / 798 Gpairs/s, and 726 / 784 / 807 / 808 on one 1.25M-row shard of the
    // 8-GPU split (shorter chunks spend more of their rows settling) -- with the ALU pipe relieved the 4 POPC per pair (a quarter-rate pipe: 32 issue cycles per pair against
    // 33 on the ALU pipe) co-limit the loop.
    for (; GUARD && r + kBfGroup <= rows; r += kBfGroup, rowKey += kBfGroup) {
      unsigned m[kBfQPT];
#pragma unroll
      for (int k = 0; k < kBfQPT; k++) m[k] = kKeyNone;
#pragma unroll
      for (int i = 0; i < kBfGroup; i += 2) {
        const uint4 d0 = tile[2 * (r + i)], d1 = tile[2 * (r + i) + 1], e0 = tile[2 * (r + i) + 2], e1 = tile[2 * (r + i) + 3];
#pragma unroll
        for (int k = 0; k < kBfQPT; k++) {
          const unsigned ka = hamming_key(qa[k], qb[k], d0, d1, rowKey + i, km);
          const unsigned kb = hamming_key(qa[k], qb[k], e0, e1, rowKey + i + 1, km);
          m[k] = min(m[k], min(ka, kb));
        }
      }
      // exact update, only for the queries of this thread whose group minimum beats their second best
#pragma unroll
      for (int k = 0; k < kBfQPT; k++) {
        if (m[k] < k2[k]) {
#pragma unroll 1
          for (int i = 0; i < kBfGroup; i += 2) {
            const uint4 d0 = tile[2 * (r + i)], d1 = tile[2 * (r + i) + 1], e0 = tile[2 * (r + i) + 2], e1 = tile[2 * (r + i) + 3];
            const unsigned ka = hamming_key(qa[k], qb[k], d0, d1, rowKey + i, km);
            const unsigned kb = hamming_key(qa[k], qb[k], e0, e1, rowKey + i + 1, km);
            const unsigned lo = min(ka, kb), hi = max(ka, kb);
            k2[k] = min(min(k2[k], hi), max(k1[k], lo));
            k1[k] = min(k1[k], lo);
          }
        }
      }
    }
    // two rows per step: the (k1, k2) update of a sorted pair (lo, hi) is 5 min/max instead of 6
    for (; r + 1 < rows; r += 2, rowKey += 2) {
      const uint4 d0 = tile[2 * r], d1 = tile[2 * r + 1], e0 = tile[2 * r + 2], e1 = tile[2 * r + 3];
#pragma unroll
      for (int k = 0; k < kBfQPT; k++) {
        const unsigned ka = hamming_key(qa[k], qb[k], d0, d1, rowKey, km);
        const unsigned kb = hamming_key(qa[k], qb[k], e0, e1, rowKey + 1, km);
        const unsigned lo = min(ka, kb), hi = max(ka, kb);
        k2[k] = min(min(k2[k], hi), max(k1[k], lo));
        k1[k] = min(k1[k], lo);
      }
    }
    if (r < rows) {
      const uint4 d0 = tile[2 * r], d1 = tile[2 * r + 1];
#pragma unroll
      for (int k = 0; k < kBfQPT; k++) {
        const unsigned key = hamming_key(qa[k], qb[k], d0, d1, rowKey, km);
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
             long long shardStride, int nq, float ratio, int th, int* __restrict__ oIdx, int* __restrict__ oD1,
             int* __restrict__ oD2, uint8_t* __restrict__ accepted) {
  const int qi = blockIdx.x * blockDim.x + threadIdx.x;
  if (qi >= nq) return;
  int best = INT_MAX, best2 = INT_MAX, idx = -1;
  for (int s = 0; s < nshards; s++) {
    const long long o = (long long)s * shardStride + qi;
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
  // ~8 CTAs per SM overall (2.7 waves of the 3 resident ones). Every chunk restarts its top-2 from scratch and runs
  // the exact update until its second best has settled, so fewer chunks would be cheaper per pair, but exactly one
  // wave of equal CTAs measured slower (config 5 on one GPU: 829 vs 868 Gpairs/s; one of its 8 shards: 799 vs 806):
  // the dynamic hand-out of many CTAs balances the SMs better than one CTA set that finishes together.
  const int qGroups = (nq + kBfQueriesPerCta - 1) / kBfQueriesPerCta;
  long long chunks = (8LL * smCount + qGroups - 1) / qGroups;
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
  const KeyMul km = {1u << 23, 1u << 24, 1u << 25};
  // A chunk runs the exact update until its second-best keys have settled (a thousand rows or so per warp): short
  // chunks (small databases, or few query groups and hence many chunks) are faster without the guard -- 2048 x 1M
  // rows in 592 chunks of 1.7 k rows: 2.64 ms unguarded, 3.13 ms guarded; 10 k x 10 M in 119 chunks of 84 k: 125 / 115 ms.
  if (rowsPerChunk >= kBfGuardMinRows)
    bf2_kernel<true><<<grid, kBfThreads, 0, s>>>(reinterpret_cast<const uint4*>(q), nq, reinterpret_cast<const uint4*>(db), ndb,
                                                 rowsPerChunk, idxBase, pIdx, pD1, pD2, km);
  else
    bf2_kernel<false><<<grid, kBfThreads, 0, s>>>(reinterpret_cast<const uint4*>(q), nq, reinterpret_cast<const uint4*>(db), ndb,
                                                  rowsPerChunk, idxBase, pIdx, pD1, pD2, km);
}

void launch_merge(const int* pIdx, const int* pD1, const int* pD2, int nshards, int nq, float ratio, int th,
                  int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s) {
  merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(pIdx, pD1, pD2, nshards, nq, nq, ratio, th, oIdx, oD1, oD2, accepted);
}

// shard s of array X starts at X + s*shardStride (the all-gathered layout [rank][idx | d1 | d2][nq] is read in place)
void launch_merge_strided(const int* pIdx, const int* pD1, const int* pD2, int nshards, long long shardStride, int nq,
                          float ratio, int th, int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s) {
  merge_kernel<<<(nq + 255) / 256, 256, 0, s>>>(pIdx, pD1, pD2, nshards, shardStride, nq, ratio, th, oIdx, oD1, oD2,
                                                accepted);
}

void launch_distance(const uint8_t* a, const uint8_t* b, int* d, long long n, cudaStream_t s) {
  if (n > 0)
    distance_kernel<<<(unsigned)((n + 255) / 256), 256, 0, s>>>(reinterpret_cast<const uint4*>(a),
                                                                 reinterpret_cast<const uint4*>(b), d, n);
}

}  // namespace orbx

// ---- allocation-free brute-force top-2 (include/orbx.h: orbx_match_bf2_noalloc) -----------------------
// The per-chunk partial results of bf2_kernel live in a file-scope device array instead of the matcher's
// grow-only cudaMalloc'd buffer, for callers that may not allocate (CUDA graph capture, serving loops).
// In host code `g_bf2_scratch` is only the host-side stand-in of the symbol, never a device address: the
// address the kernels get comes from cudaGetSymbolAddress, looked up once per device with that device current.
// Every GPU that loads the library holds the array (12 MiB) until the process exits.
namespace orbx {

constexpr int kNoAllocMaxQueries = ORBX_NOALLOC_MAX_QUERIES;
constexpr long long kNoAllocTriples = 1LL << 20;  // chunks * nq <= this; 3 ints per triple = 12 MiB
static __device__ __align__(256) int g_bf2_scratch[3 * kNoAllocTriples];

namespace {
constexpr int kNoAllocMaxDevices = 64;
std::mutex g_noallocMutex;
int* g_noallocPtr[kNoAllocMaxDevices] = {};
int g_noallocSms[kNoAllocMaxDevices] = {};

// device must be current. No allocation: cudaGetSymbolAddress only resolves the module's symbol.
int noalloc_setup(int device, int** ptr, int* sms) {
  std::lock_guard<std::mutex> lock(g_noallocMutex);
  if (!g_noallocPtr[device]) {
    int* p = nullptr;
    ORBX_CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_bf2_scratch));
    int n = 0;
    ORBX_CUDA_TRY(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, device));
    g_noallocSms[device] = n;
    g_noallocPtr[device] = p;
  }
  *ptr = g_noallocPtr[device];
  *sms = g_noallocSms[device];
  return ORBX_OK;
}
}  // namespace
}  // namespace orbx

extern "C" int orbx_match_bf2_noalloc(int32_t device, const uint8_t* d_q, int32_t nq, const uint8_t* d_db, int64_t ndb,
                                      float ratio, int32_t th, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2,
                                      uint8_t* d_accepted, void* stream) {
  using namespace orbx;
  if (!d_q || !d_db || !d_idx || !d_d1 || !d_d2 || nq <= 0 || ndb <= 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (device < 0 || device >= kNoAllocMaxDevices) { set_error("device %d out of range", device); return ORBX_ERR_ARG; }
  if (nq > kNoAllocMaxQueries) {
    set_error("orbx_match_bf2_noalloc: %d queries exceed the static scratch limit of %d (use orbx_match_bf2_device)", nq,
              kNoAllocMaxQueries);
    return ORBX_ERR_ARG;
  }
  if (ndb > 0x7fffffffLL) { set_error("database index exceeds int32"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  int* scratch = nullptr;
  int sms = 0;
  int rc = noalloc_setup(device, &scratch, &sms);
  if (rc) return rc;
  // same chunk plan as orbx_match_bf2_device, then fewer / longer chunks when the partials would not fit
  int rowsPerChunk = 0;
  int chunks = bf2_plan_chunks(nq, ndb, sms, &rowsPerChunk);
  const long long maxChunks = kNoAllocTriples / nq;  // >= 64 for nq <= kNoAllocMaxQueries
  if (chunks > maxChunks) {
    long long rpc = (ndb + maxChunks - 1) / maxChunks;
    rpc = (rpc + kBfTileRows - 1) / kBfTileRows * kBfTileRows;
    if (rpc > (1LL << 23)) {
      set_error("orbx_match_bf2_noalloc: %lld rows x %d queries exceed the static scratch (%lld partial triples)",
                (long long)ndb, nq, kNoAllocTriples);
      return ORBX_ERR_ARG;
    }
    rowsPerChunk = (int)rpc;
    chunks = (int)((ndb + rpc - 1) / rpc);
  }
  const size_t per = (size_t)chunks * nq;
  int* pIdx = scratch;
  int* pD1 = pIdx + per;
  int* pD2 = pD1 + per;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  // bf2_kernel writes every (chunk, query) slot it is launched for and merge_kernel reads only those: nothing in
  // the shared array has to start at zero, stale contents of earlier calls are never read.
  launch_bf2(d_q, nq, d_db, ndb, rowsPerChunk, chunks, 0, pIdx, pD1, pD2, s);
  launch_merge(pIdx, pD1, pD2, chunks, nq, ratio, th, d_idx, d_d1, d_d2, d_accepted, s);
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}
