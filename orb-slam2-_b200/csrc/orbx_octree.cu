// DistributeOctTree on the GPU (R/src/ORBextractor.cpp:513-817).
//
// The reference walks a std::list of quadtree leaves serially. What it computes can be stated
// without the list (this is what the kernel implements; tests/test_gpu_extractor.py::test_octree_matches_oracle_in_order
// and tests/test_gpu_fuzz.py check it against the list-walking oracle, order included):
//
//  * The leaf list is always: [children created in the latest round, newest first] ++ [older leaves
//    in their previous order]  (children are push_front'ed in creation order, parents erased in
//    place). So the list is an array rebuilt once per round with two prefix sums.
//  * A phase-1 sweep (:637-717) splits EVERY leaf holding >1 point, visiting them in list order.
//  * A phase-2 round (:726-791) visits the leaves holding >1 point in order of (point count
//    descending, then list order) -- list order is creation order reversed, i.e. exactly the
//    canonical (size, creation sequence) key of the oracle -- and stops right after the split that
//    brings the leaf count to >= N.  With every candidate split evaluated speculatively in parallel,
//    the stopping index is the first position where  size + prefix_sum(nonEmptyChildren-1) >= N.
//  * DivideNode (:513-569) only needs per-child point COUNTS (4 shared-memory atomics per point);
//    the points themselves are never moved: each carries the index of its current leaf.
//  * The kept point of a leaf is its maximum response, first in cv::FAST emission order on ties
//    (:798-814). Emission order is (cell row, cell col, y, x); it is folded into a 64-bit key so one
//    atomicMax per point selects it regardless of the order candidates were appended in.
//
// One CTA per (frame, level); node tables live in shared memory, point->leaf indices in an L2-resident
// scratch array.
#include "orbx_common.cuh"

namespace orbx {

constexpr int kOctThreads = 512;       // launch bound; a lone frame runs its 8 CTAs this wide (latency)
// A batch runs narrow CTAs instead: the selection is a chain of short barrier-separated steps, so what pays is many
// resident CTAs per SM, not threads per CTA (512 frames on a B200: 0.35 ms at 512 threads, 0.21 at 256, 0.165 at
// 128, 0.23 at 64; one frame: 138 us per extraction at 512 threads, 155 at 128).
constexpr int kOctThreadsBatch = 128;
constexpr int kOctBatchFrames = 32;    // from this many frames per launch on
constexpr int kMaxRoots = 64;

struct OctTask {
  const uint32_t* cand;   // packed candidates of this (frame, level)
  uint16_t* node;         // scratch: leaf index per candidate
  uint32_t* kept;         // output, list order
  int* keptCount;
  int n;                  // candidates
  int width, height;      // maxBorder-minBorder extents (:576-580)
  int N;                  // mnFeaturesPerLevel[level]
  int cap;                // node/kept capacity
  int wCell, hCell, nCols;  // FAST cell grid, for the emission-order key
};

// exclusive scan of data[0..n) in place; returns the total. All threads of the CTA must call it.
__device__ int block_exclusive_scan(int* data, int n, int* warpSums) {
  const int tid = threadIdx.x, T = blockDim.x;
  const int chunk = (n + T - 1) / T;
  const int b = min(tid * chunk, n), e = min(b + chunk, n);
  int s = 0;
  for (int i = b; i < e; i++) s += data[i];
  // inclusive scan of s across the CTA
  const int lane = tid & 31, warp = tid >> 5;
  int v = s;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, v, o);
    if (lane >= o) v += t;
  }
  if (lane == 31) warpSums[warp] = v;
  __syncthreads();
  if (warp == 0) {
    int w = (lane < (T >> 5)) ? warpSums[lane] : 0;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, w, o);
      if (lane >= o) w += t;
    }
    warpSums[lane] = w;  // inclusive over warps
  }
  __syncthreads();
  const int total = warpSums[(T >> 5) - 1];
  int run = v - s + (warp ? warpSums[warp - 1] : 0);  // exclusive prefix of this thread's chunk
  __syncthreads();  // warpSums may be reused by the caller right after
  for (int i = b; i < e; i++) {
    const int t = data[i];
    data[i] = run;
    run += t;
  }
  __syncthreads();
  return total;
}

__device__ __forceinline__ int quadrant(uint32_t c, short4 box) {
  // box = (x0, y0, x1, y1); halfX = ceil((x1-x0)/2) (:515-516); tests pt.x < x0+halfX, pt.y < y0+halfY
  const int hx = (box.z - box.x + 1) >> 1, hy = (box.w - box.y + 1) >> 1;
  const int right = cand_x(c) >= box.x + hx, low = cand_y(c) >= box.y + hy;
  return right | (low << 1);  // 0:n1 1:n2 2:n3 3:n4
}

__device__ void octree_run(const OctTask& t, unsigned char* smemRaw) {
  const int tid = threadIdx.x, T = blockDim.x;
  const int cap = t.cap;
  // ---- shared-memory carve-up (all arrays `cap` long unless noted) ----
  short4* box[2];
  int* cnt[2];
  box[0] = reinterpret_cast<short4*>(smemRaw);
  box[1] = box[0] + cap;
  cnt[0] = reinterpret_cast<int*>(box[1] + cap);
  cnt[1] = cnt[0] + cap;
  int* childCnt = cnt[1] + cap;       // [cap][4]
  int* scanA = childCnt + 4 * cap;    // generic scan buffer
  int* scanB = scanA + cap;
  int* order = scanB + cap;           // visiting order -> node
  int* vis = order + cap;             // node -> visiting index (or -1)
  unsigned short* newPos = reinterpret_cast<unsigned short*>(vis + cap);
  unsigned short* childPos = newPos + cap;  // [cap][4]
  __shared__ int warpSums[32];
  __shared__ int sS, sPhase, sExpand, sStop, sDone;

  const int n = t.n;
  if (n <= 0) {
    if (tid == 0) *t.keptCount = 0;
    return;
  }
  // ---- roots (:577-627) ----
  const int nIni = max(1, min(kMaxRoots, (int)roundf((float)t.width / (float)t.height)));
  const float hX = __fdiv_rn((float)t.width, (float)nIni);
  for (int i = tid; i < nIni; i += T) scanA[i] = 0;
  __syncthreads();
  for (int i = tid; i < n; i += T) {
    const int r = min((int)__fdiv_rn((float)cand_x(t.cand[i]), hX), nIni - 1);
    atomicAdd(&scanA[r], 1);
  }
  __syncthreads();
  for (int i = tid; i < nIni; i += T) {
    scanB[i] = scanA[i] > 0;
    cnt[1][i] = scanA[i];  // stash counts
  }
  __syncthreads();
  int S = block_exclusive_scan(scanB, nIni, warpSums);  // scanB[r] = list position of root r
  for (int r = tid; r < nIni; r += T) {
    if (cnt[1][r] > 0) {
      const int pos = scanB[r];
      box[0][pos] = make_short4((short)(int)__fmul_rn(hX, (float)r), 0, (short)(int)__fmul_rn(hX, (float)(r + 1)),
                                (short)t.height);
      cnt[0][pos] = cnt[1][r];
    }
  }
  for (int i = tid; i < n; i += T) {
    const int r = min((int)__fdiv_rn((float)cand_x(t.cand[i]), hX), nIni - 1);
    t.node[i] = (uint16_t)scanB[r];
  }
  if (tid == 0) { sPhase = 1; sDone = 0; }
  __syncthreads();

  int cur = 0;
  const int N = t.N;
  while (true) {
    const int prevS = S;
    short4* B = box[cur];
    int* Cn = cnt[cur];
    short4* B2 = box[cur ^ 1];
    int* Cn2 = cnt[cur ^ 1];
    // A: per-child point counts of every expandable leaf
    for (int i = tid; i < 4 * S; i += T) childCnt[i] = 0;
    if (tid == 0) sExpand = 0;
    __syncthreads();
    for (int i = tid; i < n; i += T) {
      const int nd = t.node[i];
      if (Cn[nd] > 1) atomicAdd(&childCnt[4 * nd + quadrant(t.cand[i], B[nd])], 1);
    }
    __syncthreads();
    // B: expandable flags -> compact list of expandable leaves in list order
    for (int i = tid; i < S; i += T) scanA[i] = Cn[i] > 1;
    __syncthreads();
    const int E = block_exclusive_scan(scanA, S, warpSums);  // scanA[i] = index among expandables
    const int phase = sPhase;
    int nsplit = E;
    if (phase == 1) {
      for (int i = tid; i < S; i += T) {
        if (Cn[i] > 1) { vis[i] = scanA[i]; order[scanA[i]] = i; }
        else vis[i] = -1;
      }
      __syncthreads();
    } else {
      // phase 2: rank by (count desc, list position asc) -- O(E^2/T) compares, E <= N
      for (int i = tid; i < S; i += T) {
        if (Cn[i] > 1) scanB[scanA[i]] = i;  // scanB[k] = k-th expandable leaf
        vis[i] = -1;
      }
      __syncthreads();
      for (int k = tid; k < E; k += T) {
        const int me = scanB[k], c = Cn[me];
        int rank = 0;
        for (int m = 0; m < E; m++) {
          const int cm = Cn[scanB[m]];
          rank += (cm > c) | ((cm == c) & (m < k));
        }
        order[rank] = me;
        vis[me] = rank;
      }
      __syncthreads();
      // stopping point: first visiting index r with S + sum_{k<=r}(ne_k - 1) >= N  (:783)
      for (int k = tid; k < E; k += T) {
        const int* cc = &childCnt[4 * order[k]];
        scanA[k] = (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0) - 1;
      }
      if (tid == 0) sStop = E;
      __syncthreads();
      block_exclusive_scan(scanA, E, warpSums);
      for (int k = tid; k < E; k += T) {
        const int* cc = &childCnt[4 * order[k]];
        const int incl = scanA[k] + (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0) - 1;
        if (S + incl >= N) atomicMin(&sStop, k + 1);
      }
      __syncthreads();
      nsplit = sStop;
    }
    // D: child slots in creation order (visiting order, n1..n4) and positions of surviving leaves
    for (int k = tid; k < E; k += T) {
      const int* cc = &childCnt[4 * order[k]];
      scanA[k] = (k < nsplit) ? (cc[0] > 0) + (cc[1] > 0) + (cc[2] > 0) + (cc[3] > 0) : 0;
    }
    __syncthreads();
    const int C = block_exclusive_scan(scanA, E, warpSums);  // scanA[k] = first creation index of parent k
    for (int i = tid; i < S; i += T) scanB[i] = !(vis[i] >= 0 && vis[i] < nsplit);
    __syncthreads();
    const int U = block_exclusive_scan(scanB, S, warpSums);  // scanB[i] = rank among survivors
    if (C + U > cap) {  // cannot happen for keptCap >= max(N+3, 4*nIni); guard against corrupting smem
      if (tid == 0) *t.keptCount = -1;
      return;
    }
    // E: write the new list: [children, newest first] ++ [survivors in old order]
    for (int i = tid; i < S; i += T) {
      const int v = vis[i];
      if (v >= 0 && v < nsplit) {
        const short4 b = B[i];
        const int hx = (b.z - b.x + 1) >> 1, hy = (b.w - b.y + 1) >> 1;
        const int mx = b.x + hx, my = b.y + hy;
        int ci = scanA[v];
        int nExp = 0;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const int c = childCnt[4 * i + q];
          if (c > 0) {
            const int pos = C - 1 - ci;
            ci++;
            B2[pos] = make_short4((short)((q & 1) ? mx : b.x), (short)((q & 2) ? my : b.y),
                                  (short)((q & 1) ? b.z : mx), (short)((q & 2) ? b.w : my));
            Cn2[pos] = c;
            childPos[4 * i + q] = (unsigned short)pos;
            nExp += c > 1;
          }
        }
        if (nExp) atomicAdd(&sExpand, nExp);
      } else {
        const int pos = C + scanB[i];
        B2[pos] = B[i];
        Cn2[pos] = Cn[i];
        newPos[i] = (unsigned short)pos;
      }
    }
    __syncthreads();
    // F: re-label the points
    for (int i = tid; i < n; i += T) {
      const int nd = t.node[i];
      const int v = vis[nd];
      t.node[i] = (v >= 0 && v < nsplit) ? childPos[4 * nd + quadrant(t.cand[i], B[nd])] : newPos[nd];
    }
    S = C + U;
    cur ^= 1;
    // G: termination (:722-726, :789-790)
    if (tid == 0) {
      if (S >= N || S == prevS) sDone = 1;
      else if (sPhase == 1 && S + 3 * sExpand > N) sPhase = 2;
    }
    __syncthreads();
    if (sDone) break;
  }

  // ---- keep the best point of every leaf (:798-814), output in list order ----
  unsigned long long* best = reinterpret_cast<unsigned long long*>(childCnt);  // cap*16 bytes available
  for (int i = tid; i < S; i += T) best[i] = 0ull;
  __syncthreads();
  for (int i = tid; i < n; i += T) {
    const uint32_t c = t.cand[i];
    const int x = cand_x(c), y = cand_y(c);
    const unsigned long long cell = (unsigned long long)(((y - 3) / t.hCell) * t.nCols + (x - 3) / t.wCell);
    const unsigned long long ord = (cell << 24) | ((unsigned long long)y << 12) | (unsigned long long)x;
    const unsigned long long key = ((unsigned long long)cand_score(c) << 40) | ((1ull << 40) - 1 - ord);
    atomicMax(&best[t.node[i]], key);
  }
  __syncthreads();
  for (int i = tid; i < S; i += T) {
    const unsigned long long key = best[i];
    const unsigned long long ord = (1ull << 40) - 1 - (key & ((1ull << 40) - 1));
    t.kept[i] = pack_cand((int)(ord & 0xfff), (int)((ord >> 12) & 0xfff), (int)(key >> 40));
  }
  if (tid == 0) *t.keptCount = S;
}

size_t octree_smem_bytes(int cap) {
  // box 2x8 + cnt 2x4 + childCnt 16 + scanA/scanB/order/vis 16 + newPos 2 + childPos 8
  return (size_t)cap * (16 + 8 + 16 + 16 + 2 + 8) + 64;
}

__global__ void __launch_bounds__(kOctThreads)
octree_kernel(const FrameSetGeom* __restrict__ g, const uint32_t* __restrict__ cand,
              const int* __restrict__ candCount, uint16_t* __restrict__ nodeScratch,
              uint32_t* __restrict__ kept, int* __restrict__ keptCount, int cap, int levelBase) {
  pdl_wait();
  extern __shared__ __align__(16) unsigned char smemRaw[];
  const int level = levelBase + blockIdx.x, f = blockIdx.y, nl = g->nlevels;
  const LevelGeom& L = g->lv[level];
  OctTask t;
  const long long coff = L.candOff + (long long)f * L.candCap;
  t.cand = cand + coff;
  t.node = nodeScratch + coff;
  t.kept = kept + L.keptOff + (long long)f * L.keptCap;
  t.keptCount = &keptCount[f * nl + level];
  t.n = min(candCount[f * nl + level], L.candCap);
  t.width = L.maxBX - kMinBorder;
  t.height = L.maxBY - kMinBorder;
  t.N = L.nfeat;
  t.cap = min(cap, L.keptCap);
  t.wCell = L.wCell; t.hCell = L.hCell; t.nCols = L.nCols;
  octree_run(t, smemRaw);
}

__global__ void __launch_bounds__(kOctThreads)
octree_single_kernel(OctTask t) {
  extern __shared__ __align__(16) unsigned char smemRaw[];
  octree_run(t, smemRaw);
}

int launch_octree(const FrameSetGeom* dGeom, const FrameSetGeom& hGeom, const uint32_t* cand,
                  const int* candCount, uint16_t* nodeScratch, uint32_t* kept, int* keptCount,
                  int nframes, cudaStream_t s, int levelBegin, int levelEnd) {
  if (levelEnd < 0) levelEnd = hGeom.nlevels;
  int cap = 0;
  for (int l = 0; l < hGeom.nlevels; l++) cap = max(cap, hGeom.lv[l].keptCap);
  const size_t smem = octree_smem_bytes(cap);
  static SmemLimit limit;
  cudaError_t e = raise_dynamic_smem(reinterpret_cast<const void*>(octree_kernel), smem, &limit);
  if (e != cudaSuccess) return (int)e;
  dim3 grid(levelEnd - levelBegin, nframes);
  launch_pdl(octree_kernel, grid, dim3(nframes >= kOctBatchFrames ? kOctThreadsBatch : kOctThreads), smem, s, dGeom, cand, candCount, nodeScratch, kept, keptCount, cap, levelBegin);
  return 0;
}

int launch_octree_single(const uint32_t* cand, int n, int width, int height, int nfeat, int wCell, int hCell,
                         int nCols, uint16_t* nodeScratch, uint32_t* kept, int* keptCount, int keptCap,
                         cudaStream_t s) {
  OctTask t;
  t.cand = cand; t.node = nodeScratch; t.kept = kept; t.keptCount = keptCount;
  t.n = n; t.width = width; t.height = height; t.N = nfeat; t.cap = keptCap;
  t.wCell = wCell; t.hCell = hCell; t.nCols = nCols;
  const size_t smem = octree_smem_bytes(keptCap);
  static SmemLimit limit;
  cudaError_t e = raise_dynamic_smem(reinterpret_cast<const void*>(octree_single_kernel), smem, &limit);
  if (e != cudaSuccess) return (int)e;
  octree_single_kernel<<<1, kOctThreads, smem, s>>>(t);
  return 0;
}

}  // namespace orbx
