// C ABI of liborbx.so, multi-GPU half (include/orbx.h: orbx_shard_*). Host code only.
//
// the path shards two ways on one 8 x B200 box.
//  * Extraction: frames are independent units -> contiguous, pair-aligned blocks of a sequence per GPU
//    (a stereo pair's L and R frame stay on one GPU, R/src/Frame.cpp:86-101), NO collective.
//  * Brute-force Hamming (config 5): the descriptor database is cut into contiguous row ranges, queries are
//    replicated, every GPU produces (idx, d1, d2) per query, ONE ncclAllGather of 12 B per query and rank, then the
//    merge kernel applies the sequential rule of R/src/ORBmatcher.cpp:282-291 and the ratio test -- on every rank,
//    reading the gathered [rank][idx|d1|d2][nq] block in place.
//
// A group is either all the ranks of one process (orbx_shard_group_create: one host thread per device inside the
// calls below) or one rank of a multi-process job (orbx_shard_group_create_rank with an id from
// orbx_shard_unique_id broadcast by the launcher, e.g. one process per GPU under torchrun).
//
// NCCL is bound at run time (dlopen of libnccl.so.2, preferring a copy the process has already loaded): the library
// keeps loading on a box without NCCL, and a single-rank group never touches it.
#include <dlfcn.h>
#include <nccl.h>

#include <algorithm>
#include <cstring>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "orbx_host.h"

namespace orbx {
void launch_merge_strided(const int* pIdx, const int* pD1, const int* pD2, int nshards, long long shardStride, int nq,
                          float ratio, int th, int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s);
}
using namespace orbx;

namespace {

struct NcclApi {
  void* lib = nullptr;
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;
};

std::mutex g_ncclMutex;
NcclApi g_nccl;

int nccl_api(NcclApi** out) {
  std::lock_guard<std::mutex> lock(g_ncclMutex);
  if (!g_nccl.lib) {
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD);  // e.g. the copy torch has loaded
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { set_error("multi-GPU groups need NCCL: %s", dlerror()); return ORBX_ERR_CUDA; }
    NcclApi a;
    a.lib = lib;
    bool ok = true;
    auto sym = [&](const char* name) { void* p = dlsym(lib, name); ok = ok && p; return p; };
    a.GetUniqueId = reinterpret_cast<decltype(a.GetUniqueId)>(sym("ncclGetUniqueId"));
    a.CommInitRank = reinterpret_cast<decltype(a.CommInitRank)>(sym("ncclCommInitRank"));
    a.CommDestroy = reinterpret_cast<decltype(a.CommDestroy)>(sym("ncclCommDestroy"));
    a.AllGather = reinterpret_cast<decltype(a.AllGather)>(sym("ncclAllGather"));
    a.GroupStart = reinterpret_cast<decltype(a.GroupStart)>(sym("ncclGroupStart"));
    a.GroupEnd = reinterpret_cast<decltype(a.GroupEnd)>(sym("ncclGroupEnd"));
    a.GetErrorString = reinterpret_cast<decltype(a.GetErrorString)>(sym("ncclGetErrorString"));
    a.GetVersion = reinterpret_cast<decltype(a.GetVersion)>(sym("ncclGetVersion"));
    if (!ok) { set_error("libnccl.so.2 lacks a required symbol"); return ORBX_ERR_CUDA; }
    g_nccl = a;
  }
  *out = &g_nccl;
  return ORBX_OK;
}

#define ORBX_NCCL_TRY(api, expr)                                                                      \
  do {                                                                                                \
    ncclResult_t _r = (expr);                                                                         \
    if (_r != ncclSuccess) {                                                                          \
      ::orbx::set_error("%s failed: %s (%s:%d)", #expr, (api)->GetErrorString(_r), __FILE__, __LINE__); \
      return ORBX_ERR_CUDA;                                                                           \
    }                                                                                                 \
  } while (0)

static_assert(sizeof(ncclUniqueId) == ORBX_SHARD_ID_BYTES, "orbx.h: ORBX_SHARD_ID_BYTES");

struct Rank {
  int device = 0, rank = 0;
  ncclComm_t comm = nullptr;
  orbx_matcher* matcher = nullptr;
  cudaStream_t stream = nullptr;
  // database shard: rows [rowLo, rowHi) of the global database
  const uint8_t* dDb = nullptr;
  long long rowLo = 0, rowHi = 0;
  DeviceBuf ownedDb, dQ, part, gathered, out;
  orbx_extractor* ex = nullptr;
  orbx_params exParams{};
  long long launches = 0;
  std::string err;   // worker threads report through here (set_error is thread-local)
  int rc = ORBX_OK;
};

}  // namespace

struct orbx_shard_group {
  int nranks = 1;
  long long ndb = 0;
  std::vector<Rank> local;
};

namespace {

void block_range(long long total, int rank, int nranks, int group, long long* lo, long long* hi) {
  // contiguous blocks of whole groups (group = 2 keeps a stereo pair on one rank); the first `rem` ranks get one more
  const long long units = (total + group - 1) / group;
  const long long base = units / nranks, rem = units % nranks;
  const long long u0 = rank * base + std::min<long long>(rank, rem);
  const long long u1 = u0 + base + (rank < rem ? 1 : 0);
  *lo = std::min(u0 * group, total);
  *hi = std::min(u1 * group, total);
}

int init_rank(Rank& r) {
  ORBX_CUDA_TRY(cudaSetDevice(r.device));
  int rc = orbx_matcher_create(r.device, &r.matcher);
  if (rc) return rc;
  ORBX_CUDA_TRY(cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking));
  return ORBX_OK;
}

// Run fn(rank) for every local rank: inline for one rank, one std::thread per device otherwise (every device needs
// its own host thread to keep its copy queue fed). The first error wins; its message is re-raised on this thread.
template <class F>
int for_each_rank(orbx_shard_group* g, F fn) {
  if (g->local.size() == 1) return fn(g->local[0]);
  std::vector<std::thread> th;
  for (Rank& r : g->local)
    th.emplace_back([&r, &fn] {
      r.rc = fn(r);
      r.err = r.rc ? last_error() : "";
    });
  for (auto& t : th) t.join();
  for (Rank& r : g->local)
    if (r.rc) { set_error("rank %d (device %d): %s", r.rank, r.device, r.err.c_str()); return r.rc; }
  return ORBX_OK;
}

// bf2 on the rank's shard, all-gather of the per-rank triple block, merge + ratio test; all on stream s.
// With several local ranks the caller brackets the calls with ncclGroupStart / End (one thread issues them all).
int match_enqueue_bf2(orbx_shard_group* g, Rank& r, const uint8_t* dQ, int nq, cudaStream_t s) {
  ORBX_CUDA_TRY(cudaSetDevice(r.device));
  if (!r.dDb || r.rowHi <= r.rowLo) { set_error("rank %d holds no database shard (orbx_shard_set_database*)", r.rank); return ORBX_ERR_STATE; }
  int rc;
  if ((rc = r.part.ensure((size_t)3 * nq * sizeof(int)))) return rc;
  if ((rc = r.gathered.ensure((size_t)g->nranks * 3 * nq * sizeof(int)))) return rc;
  int* p = static_cast<int*>(r.part.p);
  rc = orbx_match_bf2_device(r.matcher, dQ, nq, r.dDb, r.rowHi - r.rowLo, r.rowLo, p, p + nq, p + 2 * (size_t)nq, s);
  r.launches += 2;
  return rc;
}

int match_enqueue_gather(orbx_shard_group* g, Rank& r, int nq, cudaStream_t s) {
  if (g->nranks == 1) return ORBX_OK;
  NcclApi* api;
  int rc = nccl_api(&api);
  if (rc) return rc;
  ORBX_CUDA_TRY(cudaSetDevice(r.device));
  ORBX_NCCL_TRY(api, api->AllGather(r.part.p, r.gathered.p, (size_t)3 * nq, ncclInt32, r.comm, s));
  return ORBX_OK;
}

int match_enqueue_merge(orbx_shard_group* g, Rank& r, int nq, float ratio, int th, int* dIdx, int* dD1, int* dD2,
                        uint8_t* dAcc, cudaStream_t s) {
  ORBX_CUDA_TRY(cudaSetDevice(r.device));
  const int* src = static_cast<const int*>(g->nranks == 1 ? r.part.p : r.gathered.p);
  launch_merge_strided(src, src + nq, src + 2 * (size_t)nq, g->nranks, 3LL * nq, nq, ratio, th, dIdx, dD1, dD2, dAcc, s);
  r.launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_shard_unique_id(uint8_t* id) {
  if (!id) { set_error("null argument"); return ORBX_ERR_ARG; }
  NcclApi* api;
  int rc = nccl_api(&api);
  if (rc) return rc;
  ncclUniqueId u;
  ORBX_NCCL_TRY(api, api->GetUniqueId(&u));
  std::memcpy(id, &u, sizeof(u));
  return ORBX_OK;
}

void orbx_shard_group_destroy(orbx_shard_group* g) {
  if (!g) return;
  for (Rank& r : g->local) {
    cudaSetDevice(r.device);
    if (r.stream) cudaStreamSynchronize(r.stream);
    if (r.comm && g_nccl.CommDestroy) g_nccl.CommDestroy(r.comm);
    if (r.ex) orbx_extractor_destroy(r.ex);
    if (r.matcher) orbx_matcher_destroy(r.matcher);
    r.ownedDb.release(); r.dQ.release(); r.part.release(); r.gathered.release(); r.out.release();
    if (r.stream) cudaStreamDestroy(r.stream);
  }
  delete g;
}

int orbx_shard_group_create(const int32_t* devices, int32_t ndev, orbx_shard_group** out) {
  if (!devices || ndev <= 0 || !out) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  *out = nullptr;
  orbx_shard_group* g = new (std::nothrow) orbx_shard_group();
  if (!g) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  g->nranks = ndev;
  g->local.resize(ndev);
  int rc = ORBX_OK;
  for (int i = 0; i < ndev && !rc; i++) {
    g->local[i].device = devices[i];
    g->local[i].rank = i;
    rc = init_rank(g->local[i]);
  }
  if (!rc && ndev > 1) {
    NcclApi* api = nullptr;
    rc = nccl_api(&api);
    ncclUniqueId id;
    if (!rc && api->GetUniqueId(&id) != ncclSuccess) { set_error("ncclGetUniqueId failed"); rc = ORBX_ERR_CUDA; }
    if (!rc) {
      // ncclCommInitAll's recipe: one thread, all ranks inside a group
      ncclResult_t r0 = api->GroupStart();
      for (int i = 0; i < ndev && r0 == ncclSuccess; i++) {
        cudaSetDevice(devices[i]);
        r0 = api->CommInitRank(&g->local[i].comm, ndev, id, i);
      }
      ncclResult_t r1 = api->GroupEnd();
      if (r0 != ncclSuccess || r1 != ncclSuccess) {
        set_error("NCCL communicator over %d devices failed: %s", ndev, api->GetErrorString(r0 != ncclSuccess ? r0 : r1));
        rc = ORBX_ERR_CUDA;
      }
    }
  }
  if (rc) { std::string keep = last_error(); orbx_shard_group_destroy(g); set_error("%s", keep.c_str()); return rc; }
  *out = g;
  return ORBX_OK;
}

int orbx_shard_group_create_rank(int32_t device, int32_t rank, int32_t nranks, const uint8_t* id, orbx_shard_group** out) {
  if (!out || nranks <= 0 || rank < 0 || rank >= nranks || (nranks > 1 && !id)) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  *out = nullptr;
  orbx_shard_group* g = new (std::nothrow) orbx_shard_group();
  if (!g) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  g->nranks = nranks;
  g->local.resize(1);
  g->local[0].device = device;
  g->local[0].rank = rank;
  int rc = init_rank(g->local[0]);
  if (!rc && nranks > 1) {
    NcclApi* api = nullptr;
    rc = nccl_api(&api);
    if (!rc) {
      ncclUniqueId u;
      std::memcpy(&u, id, sizeof(u));
      cudaSetDevice(device);
      ncclResult_t r = api->CommInitRank(&g->local[0].comm, nranks, u, rank);
      if (r != ncclSuccess) { set_error("ncclCommInitRank(rank %d of %d) failed: %s", rank, nranks, api->GetErrorString(r)); rc = ORBX_ERR_CUDA; }
    }
  }
  if (rc) { std::string keep = last_error(); orbx_shard_group_destroy(g); set_error("%s", keep.c_str()); return rc; }
  *out = g;
  return ORBX_OK;
}

int orbx_shard_group_info(const orbx_shard_group* g, int32_t* nranks, int32_t* nlocal, int32_t* first_rank, int32_t* nccl_version) {
  if (!g) { set_error("null group"); return ORBX_ERR_ARG; }
  if (nranks) *nranks = g->nranks;
  if (nlocal) *nlocal = (int)g->local.size();
  if (first_rank) *first_rank = g->local[0].rank;
  if (nccl_version) {
    *nccl_version = 0;
    if (g_nccl.GetVersion) { int v = 0; if (g_nccl.GetVersion(&v) == ncclSuccess) *nccl_version = v; }
  }
  return ORBX_OK;
}

int64_t orbx_shard_launch_count(const orbx_shard_group* g) {
  long long t = 0;
  if (g) for (const Rank& r : g->local) t += r.launches;
  return t;
}

int orbx_shard_block_range(int64_t total, int32_t rank, int32_t nranks, int32_t group, int64_t* lo, int64_t* hi) {
  if (total < 0 || nranks <= 0 || rank < 0 || rank >= nranks || group <= 0 || !lo || !hi) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  long long a, b;
  block_range(total, rank, nranks, group, &a, &b);
  *lo = a; *hi = b;
  return ORBX_OK;
}

int orbx_shard_set_database(orbx_shard_group* g, const uint8_t* db, int64_t ndb) {
  if (!g || !db || ndb <= 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (ndb > 0x7fffffffLL) { set_error("global index exceeds int32"); return ORBX_ERR_ARG; }
  g->ndb = ndb;
  const int nranks = g->nranks;
  return for_each_rank(g, [&](Rank& r) -> int {
    ORBX_CUDA_TRY(cudaSetDevice(r.device));
    block_range(ndb, r.rank, nranks, 1, &r.rowLo, &r.rowHi);
    const size_t bytes = (size_t)(r.rowHi - r.rowLo) * 32;
    int rc = r.ownedDb.ensure(std::max<size_t>(bytes, 32));
    if (rc) return rc;
    ORBX_CUDA_TRY(cudaMemcpyAsync(r.ownedDb.p, db + (size_t)r.rowLo * 32, bytes, cudaMemcpyHostToDevice, r.stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(r.stream));
    r.dDb = static_cast<const uint8_t*>(r.ownedDb.p);
    return ORBX_OK;
  });
}

int orbx_shard_set_database_device(orbx_shard_group* g, int32_t local_index, const uint8_t* d_rows, int64_t row_lo,
                                   int64_t row_hi, int64_t ndb) {
  if (!g || local_index < 0 || local_index >= (int)g->local.size() || !d_rows || row_lo < 0 || row_hi <= row_lo || row_hi > ndb) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  if (ndb > 0x7fffffffLL) { set_error("global index exceeds int32"); return ORBX_ERR_ARG; }
  Rank& r = g->local[local_index];
  r.dDb = d_rows; r.rowLo = row_lo; r.rowHi = row_hi;
  g->ndb = ndb;
  return ORBX_OK;
}

int orbx_shard_match_bf2_device(orbx_shard_group* g, int32_t local_index, const uint8_t* d_q, int32_t nq, float ratio,
                                int32_t th, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2, uint8_t* d_accepted, void* stream) {
  if (!g || local_index < 0 || local_index >= (int)g->local.size() || !d_q || nq <= 0 || !d_idx || !d_d1 || !d_d2) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  if (g->local.size() > 1 && g->nranks > 1) {
    set_error("orbx_shard_match_bf2_device serves one-rank-per-process groups; a group that owns several devices matches with orbx_shard_match_bf2");
    return ORBX_ERR_STATE;
  }
  Rank& r = g->local[local_index];
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : r.stream;
  int rc;
  if ((rc = match_enqueue_bf2(g, r, d_q, nq, s))) return rc;
  if ((rc = match_enqueue_gather(g, r, nq, s))) return rc;
  return match_enqueue_merge(g, r, nq, ratio, th, d_idx, d_d1, d_d2, d_accepted, s);
}

int orbx_shard_match_bf2(orbx_shard_group* g, const uint8_t* q, int32_t nq, float ratio, int32_t th, int32_t* idx,
                         int32_t* d1, int32_t* d2, uint8_t* accepted) {
  if (!g || !q || nq <= 0 || !idx || !d1 || !d2) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  int rc;
  // queries to every local rank, per-rank top-2
  for (Rank& r : g->local) {
    ORBX_CUDA_TRY(cudaSetDevice(r.device));
    if ((rc = r.dQ.ensure((size_t)nq * 32)) || (rc = r.out.ensure((size_t)nq * 13 + 64))) return rc;
    ORBX_CUDA_TRY(cudaMemcpyAsync(r.dQ.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice, r.stream));
    if ((rc = match_enqueue_bf2(g, r, static_cast<const uint8_t*>(r.dQ.p), nq, r.stream))) return rc;
  }
  if (g->nranks > 1) {
    NcclApi* api;
    if ((rc = nccl_api(&api))) return rc;
    // one host thread drives several communicators: the collective calls of all local ranks form one group
    ORBX_NCCL_TRY(api, api->GroupStart());
    for (Rank& r : g->local)
      if ((rc = match_enqueue_gather(g, r, nq, r.stream))) { api->GroupEnd(); return rc; }
    ORBX_NCCL_TRY(api, api->GroupEnd());
  }
  for (Rank& r : g->local) {
    int* o = static_cast<int*>(r.out.p);
    uint8_t* acc = reinterpret_cast<uint8_t*>(o + 3 * (size_t)nq);
    if ((rc = match_enqueue_merge(g, r, nq, ratio, th, o, o + nq, o + 2 * (size_t)nq, acc, r.stream))) return rc;
  }
  // every rank holds the full answer; the caller's arrays are filled from the first local rank
  Rank& r0 = g->local[0];
  ORBX_CUDA_TRY(cudaSetDevice(r0.device));
  const int* o = static_cast<const int*>(r0.out.p);
  ORBX_CUDA_TRY(cudaMemcpyAsync(idx, o, (size_t)nq * 4, cudaMemcpyDeviceToHost, r0.stream));
  ORBX_CUDA_TRY(cudaMemcpyAsync(d1, o + nq, (size_t)nq * 4, cudaMemcpyDeviceToHost, r0.stream));
  ORBX_CUDA_TRY(cudaMemcpyAsync(d2, o + 2 * (size_t)nq, (size_t)nq * 4, cudaMemcpyDeviceToHost, r0.stream));
  if (accepted) ORBX_CUDA_TRY(cudaMemcpyAsync(accepted, o + 3 * (size_t)nq, (size_t)nq, cudaMemcpyDeviceToHost, r0.stream));
  for (Rank& r : g->local) {
    ORBX_CUDA_TRY(cudaSetDevice(r.device));
    ORBX_CUDA_TRY(cudaStreamSynchronize(r.stream));
  }
  return ORBX_OK;
}

int orbx_shard_extract_sequence(orbx_shard_group* g, const orbx_params* params, const uint8_t* const* imgs, int64_t count,
                                int32_t group, int32_t width, int32_t height, size_t stride, orbx_keypoint* kps,
                                uint8_t* desc, int32_t cap, int32_t* n, float mbf, float mb, float* u_right, float* depth,
                                int32_t* n_matched) {
  if (!g || !params || !imgs || count <= 0 || (group != 1 && group != 2) || !kps || !desc || !n || cap <= 0) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  const bool stereo = u_right || depth || n_matched;
  if (stereo && (group != 2 || !u_right || !depth || !n_matched || (count & 1))) {
    set_error("stereo sequences are L R L R ...: group 2, an even frame count and all three result arrays");
    return ORBX_ERR_ARG;
  }
  const int nranks = g->nranks;
  return for_each_rank(g, [&](Rank& r) -> int {
    long long lo, hi;
    block_range(count, r.rank, nranks, group, &lo, &hi);
    if (hi <= lo) return ORBX_OK;
    ORBX_CUDA_TRY(cudaSetDevice(r.device));
    orbx_params p = *params;
    p.device = r.device;
    if (!r.ex || std::memcmp(&p, &r.exParams, sizeof(p)) != 0) {
      if (r.ex) { orbx_extractor_destroy(r.ex); r.ex = nullptr; }
      int rc = orbx_extractor_create(&p, &r.ex);
      if (rc) return rc;
      r.exParams = p;
    }
    // steps of max_batch frames (even for pairs), two submissions in flight
    long long step = std::max(group, p.max_batch - (p.max_batch % group));
    int pending[2] = {0, 0};
    int k = 0, rc = ORBX_OK;
    const long long launches0 = orbx_launch_count(r.ex);
    for (long long f0 = lo; f0 < hi && !rc; f0 += step, k++) {
      const int cnt = (int)std::min(step, hi - f0);
      if (pending[k & 1]) { rc = orbx_extract_batch_wait(r.ex, pending[k & 1]); pending[k & 1] = 0; if (rc) break; }
      if (stereo)
        rc = orbx_stereo_batch_submit(r.ex, imgs + f0, cnt, width, height, stride, kps + (size_t)f0 * cap, desc + (size_t)f0 * cap * 32,
                                      cap, n + f0, mbf, mb, u_right + (size_t)(f0 / 2) * cap, depth + (size_t)(f0 / 2) * cap,
                                      n_matched + f0 / 2, &pending[k & 1]);
      else
        rc = orbx_extract_batch_submit(r.ex, imgs + f0, cnt, width, height, stride, kps + (size_t)f0 * cap,
                                       desc + (size_t)f0 * cap * 32, cap, n + f0, &pending[k & 1]);
    }
    // wait in submission order; keep the first error but always drain both tickets
    for (int j = 0; j < 2; j++) {
      const int t = pending[(k + j) & 1];
      if (t) { const int w = orbx_extract_batch_wait(r.ex, t); if (!rc) rc = w; }
    }
    r.launches += orbx_launch_count(r.ex) - launches0;
    return rc;
  });
}

}  // extern "C"
