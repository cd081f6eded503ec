// C ABI of liborbx.so, matcher half (include/orbx.h): DescriptorDistance and brute-force top-2.
#include <algorithm>
#include <new>

#include "orbx_host.h"

namespace orbx {
int bf2_plan_chunks(int nq, long long ndb, int smCount, int* rowsPerChunk);
void launch_bf2(const uint8_t* q, int nq, const uint8_t* db, long long ndb, int rowsPerChunk, int chunks,
                long long idxBase, int* pIdx, int* pD1, int* pD2, cudaStream_t s);
void launch_merge(const int* pIdx, const int* pD1, const int* pD2, int nshards, int nq, float ratio, int th,
                  int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s);
void launch_distance(const uint8_t* a, const uint8_t* b, int* d, long long n, cudaStream_t s);
}  // namespace orbx

using namespace orbx;

struct orbx_matcher {
  int device = 0;
  int smCount = 148;
  cudaStream_t stream = nullptr;
  long long launches = 0;
  DeviceBuf partial;           // [3][chunks][nq] ints
  DeviceBuf dQ, dDb, dOut;     // staging for the host-buffer entry points
};

namespace {
int set_dev(orbx_matcher* h) {
  if (!h) { set_error("null matcher handle"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->device));
  return ORBX_OK;
}
}  // namespace

extern "C" {

int orbx_matcher_create(int32_t device, orbx_matcher** out) {
  if (!out) { set_error("null argument"); return ORBX_ERR_ARG; }
  *out = nullptr;
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("no usable CUDA device (%s); liborbx has no CPU fallback", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    return ORBX_ERR_CUDA;
  }
  if (device < 0 || device >= n) { set_error("device %d out of range (have %d)", device, n); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  orbx_matcher* h = new (std::nothrow) orbx_matcher();
  if (!h) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  h->device = device;
  cudaDeviceGetAttribute(&h->smCount, cudaDevAttrMultiProcessorCount, device);
  e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  if (e != cudaSuccess) { set_error("stream creation failed: %s", cudaGetErrorString(e)); delete h; return ORBX_ERR_CUDA; }
  *out = h;
  return ORBX_OK;
}

void orbx_matcher_destroy(orbx_matcher* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->stream) { cudaStreamSynchronize(h->stream); cudaStreamDestroy(h->stream); }
  h->partial.release(); h->dQ.release(); h->dDb.release(); h->dOut.release();
  delete h;
}

int64_t orbx_matcher_launch_count(const orbx_matcher* h) { return h ? h->launches : 0; }

int orbx_descriptor_distance(orbx_matcher* h, const uint8_t* a, const uint8_t* b, int32_t* dist, int64_t n) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!a || !b || !dist || n < 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  if ((rc = h->dQ.ensure((size_t)n * 32)) || (rc = h->dDb.ensure((size_t)n * 32)) || (rc = h->dOut.ensure((size_t)n * 4))) return rc;
  ORBX_CUDA_TRY(cudaMemcpyAsync(h->dQ.p, a, (size_t)n * 32, cudaMemcpyHostToDevice, h->stream));
  ORBX_CUDA_TRY(cudaMemcpyAsync(h->dDb.p, b, (size_t)n * 32, cudaMemcpyHostToDevice, h->stream));
  launch_distance(static_cast<const uint8_t*>(h->dQ.p), static_cast<const uint8_t*>(h->dDb.p), static_cast<int*>(h->dOut.p), n, h->stream);
  h->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(dist, h->dOut.p, (size_t)n * 4, cudaMemcpyDeviceToHost, h->stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));
  return ORBX_OK;
}

int orbx_match_bf2_device(orbx_matcher* h, const uint8_t* d_q, int32_t nq, const uint8_t* d_db, int64_t ndb,
                          int64_t idx_base, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2, void* stream) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!d_q || !d_db || !d_idx || !d_d1 || !d_d2 || nq <= 0 || ndb <= 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (idx_base + ndb > 0x7fffffffLL) { set_error("global index exceeds int32"); return ORBX_ERR_ARG; }
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : h->stream;
  int rowsPerChunk = 0;
  const int chunks = bf2_plan_chunks(nq, ndb, h->smCount, &rowsPerChunk);
  const size_t per = (size_t)chunks * nq;
  if ((rc = h->partial.ensure(per * 3 * sizeof(int)))) return rc;
  int* pIdx = static_cast<int*>(h->partial.p);
  int* pD1 = pIdx + per;
  int* pD2 = pD1 + per;
  launch_bf2(d_q, nq, d_db, ndb, rowsPerChunk, chunks, idx_base, pIdx, pD1, pD2, s);
  launch_merge(pIdx, pD1, pD2, chunks, nq, 0.f, 0, d_idx, d_d1, d_d2, nullptr, s);
  h->launches += 2;
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

int orbx_match_merge_device(orbx_matcher* h, const int32_t* d_idx, const int32_t* d_d1, const int32_t* d_d2,
                            int32_t nshards, int32_t nq, float ratio, int32_t th, int32_t* d_out_idx,
                            int32_t* d_out_d1, int32_t* d_out_d2, uint8_t* d_accepted, void* stream) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!d_idx || !d_d1 || !d_d2 || !d_out_idx || !d_out_d1 || !d_out_d2 || nshards <= 0 || nq <= 0) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : h->stream;
  launch_merge(d_idx, d_d1, d_d2, nshards, nq, ratio, th, d_out_idx, d_out_d1, d_out_d2, d_accepted, s);
  h->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

int orbx_match_bf2(orbx_matcher* h, const uint8_t* q, int32_t nq, const uint8_t* db, int64_t ndb, float ratio,
                   int32_t th, int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!q || !db || !idx || !d1 || !d2 || nq <= 0 || ndb <= 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if ((rc = h->dQ.ensure((size_t)nq * 32)) || (rc = h->dDb.ensure((size_t)ndb * 32)) ||
      (rc = h->dOut.ensure((size_t)nq * 32))) return rc;
  cudaStream_t s = h->stream;
  ORBX_CUDA_TRY(cudaMemcpyAsync(h->dQ.p, q, (size_t)nq * 32, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(h->dDb.p, db, (size_t)ndb * 32, cudaMemcpyHostToDevice, s));
  int* oIdx = static_cast<int*>(h->dOut.p);
  int* oD1 = oIdx + nq;
  int* oD2 = oD1 + nq;
  int* tIdx = oD2 + nq;  // second triple: the ratio-tested merge of a single "shard"
  int* tD1 = tIdx + nq;
  int* tD2 = tD1 + nq;
  uint8_t* dAcc = reinterpret_cast<uint8_t*>(tD2 + nq);
  rc = orbx_match_bf2_device(h, static_cast<const uint8_t*>(h->dQ.p), nq, static_cast<const uint8_t*>(h->dDb.p), ndb, 0,
                             oIdx, oD1, oD2, s);
  if (rc) return rc;
  launch_merge(oIdx, oD1, oD2, 1, nq, ratio, th, tIdx, tD1, tD2, dAcc, s);
  h->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(idx, tIdx, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(d1, tD1, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(d2, tD2, (size_t)nq * 4, cudaMemcpyDeviceToHost, s));
  if (accepted) ORBX_CUDA_TRY(cudaMemcpyAsync(accepted, dAcc, (size_t)nq, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

}  // extern "C"
