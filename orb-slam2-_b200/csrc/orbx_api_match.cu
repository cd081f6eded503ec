// C ABI of liborbx.so, matcher half (include/orbx.h): DescriptorDistance and brute-force top-2.
#include <algorithm>
#include <cstring>
#include <initializer_list>
#include <new>
#include <vector>

#include "orbx_host.h"
#include "orbx_stereo.cuh"

namespace orbx {
int bf2_plan_chunks(int nq, long long ndb, int smCount, int* rowsPerChunk);
void launch_bf2(const uint8_t* q, int nq, const uint8_t* db, long long ndb, int rowsPerChunk, int chunks,
                long long idxBase, int* pIdx, int* pD1, int* pD2, cudaStream_t s);
void launch_merge(const int* pIdx, const int* pD1, const int* pD2, int nshards, int nq, float ratio, int th,
                  int* oIdx, int* oD1, int* oD2, uint8_t* accepted, cudaStream_t s);
void launch_distance(const uint8_t* a, const uint8_t* b, int* d, long long n, cudaStream_t s);
void launch_grid_build(const orbx_keypoint* kps, int n, float minX, float maxX, float minY, float maxY, int* cellStart,
                       int* items, cudaStream_t s);
void launch_window(bool fill, const orbx_keypoint* kps, const uint8_t* desc, float minX, float maxX, float minY, float maxY,
                   const int* cellStart, const int* items, const float* qxy, const float* qr, const int* qMinLevel,
                   const int* qMaxLevel, const uint8_t* qdesc, int nq, int* count, const int* off, int* candIdx,
                   int* candDist, cudaStream_t s);
void launch_scan(const int* count, int n, int* off, cudaStream_t s);
int grid_cells();
void launch_block_distance(const uint8_t* descA, const uint8_t* descB, const int* aOff, const int* aIdx, const int* bOff,
                           const int* bIdx, const int* entryBlock, const long long* dOff, int nEntries,
                           unsigned short* dist, cudaStream_t s);
void launch_distinctive(const uint8_t* desc, const int* obsOff, int nPoints, int* best, int* bestMedian, cudaStream_t s);
void launch_voc_descend(const uint8_t* feat, int n, const uint8_t* nodeDesc, const int* childOff, const int* childIdx,
                        int nidLevel, int* leafNode, int* levelNode, cudaStream_t s);
void launch_undistort_keys(const orbx_keypoint* in, orbx_keypoint* out, int n, const float* K4, const float* dist, int ndist,
                           cudaStream_t s);
void launch_undistort_points(const float* in, float* out, long long n, const float* K4, const float* dist, int ndist,
                             cudaStream_t s);
}  // namespace orbx

using namespace orbx;

struct orbx_matcher {
  int device = 0;
  int smCount = 148;
  cudaStream_t stream = nullptr;
  long long launches = 0;
  DeviceBuf partial;           // [3][chunks][nq] ints
  DeviceBuf dQ, dDb, dOut;     // staging for the host-buffer entry points
  DeviceBuf dWin;              // windowed search / stereo: one arena carved per call
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
  h->partial.release(); h->dQ.release(); h->dDb.release(); h->dOut.release(); h->dWin.release();
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

namespace {
// bump allocation inside h->dWin (256-byte aligned pieces)
struct Carver {
  uint8_t* base; size_t off = 0;
  template <class T> T* take(size_t count) {
    T* p = reinterpret_cast<T*>(base + off);
    off += (count * sizeof(T) + 255) & ~size_t(255);
    return p;
  }
};
size_t carve_size(std::initializer_list<size_t> bytes) {
  size_t t = 0;
  for (size_t b : bytes) t += (b + 255) & ~size_t(255);
  return t + 256;
}
}  // namespace

extern "C" {

int orbx_match_window(orbx_matcher* h, const orbx_keypoint* kps_un, const uint8_t* desc, int32_t n, float min_x,
                      float max_x, float min_y, float max_y, const float* q_xy, const float* q_r,
                      const int32_t* q_min_level, const int32_t* q_max_level, const uint8_t* q_desc, int32_t nq,
                      int32_t* cand_off, int32_t* cand_idx, int32_t* cand_dist, int32_t cand_cap, int32_t* n_cand) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!kps_un || !desc || !q_xy || !q_r || !q_min_level || !q_max_level || !q_desc || !cand_off || !n_cand || n < 0 ||
      nq < 0 || cand_cap < 0 || !(max_x > min_x) || !(max_y > min_y)) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  *n_cand = 0;
  if (nq == 0) { cand_off[0] = 0; return ORBX_OK; }
  const int cells = grid_cells();
  const size_t N = (size_t)std::max(n, 1), Q = (size_t)nq;
  if ((rc = h->dWin.ensure(carve_size({N * sizeof(orbx_keypoint), N * 32, (size_t)(cells + 1) * 4, N * 4, Q * 8, Q * 4, Q * 4,
                                       Q * 4, Q * 32, Q * 4, (Q + 1) * 4, (size_t)cand_cap * 4, (size_t)cand_cap * 4}))))
    return rc;
  Carver c{static_cast<uint8_t*>(h->dWin.p)};
  orbx_keypoint* dK = c.take<orbx_keypoint>(N);
  uint8_t* dD = c.take<uint8_t>(N * 32);
  int* dCell = c.take<int>(cells + 1);
  int* dItems = c.take<int>(N);
  float* dQxy = c.take<float>(Q * 2);
  float* dQr = c.take<float>(Q);
  int* dQmin = c.take<int>(Q);
  int* dQmax = c.take<int>(Q);
  uint8_t* dQd = c.take<uint8_t>(Q * 32);
  int* dCount = c.take<int>(Q);
  int* dOff = c.take<int>(Q + 1);
  int* dCi = c.take<int>(std::max(cand_cap, 1));
  int* dCd = c.take<int>(std::max(cand_cap, 1));
  cudaStream_t s = h->stream;
  if (n > 0) {
    ORBX_CUDA_TRY(cudaMemcpyAsync(dK, kps_un, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(dD, desc, (size_t)n * 32, cudaMemcpyHostToDevice, s));
  }
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQxy, q_xy, Q * 8, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQr, q_r, Q * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQmin, q_min_level, Q * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQmax, q_max_level, Q * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQd, q_desc, Q * 32, cudaMemcpyHostToDevice, s));
  launch_grid_build(dK, n, min_x, max_x, min_y, max_y, dCell, dItems, s);
  launch_window(false, dK, dD, min_x, max_x, min_y, max_y, dCell, dItems, dQxy, dQr, dQmin, dQmax, dQd, nq, dCount, dOff,
                dCi, dCd, s);
  launch_scan(dCount, nq, dOff, s);
  h->launches += 3;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(cand_off, dOff, (Q + 1) * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  const int total = cand_off[nq];
  *n_cand = total;
  if (total > cand_cap || (total > 0 && (!cand_idx || !cand_dist))) {
    set_error("candidate capacity %d < %d", cand_cap, total);
    return ORBX_ERR_CAPACITY;
  }
  if (total > 0) {
    launch_window(true, dK, dD, min_x, max_x, min_y, max_y, dCell, dItems, dQxy, dQr, dQmin, dQmax, dQd, nq, dCount, dOff,
                  dCi, dCd, s);
    h->launches++;
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(cand_idx, dCi, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(cand_dist, dCd, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  }
  return ORBX_OK;
}

// pyramids of the frames the two handles processed last (single-chunk batches, lane 0)
static int stereo_levels_of(const orbx_extractor* left, const orbx_extractor* right, StereoLevels* lv) {
  const orbx::Lane& LL = left->lane[0];
  const orbx::Lane& LR = right->lane[0];
  const int nlv = LL.geom.nlevels;
  if (LR.geom.nlevels != nlv || LL.w != LR.w || LL.h != LR.h) { set_error("left/right shapes differ"); return ORBX_ERR_ARG; }
  for (int l = 0; l < kMaxLevels; l++) {
    lv->left[l] = lv->right[l] = nullptr;
    lv->strideL[l] = lv->strideR[l] = 0;
    lv->pitchL[l] = lv->pitchR[l] = lv->width[l] = 0;
    lv->scale[l] = lv->invScale[l] = 1.f;
  }
  for (int l = 0; l < nlv; l++) {
    const LevelGeom& G = LL.geom.lv[l];
    const LevelGeom& GR = LR.geom.lv[l];
    if (l == 0) {
      lv->left[0] = left->lastLevel0; lv->strideL[0] = left->lastL0FrameStride; lv->pitchL[0] = left->lastL0Pitch;
      lv->right[0] = right->lastLevel0; lv->strideR[0] = right->lastL0FrameStride; lv->pitchR[0] = right->lastL0Pitch;
    } else {
      lv->left[l] = static_cast<const uint8_t*>(LL.pyrArena.p) + G.planeOff; lv->strideL[l] = G.planeBytes; lv->pitchL[l] = G.pitch;
      lv->right[l] = static_cast<const uint8_t*>(LR.pyrArena.p) + GR.planeOff; lv->strideR[l] = GR.planeBytes; lv->pitchR[l] = GR.pitch;
    }
    lv->width[l] = G.w;
    lv->scale[l] = left->tables.scale[l];
    lv->invScale[l] = left->tables.invScale[l];
  }
  lv->rows0 = LL.geom.lv[0].h;
  return ORBX_OK;
}

int orbx_stereo_match(orbx_matcher* h, orbx_extractor* left, orbx_extractor* right, int32_t frame_left,
                      int32_t frame_right, const orbx_keypoint* kps_l, const uint8_t* desc_l, int32_t nl,
                      const orbx_keypoint* kps_r, const uint8_t* desc_r, int32_t nr, float mbf, float mb,
                      float* u_right, float* depth, int32_t* n_matched) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!left || !right || !u_right || !depth || !n_matched || nl < 0 || nr < 0 || (nl > 0 && (!kps_l || !desc_l)) ||
      (nr > 0 && (!kps_r || !desc_r))) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  if (nr > 65535) { set_error("more than 65535 right keypoints"); return ORBX_ERR_ARG; }
  if (left->params.device != h->device || right->params.device != h->device) {
    set_error("extractors and matcher live on different devices");
    return ORBX_ERR_ARG;
  }
  if (!left->haveResult || !right->haveResult || frame_left < 0 || frame_left >= left->lastCount || frame_right < 0 ||
      frame_right >= right->lastCount) {
    set_error("the extractors hold no (single-chunk) result for frames %d / %d", frame_left, frame_right);
    return ORBX_ERR_STATE;
  }
  StereoLevels lv;
  if ((rc = stereo_levels_of(left, right, &lv))) return rc;
  const int nlv = left->lane[0].geom.nlevels;
  const size_t NL = (size_t)std::max(nl, 1), NR = (size_t)std::max(nr, 1);
  const size_t itemCap = NR * (size_t)stereo_items_per_keypoint(left->tables.scale[nlv - 1]);
  if ((rc = h->dWin.ensure(carve_size({NL * sizeof(orbx_keypoint), NL * 32, NR * sizeof(orbx_keypoint), NR * 32, NL * 4, NL * 4,
                                       NL * 4, 4, ((size_t)lv.rows0 + 1) * 4, itemCap * 4}))))
    return rc;
  Carver c{static_cast<uint8_t*>(h->dWin.p)};
  orbx_keypoint* dKL = c.take<orbx_keypoint>(NL);
  uint8_t* dDL = c.take<uint8_t>(NL * 32);
  orbx_keypoint* dKR = c.take<orbx_keypoint>(NR);
  uint8_t* dDR = c.take<uint8_t>(NR * 32);
  StereoArrays a;
  a.kpL = dKL; a.descL = reinterpret_cast<const uint4*>(dDL); a.nL = nullptr;
  a.kpR = dKR; a.descR = reinterpret_cast<const uint4*>(dDR); a.nR = nullptr;
  a.uRight = c.take<float>(NL);
  a.depth = c.take<float>(NL);
  a.sad = c.take<int>(NL);
  a.nKept = c.take<int>(1);
  a.rowStart = c.take<int>((size_t)lv.rows0 + 1);
  a.rowItems = c.take<int>(itemCap);
  a.itemCap = (int)itemCap;
  lv.pyrLeft0 = frame_left; lv.pyrRight0 = frame_right; lv.pyrStep = 0;
  lv.arrLeft0 = lv.arrRight0 = 0; lv.arrStep = 0; lv.cap = 0;  // one pair: the staged arrays start at row 0
  lv.fixedNL = nl; lv.fixedNR = nr;
  // the extractors' streams produced the pyramids; their host-buffer calls are synchronous, the device-resident
  // entry point is not: wait for both before reading
  ORBX_CUDA_TRY(cudaStreamSynchronize(left->lane[0].stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(right->lane[0].stream));
  cudaStream_t s = h->stream;
  if (nl > 0) {
    ORBX_CUDA_TRY(cudaMemcpyAsync(dKL, kps_l, (size_t)nl * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(dDL, desc_l, (size_t)nl * 32, cudaMemcpyHostToDevice, s));
  }
  if (nr > 0) {
    ORBX_CUDA_TRY(cudaMemcpyAsync(dKR, kps_r, (size_t)nr * sizeof(orbx_keypoint), cudaMemcpyHostToDevice, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(dDR, desc_r, (size_t)nr * 32, cudaMemcpyHostToDevice, s));
  }
  const int le = launch_stereo(lv, a, 1, nl, mbf, mb, s);
  if (le) { set_error("stereo kernel configuration failed: %s", cudaGetErrorString((cudaError_t)le)); return ORBX_ERR_CUDA; }
  h->launches += nl > 0 ? 3 : 2;
  ORBX_CUDA_TRY(cudaGetLastError());
  if (nl > 0) {
    ORBX_CUDA_TRY(cudaMemcpyAsync(u_right, a.uRight, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(depth, a.depth, (size_t)nl * 4, cudaMemcpyDeviceToHost, s));
  }
  ORBX_CUDA_TRY(cudaMemcpyAsync(n_matched, a.nKept, 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

int orbx_stereo_match_batch_device(orbx_matcher* h, orbx_extractor* ex, int32_t pairs, int32_t left_frame0,
                                   int32_t right_frame0, int32_t frame_step, const orbx_keypoint* d_kps,
                                   const uint8_t* d_desc, const int32_t* d_n, int32_t cap, float mbf, float mb,
                                   float* d_u_right, float* d_depth, int32_t* d_n_matched, void* stream) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!ex || pairs <= 0 || !d_kps || !d_desc || !d_n || cap <= 0 || cap > 65535 || !d_u_right || !d_depth || !d_n_matched ||
      left_frame0 < 0 || right_frame0 < 0) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  if (ex->params.device != h->device) { set_error("extractor and matcher live on different devices"); return ORBX_ERR_ARG; }
  const int lastL = left_frame0 + (pairs - 1) * frame_step, lastR = right_frame0 + (pairs - 1) * frame_step;
  if (!ex->haveResult || lastL < 0 || lastR < 0 || std::max({left_frame0, right_frame0, lastL, lastR}) >= ex->lastCount) {
    set_error("the extractor holds no (single-chunk) result covering frames up to %d", std::max(lastL, lastR));
    return ORBX_ERR_STATE;
  }
  StereoLevels lv;
  if ((rc = stereo_levels_of(ex, ex, &lv))) return rc;
  const int nlv = ex->lane[0].geom.nlevels;
  const size_t P = (size_t)pairs;
  const size_t itemCap = (size_t)cap * (size_t)stereo_items_per_keypoint(ex->tables.scale[nlv - 1]);
  if ((rc = h->dWin.ensure(carve_size({P * cap * 4, P * ((size_t)lv.rows0 + 1) * 4, P * itemCap * 4})))) return rc;
  Carver c{static_cast<uint8_t*>(h->dWin.p)};
  StereoArrays a;
  a.kpL = a.kpR = d_kps;
  a.descL = a.descR = reinterpret_cast<const uint4*>(d_desc);
  a.nL = a.nR = d_n;
  a.uRight = d_u_right; a.depth = d_depth; a.nKept = d_n_matched;
  a.sad = c.take<int>(P * cap);
  a.rowStart = c.take<int>(P * ((size_t)lv.rows0 + 1));
  a.rowItems = c.take<int>(P * itemCap);
  a.itemCap = (int)itemCap;
  lv.pyrLeft0 = lv.arrLeft0 = left_frame0;
  lv.pyrRight0 = lv.arrRight0 = right_frame0;
  lv.pyrStep = lv.arrStep = frame_step;
  lv.cap = cap;
  lv.fixedNL = lv.fixedNR = 0;
  // stream-ordered after the extraction when both run on the caller's stream; with the handles' own streams the
  // extraction is waited for here
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : h->stream;
  if (!stream) ORBX_CUDA_TRY(cudaStreamSynchronize(ex->lane[0].stream));
  const int le = launch_stereo(lv, a, pairs, cap, mbf, mb, s);
  if (le) { set_error("stereo kernel configuration failed: %s", cudaGetErrorString((cudaError_t)le)); return ORBX_ERR_CUDA; }
  h->launches += 3;
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

int orbx_match_blocks(orbx_matcher* h, const uint8_t* desc_a, int32_t na, const uint8_t* desc_b, int32_t nb,
                      const int32_t* a_off, const int32_t* a_idx, const int32_t* b_off, const int32_t* b_idx,
                      int32_t nblocks, int64_t* dist_off, uint16_t* dist, int64_t dist_cap, int64_t* n_dist) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (!dist_off || !n_dist || nblocks < 0 || na < 0 || nb < 0 || (nblocks > 0 && (!a_off || !b_off))) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  // host pass: matrix offsets, entry -> block table, index validation
  long long total = 0;
  dist_off[0] = 0;
  for (int k = 0; k < nblocks; k++) {
    const long long ca = (long long)a_off[k + 1] - a_off[k], cb = (long long)b_off[k + 1] - b_off[k];
    if (ca < 0 || cb < 0 || (k == 0 && (a_off[0] != 0 || b_off[0] != 0))) { set_error("offsets of block %d are not a CSR", k); return ORBX_ERR_ARG; }
    total += ca * cb;
    dist_off[k + 1] = total;
  }
  *n_dist = total;
  const int TA = nblocks ? a_off[nblocks] : 0, TB = nblocks ? b_off[nblocks] : 0;
  if ((TA > 0 && (!a_idx || !desc_a)) || (TB > 0 && (!b_idx || !desc_b))) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  for (int i = 0; i < TA; i++) if (a_idx[i] < 0 || a_idx[i] >= na) { set_error("a_idx[%d] = %d out of range", i, a_idx[i]); return ORBX_ERR_ARG; }
  for (int i = 0; i < TB; i++) if (b_idx[i] < 0 || b_idx[i] >= nb) { set_error("b_idx[%d] = %d out of range", i, b_idx[i]); return ORBX_ERR_ARG; }
  if (total > dist_cap || (total > 0 && !dist)) { set_error("distance capacity %lld < %lld", (long long)dist_cap, total); return ORBX_ERR_CAPACITY; }
  if (total == 0) return ORBX_OK;
  std::vector<int> entryBlock((size_t)TA);
  for (int k = 0; k < nblocks; k++) for (int e = a_off[k]; e < a_off[k + 1]; e++) entryBlock[e] = k;
  const size_t NA = (size_t)na, NB = (size_t)nb, K = (size_t)nblocks;
  if ((rc = h->dWin.ensure(carve_size({NA * 32, NB * 32, (K + 1) * 4, (size_t)TA * 4, (K + 1) * 4, (size_t)TB * 4, (size_t)TA * 4,
                                       (K + 1) * 8, (size_t)total * 2}))))
    return rc;
  Carver c{static_cast<uint8_t*>(h->dWin.p)};
  uint8_t* dA = c.take<uint8_t>(NA * 32);
  uint8_t* dB = c.take<uint8_t>(NB * 32);
  int* dAoff = c.take<int>(K + 1);
  int* dAidx = c.take<int>(TA);
  int* dBoff = c.take<int>(K + 1);
  int* dBidx = c.take<int>(TB);
  int* dEntry = c.take<int>(TA);
  long long* dDoff = c.take<long long>(K + 1);
  unsigned short* dDist = c.take<unsigned short>((size_t)total);
  cudaStream_t s = h->stream;
  ORBX_CUDA_TRY(cudaMemcpyAsync(dA, desc_a, NA * 32, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dB, desc_b, NB * 32, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dAoff, a_off, (K + 1) * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dAidx, a_idx, (size_t)TA * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dBoff, b_off, (K + 1) * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dBidx, b_idx, (size_t)TB * 4, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dEntry, entryBlock.data(), (size_t)TA * 4, cudaMemcpyHostToDevice, s));
  static_assert(sizeof(long long) == sizeof(int64_t), "offset width");
  ORBX_CUDA_TRY(cudaMemcpyAsync(dDoff, dist_off, (K + 1) * 8, cudaMemcpyHostToDevice, s));
  // entryBlock is pageable: the copy above has been staged by the time cudaMemcpyAsync returns
  launch_block_distance(dA, dB, dAoff, dAidx, dBoff, dBidx, dEntry, dDoff, TA, dDist, s);
  h->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(dist, dDist, (size_t)total * 2, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

int orbx_distinctive_device(orbx_matcher* h, const uint8_t* d_desc, const int32_t* d_obs_off, int32_t npoints,
                            int32_t* d_best, int32_t* d_best_median, void* stream) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (npoints < 0 || (npoints > 0 && (!d_desc || !d_obs_off || !d_best || !d_best_median))) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (npoints == 0) return ORBX_OK;
  launch_distinctive(d_desc, d_obs_off, npoints, d_best, d_best_median, stream ? static_cast<cudaStream_t>(stream) : h->stream);
  h->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

int orbx_distinctive(orbx_matcher* h, const uint8_t* desc, const int32_t* obs_off, int32_t npoints, int32_t* best,
                     int32_t* best_median) {
  int rc = set_dev(h);
  if (rc) return rc;
  if (npoints < 0 || (npoints > 0 && (!obs_off || !best || !best_median))) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (npoints == 0) return ORBX_OK;
  if (obs_off[0] != 0) { set_error("obs_off[0] must be 0"); return ORBX_ERR_ARG; }
  for (int p = 0; p < npoints; p++) {
    const long long n = (long long)obs_off[p + 1] - obs_off[p];
    if (n < 0 || n >= (1 << 20)) { set_error("point %d has %lld observations (supported: 0 .. 2^20-1)", p, n); return ORBX_ERR_ARG; }
  }
  const size_t R = (size_t)obs_off[npoints], P = (size_t)npoints;
  if (R > 0 && !desc) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if ((rc = h->dWin.ensure(carve_size({std::max<size_t>(R, 1) * 32, (P + 1) * 4, P * 4, P * 4})))) return rc;
  Carver c{static_cast<uint8_t*>(h->dWin.p)};
  uint8_t* dD = c.take<uint8_t>(std::max<size_t>(R, 1) * 32);
  int* dOff = c.take<int>(P + 1);
  int* dBest = c.take<int>(P);
  int* dMed = c.take<int>(P);
  cudaStream_t s = h->stream;
  if (R > 0) ORBX_CUDA_TRY(cudaMemcpyAsync(dD, desc, R * 32, cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(dOff, obs_off, (P + 1) * 4, cudaMemcpyHostToDevice, s));
  if ((rc = orbx_distinctive_device(h, dD, dOff, npoints, dBest, dMed, s))) return rc;
  ORBX_CUDA_TRY(cudaMemcpyAsync(best, dBest, P * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(best_median, dMed, P * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

// ---- device-resident frames -----------------------------------------------------------------
}  // extern "C"

struct orbx_frame {
  orbx_matcher* m = nullptr;
  int n = 0;
  bool ready = false;
  float bounds[4] = {0, 0, 0, 0};
  DeviceBuf kps, kpsUn, desc, grid, work;   // grid: cellStart | items; work: packed queries | counts | offsets | candidates
  void* hostQ = nullptr;                    // pinned staging of one call's packed queries
  size_t hostQBytes = 0;
};

extern "C" {

int orbx_undistort_points(int32_t device, const float* xy, int64_t n, const float* K4, const float* dist, int32_t n_dist,
                          float* out_xy) {
  if (!xy || !out_xy || !K4 || n < 0 || n_dist < 0 || n_dist > 12 || (n_dist > 0 && !dist)) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  int cnt = 0;
  cudaError_t e = cudaGetDeviceCount(&cnt);
  if (e != cudaSuccess || cnt <= 0) { set_error("no usable CUDA device; liborbx has no CPU fallback"); return ORBX_ERR_CUDA; }
  if (device < 0 || device >= cnt) { set_error("device %d out of range", device); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  if (n == 0) return ORBX_OK;
  float *dIn = nullptr, *dOut = nullptr;
  ORBX_CUDA_TRY(cudaMalloc(&dIn, (size_t)n * 8));
  if (cudaMalloc(&dOut, (size_t)n * 8) != cudaSuccess) { cudaFree(dIn); set_error("out of device memory"); return ORBX_ERR_CUDA; }
  cudaMemcpy(dIn, xy, (size_t)n * 8, cudaMemcpyHostToDevice);
  launch_undistort_points(dIn, dOut, n, K4, dist, n_dist, nullptr);
  cudaMemcpy(out_xy, dOut, (size_t)n * 8, cudaMemcpyDeviceToHost);
  e = cudaGetLastError();
  cudaFree(dIn); cudaFree(dOut);
  if (e != cudaSuccess) { set_error("undistort failed: %s", cudaGetErrorString(e)); return ORBX_ERR_CUDA; }
  return ORBX_OK;
}

int orbx_frame_create(orbx_matcher* m, orbx_frame** out) {
  int rc = set_dev(m);
  if (rc) return rc;
  if (!out) { set_error("null argument"); return ORBX_ERR_ARG; }
  orbx_frame* f = new (std::nothrow) orbx_frame();
  if (!f) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  f->m = m;
  *out = f;
  return ORBX_OK;
}

void orbx_frame_destroy(orbx_frame* f) {
  if (!f) return;
  cudaSetDevice(f->m->device);
  cudaStreamSynchronize(f->m->stream);
  f->kps.release(); f->kpsUn.release(); f->desc.release(); f->grid.release(); f->work.release();
  if (f->hostQ) cudaFreeHost(f->hostQ);
  delete f;
}

int orbx_frame_set(orbx_frame* f, const orbx_keypoint* kps, const uint8_t* desc, int32_t n, int32_t on_device,
                   const float* K4, const float* dist, int32_t n_dist, const float* bounds4) {
  if (!f) { set_error("null frame"); return ORBX_ERR_ARG; }
  int rc = set_dev(f->m);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!kps || !desc)) || !bounds4 || n_dist < 0 || n_dist > 12 || (n_dist > 0 && (!dist || !K4))) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  if (!(bounds4[1] > bounds4[0]) || !(bounds4[3] > bounds4[2])) { set_error("empty image bounds"); return ORBX_ERR_ARG; }
  f->ready = false;
  const size_t N = (size_t)std::max(n, 1);
  const int cells = grid_cells();
  if ((rc = f->kps.ensure(N * sizeof(orbx_keypoint))) || (rc = f->kpsUn.ensure(N * sizeof(orbx_keypoint))) ||
      (rc = f->desc.ensure(N * 32)) || (rc = f->grid.ensure(carve_size({(size_t)(cells + 1) * 4, N * 4}))))
    return rc;
  cudaStream_t s = f->m->stream;
  const cudaMemcpyKind kind = on_device ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
  orbx_keypoint* dK = static_cast<orbx_keypoint*>(f->kps.p);
  orbx_keypoint* dKU = static_cast<orbx_keypoint*>(f->kpsUn.p);
  if (n > 0) {
    ORBX_CUDA_TRY(cudaMemcpyAsync(dK, kps, (size_t)n * sizeof(orbx_keypoint), kind, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(f->desc.p, desc, (size_t)n * 32, kind, s));
    // R/src/Frame.cpp:472-476: no distortion => mvKeysUn = mvKeys
    if (n_dist > 0 && dist[0] != 0.0f) {
      launch_undistort_keys(dK, dKU, n, K4, dist, n_dist, s);
      f->m->launches++;
    } else {
      ORBX_CUDA_TRY(cudaMemcpyAsync(dKU, dK, (size_t)n * sizeof(orbx_keypoint), cudaMemcpyDeviceToDevice, s));
    }
  }
  Carver c{static_cast<uint8_t*>(f->grid.p)};
  int* dCell = c.take<int>(cells + 1);
  int* dItems = c.take<int>(N);
  launch_grid_build(dKU, n, bounds4[0], bounds4[1], bounds4[2], bounds4[3], dCell, dItems, s);   // AssignFeaturesToGrid :244-260
  f->m->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  f->n = n;
  for (int i = 0; i < 4; i++) f->bounds[i] = bounds4[i];
  f->ready = true;
  return ORBX_OK;
}

int orbx_frame_get_keys_un(orbx_frame* f, orbx_keypoint* kps_un) {
  if (!f || !f->ready) { set_error("frame holds no keypoints"); return ORBX_ERR_STATE; }
  int rc = set_dev(f->m);
  if (rc) return rc;
  if (f->n > 0) {
    if (!kps_un) { set_error("null argument"); return ORBX_ERR_ARG; }
    ORBX_CUDA_TRY(cudaMemcpyAsync(kps_un, f->kpsUn.p, (size_t)f->n * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, f->m->stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(f->m->stream));
  }
  return ORBX_OK;
}

int orbx_frame_match_window(orbx_frame* f, const float* q_xy, const float* q_r, const int32_t* q_min_level,
                            const int32_t* q_max_level, const uint8_t* q_desc, int32_t nq, int32_t* cand_off,
                            int32_t* cand_idx, int32_t* cand_dist, int32_t cand_cap, int32_t* n_cand) {
  if (!f || !f->ready) { set_error("frame holds no keypoints"); return ORBX_ERR_STATE; }
  orbx_matcher* h = f->m;
  int rc = set_dev(h);
  if (rc) return rc;
  if (nq < 0 || cand_cap < 0 || !cand_off || !n_cand || (nq > 0 && (!q_xy || !q_r || !q_min_level || !q_max_level || !q_desc))) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  *n_cand = 0;
  cand_off[0] = 0;
  if (nq == 0) return ORBX_OK;
  // one packed upload instead of five: xy | r | min | max | desc, each part 256-byte aligned
  const size_t Q = (size_t)nq;
  const size_t oXY = 0, oR = oXY + ((Q * 8 + 255) & ~size_t(255)), oMin = oR + ((Q * 4 + 255) & ~size_t(255)),
               oMax = oMin + ((Q * 4 + 255) & ~size_t(255)), oD = oMax + ((Q * 4 + 255) & ~size_t(255)),
               qBytes = oD + ((Q * 32 + 255) & ~size_t(255));
  if (qBytes > f->hostQBytes) {
    if (f->hostQ) { ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream)); cudaFreeHost(f->hostQ); f->hostQ = nullptr; f->hostQBytes = 0; }
    ORBX_CUDA_TRY(cudaMallocHost(&f->hostQ, qBytes * 2));
    f->hostQBytes = qBytes * 2;
  }
  uint8_t* hq = static_cast<uint8_t*>(f->hostQ);
  std::memcpy(hq + oXY, q_xy, Q * 8);
  std::memcpy(hq + oR, q_r, Q * 4);
  std::memcpy(hq + oMin, q_min_level, Q * 4);
  std::memcpy(hq + oMax, q_max_level, Q * 4);
  std::memcpy(hq + oD, q_desc, Q * 32);
  if ((rc = f->work.ensure(carve_size({qBytes, Q * 4, (Q + 1) * 4, (size_t)std::max(cand_cap, 1) * 4, (size_t)std::max(cand_cap, 1) * 4}))))
    return rc;
  Carver c{static_cast<uint8_t*>(f->work.p)};
  uint8_t* dQ = c.take<uint8_t>(qBytes);
  int* dCount = c.take<int>(Q);
  int* dOff = c.take<int>(Q + 1);
  int* dCi = c.take<int>(std::max(cand_cap, 1));
  int* dCd = c.take<int>(std::max(cand_cap, 1));
  Carver cg{static_cast<uint8_t*>(f->grid.p)};
  const int* dCell = cg.take<int>(grid_cells() + 1);
  const int* dItems = cg.take<int>(std::max(f->n, 1));
  cudaStream_t s = h->stream;
  const orbx_keypoint* dKU = static_cast<const orbx_keypoint*>(f->kpsUn.p);
  const uint8_t* dD = static_cast<const uint8_t*>(f->desc.p);
  const float* b = f->bounds;
  ORBX_CUDA_TRY(cudaMemcpyAsync(dQ, hq, qBytes, cudaMemcpyHostToDevice, s));
  auto window = [&](bool fill) {
    launch_window(fill, dKU, dD, b[0], b[1], b[2], b[3], dCell, dItems, reinterpret_cast<const float*>(dQ + oXY),
                  reinterpret_cast<const float*>(dQ + oR), reinterpret_cast<const int*>(dQ + oMin),
                  reinterpret_cast<const int*>(dQ + oMax), dQ + oD, nq, dCount, dOff, dCi, dCd, s);
  };
  window(false);
  launch_scan(dCount, nq, dOff, s);
  h->launches += 2;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(cand_off, dOff, (Q + 1) * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  const int total = cand_off[nq];
  *n_cand = total;
  if (total > cand_cap || (total > 0 && (!cand_idx || !cand_dist))) { set_error("candidate capacity %d < %d", cand_cap, total); return ORBX_ERR_CAPACITY; }
  if (total > 0) {
    window(true);
    h->launches++;
    ORBX_CUDA_TRY(cudaGetLastError());
    ORBX_CUDA_TRY(cudaMemcpyAsync(cand_idx, dCi, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(cand_dist, dCd, (size_t)total * 4, cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  }
  return ORBX_OK;
}

// ---- vocabulary -----------------------------------------------------------------------------
}  // extern "C"

struct orbx_vocabulary {
  orbx_matcher* m = nullptr;
  int nnodes = 0, depth = 0;
  DeviceBuf tree;   // node descriptors | child offsets | child ids
  DeviceBuf work;   // features | leaf node | level node
  size_t oOff = 0, oIdx = 0;
};

extern "C" {

int orbx_vocabulary_create(orbx_matcher* m, int32_t nnodes, const int32_t* parent, const uint8_t* desc, int32_t depth_L,
                           orbx_vocabulary** out) {
  int rc = set_dev(m);
  if (rc) return rc;
  if (!out || nnodes < 1 || !parent || !desc || depth_L < 1) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  *out = nullptr;
  // children in id order, like loadFromTextFile appends them (TemplatedVocabulary.h:1409)
  std::vector<int> off((size_t)nnodes + 1, 0), idx((size_t)std::max(nnodes - 1, 1));
  for (int i = 1; i < nnodes; i++) {
    if (parent[i] < 0 || parent[i] >= i) { set_error("parent[%d] = %d: parents must precede their children", i, parent[i]); return ORBX_ERR_ARG; }
    off[parent[i] + 1]++;
  }
  for (int i = 0; i < nnodes; i++) {
    if (off[i + 1] > 256) { set_error("node %d has %d children (supported: 256)", i, off[i + 1]); return ORBX_ERR_ARG; }
    off[i + 1] += off[i];
  }
  std::vector<int> fill(off.begin(), off.end() - 1);
  for (int i = 1; i < nnodes; i++) idx[fill[parent[i]]++] = i;
  orbx_vocabulary* v = new (std::nothrow) orbx_vocabulary();
  if (!v) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  v->m = m; v->nnodes = nnodes; v->depth = depth_L;
  const size_t N = (size_t)nnodes;
  v->oOff = (N * 32 + 255) & ~size_t(255);
  v->oIdx = v->oOff + (((N + 1) * 4 + 255) & ~size_t(255));
  if ((rc = v->tree.ensure(v->oIdx + idx.size() * 4 + 256))) { delete v; return rc; }
  uint8_t* base = static_cast<uint8_t*>(v->tree.p);
  cudaStream_t s = m->stream;
  cudaError_t e = cudaMemcpyAsync(base, desc, N * 32, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(base + v->oOff, off.data(), (N + 1) * 4, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaMemcpyAsync(base + v->oIdx, idx.data(), idx.size() * 4, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);   // off / idx are locals
  if (e != cudaSuccess) { set_error("vocabulary upload failed: %s", cudaGetErrorString(e)); v->tree.release(); delete v; return ORBX_ERR_CUDA; }
  *out = v;
  return ORBX_OK;
}

void orbx_vocabulary_destroy(orbx_vocabulary* v) {
  if (!v) return;
  cudaSetDevice(v->m->device);
  cudaStreamSynchronize(v->m->stream);
  v->tree.release(); v->work.release();
  delete v;
}

int orbx_vocabulary_descend(orbx_vocabulary* v, const uint8_t* desc, int32_t n, int32_t levelsup, int32_t* leaf_node,
                            int32_t* level_node) {
  if (!v) { set_error("null vocabulary"); return ORBX_ERR_ARG; }
  int rc = set_dev(v->m);
  if (rc) return rc;
  if (n < 0 || (n > 0 && (!desc || !leaf_node || !level_node))) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  if (n == 0) return ORBX_OK;
  const size_t N = (size_t)n;
  if ((rc = v->work.ensure(carve_size({N * 32, N * 4, N * 4})))) return rc;
  Carver c{static_cast<uint8_t*>(v->work.p)};
  uint8_t* dF = c.take<uint8_t>(N * 32);
  int* dLeaf = c.take<int>(N);
  int* dLevel = c.take<int>(N);
  cudaStream_t s = v->m->stream;
  const uint8_t* base = static_cast<const uint8_t*>(v->tree.p);
  ORBX_CUDA_TRY(cudaMemcpyAsync(dF, desc, N * 32, cudaMemcpyHostToDevice, s));
  launch_voc_descend(dF, n, base, reinterpret_cast<const int*>(base + v->oOff), reinterpret_cast<const int*>(base + v->oIdx),
                     v->depth - levelsup, dLeaf, dLevel, s);
  v->m->launches++;
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpyAsync(leaf_node, dLeaf, N * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(level_node, dLevel, N * 4, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

}  // extern "C"
