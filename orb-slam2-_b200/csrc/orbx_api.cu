// C ABI of liborbx.so, extractor half (include/orbx.h). Host orchestration only: shape-dependent
// setup, arena management, the launch sequence of the extraction pipeline, and host<->device copies
// for the host-buffer entry points. No CPU implementation of any stage lives here.
#include <cstring>
#include <new>

#include "orbx_host.h"

using namespace orbx;

namespace {

int check_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("no usable CUDA device (%s); liborbx has no CPU fallback", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    return ORBX_ERR_CUDA;
  }
  if (device < 0 || device >= n) {
    set_error("device %d out of range (have %d)", device, n);
    return ORBX_ERR_ARG;
  }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  return ORBX_OK;
}

// (Re)build geometry, coefficient tables and arenas for a frame shape / batch size.
int ensure_shape(orbx_extractor* h, int w, int hh, int count) {
  if (w <= 0 || hh <= 0 || count <= 0) { set_error("bad shape %dx%d x%d", w, hh, count); return ORBX_ERR_ARG; }
  if (w > h->params.max_width || hh > h->params.max_height || count > h->params.max_batch) {
    set_error("frame %dx%d x%d exceeds the handle's limits %dx%d x%d", w, hh, count, h->params.max_width,
              h->params.max_height, h->params.max_batch);
    return ORBX_ERR_ARG;
  }
  if (w == h->w && hh == h->h && count <= h->batchCap) return ORBX_OK;
  const int batchCap = h->params.max_batch;
  FrameSetGeom g;
  int rc = build_geometry(h->params, h->tables, w, hh, batchCap, &g);
  if (rc) return rc;
  const int nl = g.nlevels;
  long long candTotal = 0, keptTotal = 0;
  for (int l = 0; l < nl; l++) {
    candTotal += (long long)g.lv[l].candCap * batchCap;
    keptTotal += (long long)g.lv[l].keptCap * batchCap;
  }
  if ((rc = h->pyrArena.ensure((size_t)g.pyrFrameStride + 64))) return rc;
  if ((rc = h->blurArena.ensure((size_t)g.pyrFrameStride + 64))) return rc;
  if ((rc = h->candArena.ensure((size_t)candTotal * 4))) return rc;
  if ((rc = h->nodeArena.ensure((size_t)candTotal * 2))) return rc;
  if ((rc = h->keptArena.ensure((size_t)keptTotal * 4))) return rc;
  if ((rc = h->counts.ensure((size_t)(2 * batchCap * nl + 4) * sizeof(int)))) return rc;
  if ((rc = h->dGeom.ensure(sizeof(FrameSetGeom)))) return rc;
  if ((rc = h->outKps.ensure((size_t)batchCap * g.keptPerFrame * sizeof(orbx_keypoint)))) return rc;
  if ((rc = h->outDesc.ensure((size_t)batchCap * g.keptPerFrame * 32))) return rc;
  if ((rc = h->outN.ensure((size_t)batchCap * sizeof(int)))) return rc;
  const size_t stageBytes = (size_t)batchCap * g.keptPerFrame * (sizeof(orbx_keypoint) + 32) + (size_t)batchCap * 4 + 64;
  if (stageBytes > h->hostStageBytes) {
    if (h->hostStage) cudaFreeHost(h->hostStage);
    h->hostStage = nullptr;
    h->hostStageBytes = 0;
    ORBX_CUDA_TRY(cudaMallocHost(&h->hostStage, stageBytes));
    h->hostStageBytes = stageBytes;
  }
  // resize coefficient tables per destination level: x/y taps and (when they exist) 4-pixel groups
  size_t tapCount = 0, groupCount = 0;
  for (int l = 1; l < nl; l++) { tapCount += g.lv[l].w + g.lv[l].h; groupCount += (g.lv[l].w + 3) / 4; }
  const size_t groupBytes = groupCount * sizeof(ResizeGroup);
  if ((rc = h->tapArena.ensure(groupBytes + std::max<size_t>(tapCount, 1) * sizeof(ResizeTap)))) return rc;
  std::vector<ResizeGroup> groups, xg;
  std::vector<ResizeTap> all, tx, ty;
  groups.reserve(groupCount);
  all.reserve(tapCount);
  const ResizeGroup* gbase = static_cast<const ResizeGroup*>(h->tapArena.p);
  const ResizeTap* tbase = reinterpret_cast<const ResizeTap*>(static_cast<const uint8_t*>(h->tapArena.p) + groupBytes);
  for (int l = 1; l < nl; l++) {
    build_resize_axis(g.lv[l - 1].w, g.lv[l].w, &tx);
    build_resize_axis_y(g.lv[l - 1].h, g.lv[l].h, &ty);
    h->taps[l].xt = tbase + all.size();
    all.insert(all.end(), tx.begin(), tx.end());
    h->taps[l].yt = tbase + all.size();
    all.insert(all.end(), ty.begin(), ty.end());
    h->taps[l].xg = nullptr;
    if (build_resize_groups(tx, &xg)) {
      h->taps[l].xg = gbase + groups.size();
      groups.insert(groups.end(), xg.begin(), xg.end());
    }
  }
  if (!groups.empty())
    ORBX_CUDA_TRY(cudaMemcpyAsync(h->tapArena.p, groups.data(), groups.size() * sizeof(ResizeGroup), cudaMemcpyHostToDevice, h->stream));
  ORBX_CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(h->tapArena.p) + groupBytes, all.data(), all.size() * sizeof(ResizeTap),
                                cudaMemcpyHostToDevice, h->stream));
  ORBX_CUDA_TRY(cudaMemcpyAsync(h->dGeom.p, &g, sizeof(g), cudaMemcpyHostToDevice, h->stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(h->stream));  // `all`, `groups` and `g` are stack/heap temporaries
  h->geom = g;
  h->w = w; h->h = hh; h->batchCap = batchCap;
  h->haveResult = false;
  return ORBX_OK;
}

// The extraction pipeline on `count` frames whose level 0 is already on the device.
int run_pipeline(orbx_extractor* h, const uint8_t* dL0, int l0Pitch, long long l0FrameStride, int count,
                 orbx_keypoint* dKps, uint8_t* dDesc, int cap, int* dN, cudaStream_t s) {
  const FrameSetGeom& g = h->geom;
  const int nl = g.nlevels;
  PyramidPtrs pyr;
  for (int l = 0; l < kMaxLevels; l++) { pyr.lvl[l] = nullptr; pyr.frameStride[l] = 0; pyr.pitch[l] = 0; }
  pyr.lvl[0] = dL0; pyr.frameStride[0] = l0FrameStride; pyr.pitch[0] = l0Pitch;
  uint8_t* arena = static_cast<uint8_t*>(h->pyrArena.p);
  for (int l = 1; l < nl; l++) {
    pyr.lvl[l] = arena + g.lv[l].planeOff;
    pyr.frameStride[l] = g.lv[l].planeBytes;
    pyr.pitch[l] = g.lv[l].pitch;
  }
  int* candCount = static_cast<int*>(h->counts.p);
  int* keptCount = candCount + h->batchCap * nl;
  int* overflow = keptCount + h->batchCap * nl;
  const FrameSetGeom* dG = static_cast<const FrameSetGeom*>(h->dGeom.p);
  const bool prof = h->profiling;

  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[0], s));
  ORBX_CUDA_TRY(cudaMemsetAsync(candCount, 0, (size_t)(2 * h->batchCap * nl + 1) * sizeof(int), s));
  for (int l = 1; l < nl; l++) {
    launch_resize(pyr.lvl[l - 1], pyr.frameStride[l - 1], pyr.pitch[l - 1], g.lv[l - 1].w, g.lv[l - 1].h,
                  arena + g.lv[l].planeOff, g.lv[l].planeBytes, g.lv[l].pitch, g.lv[l].w, g.lv[l].h, h->taps[l],
                  count, s);
    h->launches++;
  }
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[1], s));
  int fe = launch_fast(pyr, g, static_cast<uint32_t*>(h->candArena.p), candCount, overflow, count, s);
  if (fe) { set_error("FAST kernel configuration failed: %s", cudaGetErrorString((cudaError_t)fe)); return ORBX_ERR_CUDA; }
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[2], s));
  int oe = launch_octree(dG, g, static_cast<const uint32_t*>(h->candArena.p), candCount,
                         static_cast<uint16_t*>(h->nodeArena.p), static_cast<uint32_t*>(h->keptArena.p), keptCount,
                         count, s);
  if (oe) { set_error("octree kernel configuration failed: %s", cudaGetErrorString((cudaError_t)oe)); return ORBX_ERR_CUDA; }
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[3], s));
  launch_blur(pyr, g, static_cast<uint8_t*>(h->blurArena.p), count, s);
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[4], s));
  launch_describe(pyr, static_cast<const uint8_t*>(h->blurArena.p), g, static_cast<const uint32_t*>(h->keptArena.p),
                  keptCount, dKps, dDesc, cap, dN, count, s);
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(h->ev[5], s));
  ORBX_CUDA_TRY(cudaGetLastError());
  h->lastCount = count;
  h->lastLevel0 = dL0; h->lastL0FrameStride = l0FrameStride; h->lastL0Pitch = l0Pitch;
  h->haveResult = true;
  return ORBX_OK;
}

int collect_profile(orbx_extractor* h) {
  if (!h->profiling) return ORBX_OK;
  ORBX_CUDA_TRY(cudaEventSynchronize(h->ev[5]));
  for (int i = 0; i < 5; i++) ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[i], h->ev[i], h->ev[i + 1]));
  ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[5], h->ev[0], h->ev[5]));
  return ORBX_OK;
}

int check_overflow(orbx_extractor* h, cudaStream_t s) {
  int* overflow = static_cast<int*>(h->counts.p) + 2 * h->batchCap * h->geom.nlevels;
  int v = 0;
  ORBX_CUDA_TRY(cudaMemcpyAsync(&v, overflow, sizeof(int), cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  if (v) { set_error("FAST candidate list overflowed its capacity (corner density above 1 per 8 pixels)"); return ORBX_ERR_CAPACITY; }
  return ORBX_OK;
}

}  // namespace

extern "C" {

const char* orbx_last_error(void) { return last_error(); }
const char* orbx_version(void) { return "orbx 0.1 sm_100a"; }
int orbx_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}

int orbx_extractor_create(const orbx_params* p, orbx_extractor** out) {
  if (!p || !out) { set_error("null argument"); return ORBX_ERR_ARG; }
  *out = nullptr;
  if (p->nfeatures <= 0 || p->nlevels <= 0 || p->nlevels > kMaxLevels || !(p->scale_factor > 1.0f) ||
      p->max_width <= 0 || p->max_height <= 0 || p->max_batch <= 0) {
    set_error("bad extractor parameters (nfeatures %d, nlevels %d, scale %g, max %dx%d x%d)", p->nfeatures,
              p->nlevels, (double)p->scale_factor, p->max_width, p->max_height, p->max_batch);
    return ORBX_ERR_ARG;
  }
  int rc = check_device(p->device);
  if (rc) return rc;
  orbx_extractor* h = new (std::nothrow) orbx_extractor();
  if (!h) { set_error("out of host memory"); return ORBX_ERR_ARG; }
  h->params = *p;
  build_ctor_tables(p->nfeatures, p->scale_factor, p->nlevels, &h->tables);
  cudaError_t e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
  for (int i = 0; i < 8 && e == cudaSuccess; i++) e = cudaEventCreate(&h->ev[i]);
  if (e != cudaSuccess) {
    set_error("stream/event creation failed: %s", cudaGetErrorString(e));
    orbx_extractor_destroy(h);
    return ORBX_ERR_CUDA;
  }
  // validate the largest shape now so that unsupported parameters fail at construction
  FrameSetGeom g;
  rc = build_geometry(h->params, h->tables, p->max_width, p->max_height, 1, &g);
  if (rc) { orbx_extractor_destroy(h); return rc; }
  *out = h;
  return ORBX_OK;
}

void orbx_extractor_destroy(orbx_extractor* h) {
  if (!h) return;
  cudaSetDevice(h->params.device);
  if (h->stream) cudaStreamSynchronize(h->stream);
  h->dGeom.release(); h->pyrArena.release(); h->blurArena.release(); h->candArena.release();
  h->nodeArena.release(); h->keptArena.release(); h->counts.release(); h->tapArena.release();
  h->outKps.release(); h->outDesc.release(); h->outN.release();
  if (h->hostStage) cudaFreeHost(h->hostStage);
  for (int i = 0; i < 8; i++) if (h->ev[i]) cudaEventDestroy(h->ev[i]);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

int orbx_extractor_tables(const orbx_extractor* h, float* sc, float* isc, float* s2, float* is2, int32_t* feats,
                          int32_t* umax16) {
  if (!h) { set_error("null handle"); return ORBX_ERR_ARG; }
  const int nl = h->params.nlevels;
  for (int i = 0; i < nl; i++) {
    if (sc) sc[i] = h->tables.scale[i];
    if (isc) isc[i] = h->tables.invScale[i];
    if (s2) s2[i] = h->tables.sigma2[i];
    if (is2) is2[i] = h->tables.invSigma2[i];
    if (feats) feats[i] = h->tables.featsPerLevel[i];
  }
  if (umax16) for (int i = 0; i < 16; i++) umax16[i] = h->tables.umax[i];
  return ORBX_OK;
}

int orbx_extract_batch_device(orbx_extractor* h, const uint8_t* d_imgs, int32_t count, int32_t width, int32_t height,
                              size_t row_stride, size_t frame_stride, orbx_keypoint* d_kps, uint8_t* d_desc,
                              int32_t cap, int32_t* d_n, void* stream) {
  if (!h || !d_imgs || !d_kps || !d_desc || !d_n || cap <= 0) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  int rc = ensure_shape(h, width, height, count);
  if (rc) return rc;
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : h->stream;
  if (row_stride < (size_t)width) { set_error("row_stride %zu < width %d", row_stride, width); return ORBX_ERR_ARG; }
  if ((reinterpret_cast<uintptr_t>(d_imgs) | row_stride | frame_stride) & 3) {
    // the kernels read rows as aligned 32-bit words: stage unaligned frames in the arena's level-0 slot
    const FrameSetGeom& g = h->geom;
    uint8_t* l0 = static_cast<uint8_t*>(h->pyrArena.p);
    for (int f = 0; f < count; f++)
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0 + (size_t)f * g.lv[0].planeBytes, g.lv[0].pitch, d_imgs + (size_t)f * frame_stride,
                                      row_stride, width, height, cudaMemcpyDeviceToDevice, s));
    return run_pipeline(h, l0, g.lv[0].pitch, g.lv[0].planeBytes, count, d_kps, d_desc, cap, d_n, s);
  }
  return run_pipeline(h, d_imgs, (int)row_stride, (long long)frame_stride, count, d_kps, d_desc, cap, d_n, s);
}

int orbx_extract_batch(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height,
                       size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n) {
  if (!h || !imgs || !kps || !desc || !n || cap <= 0) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  if (stride < (size_t)width) { set_error("stride %zu < width %d", stride, width); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  int rc = ensure_shape(h, width, height, count);
  if (rc) return rc;
  const FrameSetGeom& g = h->geom;
  cudaStream_t s = h->stream;
  uint8_t* l0 = static_cast<uint8_t*>(h->pyrArena.p);  // level-0 slot of the arena (planeOff 0)
  for (int f = 0; f < count; f++) {
    if (!imgs[f]) { set_error("imgs[%d] is null", f); return ORBX_ERR_ARG; }
    ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0 + (size_t)f * g.lv[0].planeBytes, g.lv[0].pitch, imgs[f], stride, width, height,
                                    cudaMemcpyHostToDevice, s));
  }
  const int rows = g.keptPerFrame;
  orbx_keypoint* dK = static_cast<orbx_keypoint*>(h->outKps.p);
  uint8_t* dD = static_cast<uint8_t*>(h->outDesc.p);
  int* dN = static_cast<int*>(h->outN.p);
  rc = run_pipeline(h, l0, g.lv[0].pitch, g.lv[0].planeBytes, count, dK, dD, rows, dN, s);
  if (rc) return rc;
  // one D2H per output array into pinned staging, then row copies into the caller's buffers
  uint8_t* st = static_cast<uint8_t*>(h->hostStage);
  orbx_keypoint* hK = reinterpret_cast<orbx_keypoint*>(st);
  uint8_t* hD = st + (size_t)h->batchCap * rows * sizeof(orbx_keypoint);
  int* hN = reinterpret_cast<int*>(hD + (size_t)h->batchCap * rows * 32);
  ORBX_CUDA_TRY(cudaMemcpyAsync(hN, dN, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(hK, dK, (size_t)count * rows * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(hD, dD, (size_t)count * rows * 32, cudaMemcpyDeviceToHost, s));
  rc = check_overflow(h, s);  // synchronises the stream
  if (rc) return rc;
  if ((rc = collect_profile(h))) return rc;
  bool tooSmall = false;
  for (int f = 0; f < count; f++) {
    n[f] = hN[f];
    if (hN[f] > cap) tooSmall = true;
  }
  if (tooSmall) { set_error("keypoint capacity %d too small (see n[])", cap); return ORBX_ERR_CAPACITY; }
  for (int f = 0; f < count; f++) {
    std::memcpy(kps + (size_t)f * cap, hK + (size_t)f * rows, (size_t)hN[f] * sizeof(orbx_keypoint));
    std::memcpy(desc + (size_t)f * cap * 32, hD + (size_t)f * rows * 32, (size_t)hN[f] * 32);
  }
  return ORBX_OK;
}

int orbx_extract(orbx_extractor* h, const uint8_t* img, int32_t width, int32_t height, size_t stride,
                 orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n) {
  const uint8_t* imgs[1] = {img};
  return orbx_extract_batch(h, imgs, 1, width, height, stride, kps, desc, cap, n);
}

}  // extern "C"
