// C ABI of liborbx.so: lazy pyramid read-back (mvImagePyramid), introspection used by the parity
// tests, and the stand-alone primitive entry points (one kernel each, host buffers in and out).
#include <vector>

#include "orbx_host.h"

using namespace orbx;

namespace {

struct TempBuf {  // RAII device scratch for the stand-alone primitives
  void* p = nullptr;
  ~TempBuf() { if (p) cudaFree(p); }
  int alloc(size_t bytes) {
    ORBX_CUDA_TRY(cudaMalloc(&p, bytes ? bytes : 1));
    return ORBX_OK;
  }
  template <class T> T* as() { return static_cast<T*>(p); }
};

int use_device(int device) {
  int n = 0;
  cudaError_t e = cudaGetDeviceCount(&n);
  if (e != cudaSuccess || n <= 0) {
    set_error("no usable CUDA device (%s); liborbx has no CPU fallback", e == cudaSuccess ? "count 0" : cudaGetErrorString(e));
    return ORBX_ERR_CUDA;
  }
  if (device < 0 || device >= n) { set_error("device %d out of range (have %d)", device, n); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  return ORBX_OK;
}

int copy_plane_out(orbx_extractor* h, const uint8_t* dsrc, int pitch, int w, int hh, uint8_t* dst, size_t dstStride) {
  ORBX_CUDA_TRY(cudaMemcpy2DAsync(dst, dstStride, dsrc, pitch, w, hh, cudaMemcpyDeviceToHost, h->lane[0].stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(h->lane[0].stream));
  return ORBX_OK;
}

int check_result(const orbx_extractor* h, int frame, int level) {
  if (!h) { set_error("null handle"); return ORBX_ERR_ARG; }
  if (!h->haveResult) { set_error("no extraction has run on this handle yet"); return ORBX_ERR_STATE; }
  if (level < 0 || level >= h->lane[0].geom.nlevels || frame < 0 || frame >= h->lastCount) {
    set_error("frame %d / level %d out of range (%d frames, %d levels)", frame, level, h->lastCount, h->lane[0].geom.nlevels);
    return ORBX_ERR_ARG;
  }
  return ORBX_OK;
}

}  // namespace

extern "C" {

int orbx_level_size(const orbx_extractor* h, int32_t level, int32_t* width, int32_t* height) {
  if (!h || !width || !height) { set_error("null argument"); return ORBX_ERR_ARG; }
  if (h->lane[0].w == 0) { set_error("no extraction has run on this handle yet"); return ORBX_ERR_STATE; }
  if (level < 0 || level >= h->lane[0].geom.nlevels) { set_error("level %d out of range", level); return ORBX_ERR_ARG; }
  *width = h->lane[0].geom.lv[level].w;
  *height = h->lane[0].geom.lv[level].h;
  return ORBX_OK;
}

int orbx_get_level(orbx_extractor* h, int32_t frame, int32_t level, uint8_t* dst, size_t dst_stride) {
  int rc = check_result(h, frame, level);
  if (rc) return rc;
  if (!dst) { set_error("null dst"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  const LevelGeom& L = h->lane[0].geom.lv[level];
  if (level == 0)
    return copy_plane_out(h, h->lastLevel0 + (long long)frame * h->lastL0FrameStride, h->lastL0Pitch, L.w, L.h, dst, dst_stride);
  const uint8_t* src = static_cast<const uint8_t*>(h->lane[0].pyrArena.p) + L.planeOff + (long long)frame * L.planeBytes;
  return copy_plane_out(h, src, L.pitch, L.w, L.h, dst, dst_stride);
}

int orbx_get_pyramid(orbx_extractor* h, int32_t frame, uint8_t* dst, size_t dst_bytes, size_t* offsets, int32_t* pitches,
                     size_t* bytes) {
  int rc = check_result(h, frame, 0);
  if (rc) return rc;
  if (!offsets || !pitches || !bytes) { set_error("null argument"); return ORBX_ERR_ARG; }
  const FrameSetGeom& g = h->lane[0].geom;
  size_t total = 0;
  for (int l = 1; l < g.nlevels; l++) {
    offsets[l] = total;
    pitches[l] = g.lv[l].pitch;
    total += (size_t)g.lv[l].pitch * g.lv[l].h;
  }
  *bytes = total;
  if (!dst) return ORBX_OK;
  if (dst_bytes < total) { set_error("pyramid needs %zu bytes, dst has %zu", total, dst_bytes); return ORBX_ERR_CAPACITY; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  cudaStream_t s = h->lane[0].stream;
  const uint8_t* arena = static_cast<const uint8_t*>(h->lane[0].pyrArena.p);
  for (int l = 1; l < g.nlevels; l++) {
    const LevelGeom& L = g.lv[l];
    ORBX_CUDA_TRY(cudaMemcpyAsync(dst + offsets[l], arena + L.planeOff + (long long)frame * L.planeBytes,
                                  (size_t)L.pitch * L.h, cudaMemcpyDeviceToHost, s));
  }
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));
  return ORBX_OK;
}

int orbx_get_blurred_level(orbx_extractor* h, int32_t frame, int32_t level, uint8_t* dst, size_t dst_stride) {
  int rc = check_result(h, frame, level);
  if (rc) return rc;
  if (!dst) { set_error("null dst"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  const LevelGeom& L = h->lane[0].geom.lv[level];
  const uint8_t* src = static_cast<const uint8_t*>(h->lane[0].blurArena.p) + L.planeOff + (long long)frame * L.planeBytes;
  return copy_plane_out(h, src, L.pitch, L.w, L.h, dst, dst_stride);
}

int orbx_get_level_candidates(orbx_extractor* h, int32_t frame, int32_t level, int32_t* xys, int32_t cap, int32_t* n) {
  int rc = check_result(h, frame, level);
  if (rc) return rc;
  if (!n) { set_error("null n"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  const LevelGeom& L = h->lane[0].geom.lv[level];
  const int nl = h->lane[0].geom.nlevels;
  int cnt = 0;
  ORBX_CUDA_TRY(cudaMemcpyAsync(&cnt, static_cast<const int*>(h->lane[0].counts.p) + frame * nl + level, sizeof(int),
                                cudaMemcpyDeviceToHost, h->lane[0].stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(h->lane[0].stream));
  cnt = std::min(cnt, L.candCap);
  *n = cnt;
  const int m = std::min(cnt, cap);
  if (m > 0 && xys) {
    std::vector<uint32_t> tmp(m);
    ORBX_CUDA_TRY(cudaMemcpyAsync(tmp.data(), static_cast<const uint32_t*>(h->lane[0].candArena.p) + L.candOff + (long long)frame * L.candCap,
                                  (size_t)m * 4, cudaMemcpyDeviceToHost, h->lane[0].stream));
    ORBX_CUDA_TRY(cudaStreamSynchronize(h->lane[0].stream));
    for (int i = 0; i < m; i++) {
      xys[3 * i] = cand_x(tmp[i]); xys[3 * i + 1] = cand_y(tmp[i]); xys[3 * i + 2] = cand_score(tmp[i]);
    }
  }
  return ORBX_OK;
}

int64_t orbx_launch_count(const orbx_extractor* h) { return h ? h->launches : 0; }

int orbx_extractor_set_pipeline(orbx_extractor* h, int32_t chunk_frames, int32_t lanes) {
  if (!h || chunk_frames < 0 || lanes < 0 || lanes > orbx::kMaxLanes) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  h->chunkFrames = chunk_frames;
  h->laneCount = lanes;
  return ORBX_OK;
}

int orbx_set_profiling(orbx_extractor* h, int32_t enable) {
  if (!h) { set_error("null handle"); return ORBX_ERR_ARG; }
  h->profiling = enable != 0;
  return ORBX_OK;
}

int orbx_last_stage_ms(orbx_extractor* h, float* ms6) {
  if (!h || !ms6) { set_error("null argument"); return ORBX_ERR_ARG; }
  if (!h->profiling || !h->haveResult) { set_error("profiling not enabled or nothing extracted"); return ORBX_ERR_STATE; }
  const Lane& L = h->lane[0];
  ORBX_CUDA_TRY(cudaEventSynchronize(L.ev[5]));
  for (int i = 0; i < 5; i++) ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[i], L.ev[i], L.ev[i + 1]));
  ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[5], L.ev[0], L.ev[5]));
  for (int i = 0; i < 6; i++) ms6[i] = h->stageMs[i];
  return ORBX_OK;
}

// ---- stand-alone primitives ------------------------------------------------------------------

int orbx_resize_u8(int32_t device, const uint8_t* src, int32_t sw, int32_t sh, size_t sstride, uint8_t* dst,
                   int32_t dw, int32_t dh, size_t dstride) {
  if (!src || !dst || sw <= 0 || sh <= 0 || dw <= 0 || dh <= 0 || sw > 32767 || sh > 32767) {
    set_error("bad resize arguments");
    return ORBX_ERR_ARG;
  }
  int rc = use_device(device);
  if (rc) return rc;
  std::vector<ResizeTap> tx, ty;
  build_resize_axis(sw, dw, &tx);
  build_resize_axis_y(sh, dh, &ty);
  std::vector<ResizeGroup> xg;
  const bool haveGroups = build_resize_groups(tx, ty, &xg);
  TempBuf dS, dD, dT, dG;
  const int sp = (sw + 15) & ~15, dp = (dw + 15) & ~15;
  if ((rc = dS.alloc((size_t)sp * sh + 64)) || (rc = dD.alloc((size_t)dp * dh)) ||
      (rc = dT.alloc((tx.size() + ty.size()) * sizeof(ResizeTap))) || (rc = dG.alloc(xg.size() * sizeof(ResizeGroup)))) return rc;
  ORBX_CUDA_TRY(cudaMemcpy2D(dS.p, sp, src, sstride, sw, sh, cudaMemcpyHostToDevice));
  ORBX_CUDA_TRY(cudaMemcpy(dT.p, tx.data(), tx.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
  ORBX_CUDA_TRY(cudaMemcpy(dT.as<ResizeTap>() + tx.size(), ty.data(), ty.size() * sizeof(ResizeTap), cudaMemcpyHostToDevice));
  if (haveGroups) ORBX_CUDA_TRY(cudaMemcpy(dG.p, xg.data(), xg.size() * sizeof(ResizeGroup), cudaMemcpyHostToDevice));
  ResizeTables t{dT.as<ResizeTap>(), dT.as<ResizeTap>() + tx.size(), haveGroups ? dG.as<ResizeGroup>() : nullptr};
  launch_resize(dS.as<uint8_t>(), 0, sp, sw, sh, dD.as<uint8_t>(), 0, dp, dw, dh, t, 1, 0);
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpy2D(dst, dstride, dD.p, dp, dw, dh, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_blur7_u8(int32_t device, const uint8_t* src, int32_t w, int32_t h, size_t sstride, uint8_t* dst, size_t dstride) {
  if (!src || !dst || w < 4 || h < 4) { set_error("bad blur arguments (need w,h >= 4)"); return ORBX_ERR_ARG; }
  int rc = use_device(device);
  if (rc) return rc;
  TempBuf dS, dD;
  const int p = (w + 15) & ~15;
  if ((rc = dS.alloc((size_t)p * h)) || (rc = dD.alloc((size_t)p * h))) return rc;
  ORBX_CUDA_TRY(cudaMemcpy2D(dS.p, p, src, sstride, w, h, cudaMemcpyHostToDevice));
  launch_blur_single(dS.as<uint8_t>(), p, dD.as<uint8_t>(), p, w, h, 0);
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpy2D(dst, dstride, dD.p, p, w, h, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_fast_atan2(int32_t device, const float* y, const float* x, float* out, int64_t n) {
  if (!y || !x || !out || n < 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  int rc = use_device(device);
  if (rc) return rc;
  TempBuf dY, dX, dO;
  if ((rc = dY.alloc(n * 4)) || (rc = dX.alloc(n * 4)) || (rc = dO.alloc(n * 4))) return rc;
  ORBX_CUDA_TRY(cudaMemcpy(dY.p, y, n * 4, cudaMemcpyHostToDevice));
  ORBX_CUDA_TRY(cudaMemcpy(dX.p, x, n * 4, cudaMemcpyHostToDevice));
  launch_atan2(dY.as<float>(), dX.as<float>(), dO.as<float>(), n, 0);
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpy(out, dO.p, n * 4, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_sincosf(int32_t device, const float* a, float* out_sin, float* out_cos, int64_t n) {
  if (!a || !out_sin || !out_cos || n < 0) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  int rc = use_device(device);
  if (rc) return rc;
  TempBuf dA, dS, dC;
  if ((rc = dA.alloc(n * 4)) || (rc = dS.alloc(n * 4)) || (rc = dC.alloc(n * 4))) return rc;
  ORBX_CUDA_TRY(cudaMemcpy(dA.p, a, n * 4, cudaMemcpyHostToDevice));
  launch_sincosf(dA.as<float>(), dS.as<float>(), dC.as<float>(), n, 0);
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpy(out_sin, dS.p, n * 4, cudaMemcpyDeviceToHost));
  ORBX_CUDA_TRY(cudaMemcpy(out_cos, dC.p, n * 4, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_sincosf_sweep(int32_t device, uint32_t lo_bits, uint32_t hi_bits, float* out_sin, float* out_cos) {
  if (!out_sin || !out_cos || hi_bits < lo_bits) { set_error("bad arguments"); return ORBX_ERR_ARG; }
  int rc = use_device(device);
  if (rc) return rc;
  const int64_t n = (int64_t)hi_bits - lo_bits;
  TempBuf dS, dC;
  if ((rc = dS.alloc(n * 4)) || (rc = dC.alloc(n * 4))) return rc;
  launch_sincosf_sweep(lo_bits, hi_bits, dS.as<float>(), dC.as<float>(), 0);
  ORBX_CUDA_TRY(cudaGetLastError());
  ORBX_CUDA_TRY(cudaMemcpy(out_sin, dS.p, n * 4, cudaMemcpyDeviceToHost));
  ORBX_CUDA_TRY(cudaMemcpy(out_cos, dC.p, n * 4, cudaMemcpyDeviceToHost));
  return ORBX_OK;
}

int orbx_distribute(int32_t device, const int32_t* xys, int32_t n, int32_t min_x, int32_t max_x, int32_t min_y,
                    int32_t max_y, int32_t n_features, int32_t cell_w, int32_t cell_h, int32_t* out_xys,
                    int32_t cap, int32_t* n_out) {
  if (!xys || !out_xys || !n_out || n < 0 || n_features <= 0 || max_x <= min_x || max_y <= min_y || cell_w <= 0 ||
      cell_h <= 0) {
    set_error("bad arguments");
    return ORBX_ERR_ARG;
  }
  const int width = max_x - min_x, height = max_y - min_y;
  if (width > 4095 || height > 4095) { set_error("extent above 4095"); return ORBX_ERR_ARG; }
  const int nIni = (int)roundf((float)width / (float)height);
  if (nIni < 1 || nIni > 64) { set_error("aspect ratio gives %d roots", nIni); return ORBX_ERR_ARG; }
  const int keptCap = std::max(std::max(n_features + 3, 4 * nIni), 64) + 1;
  if (keptCap > kMaxNodes) { set_error("n_features too large"); return ORBX_ERR_ARG; }
  int rc = use_device(device);
  if (rc) return rc;
  std::vector<uint32_t> packed(std::max(n, 1));
  for (int i = 0; i < n; i++) {
    const int x = xys[3 * i], y = xys[3 * i + 1], s = xys[3 * i + 2];
    if (x < 0 || x >= width || y < 0 || y >= height || s < 0 || s > 255) { set_error("candidate %d out of range", i); return ORBX_ERR_ARG; }
    packed[i] = pack_cand(x, y, s);
  }
  TempBuf dC, dNode, dK, dCnt;
  if ((rc = dC.alloc((size_t)std::max(n, 1) * 4)) || (rc = dNode.alloc((size_t)std::max(n, 1) * 2)) ||
      (rc = dK.alloc((size_t)keptCap * 4)) || (rc = dCnt.alloc(4))) return rc;
  ORBX_CUDA_TRY(cudaMemcpy(dC.p, packed.data(), (size_t)std::max(n, 1) * 4, cudaMemcpyHostToDevice));
  const int nCols = (width + cell_w - 1) / cell_w + 1;
  int e = launch_octree_single(dC.as<uint32_t>(), n, width, height, n_features, cell_w, cell_h, nCols,
                               dNode.as<uint16_t>(), dK.as<uint32_t>(), dCnt.as<int>(), keptCap, 0);
  if (e) { set_error("octree launch failed: %s", cudaGetErrorString((cudaError_t)e)); return ORBX_ERR_CUDA; }
  ORBX_CUDA_TRY(cudaGetLastError());
  int cnt = 0;
  ORBX_CUDA_TRY(cudaMemcpy(&cnt, dCnt.p, 4, cudaMemcpyDeviceToHost));
  if (cnt < 0) { set_error("octree node table overflow"); return ORBX_ERR_CAPACITY; }
  *n_out = cnt;
  if (cnt > cap) { set_error("output capacity %d < %d", cap, cnt); return ORBX_ERR_CAPACITY; }
  std::vector<uint32_t> kept(std::max(cnt, 1));
  ORBX_CUDA_TRY(cudaMemcpy(kept.data(), dK.p, (size_t)cnt * 4, cudaMemcpyDeviceToHost));
  for (int i = 0; i < cnt; i++) {
    out_xys[3 * i] = cand_x(kept[i]); out_xys[3 * i + 1] = cand_y(kept[i]); out_xys[3 * i + 2] = cand_score(kept[i]);
  }
  return ORBX_OK;
}

}  // extern "C"
