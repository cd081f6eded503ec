// C ABI of liborbx.so, extractor half (include/orbx.h). Host orchestration only: shape-dependent
// setup, arena management, the launch sequence of the extraction pipeline, and host<->device copies
// for the host-buffer entry points. No CPU implementation of any stage lives here.
//
// Host-buffer batches are cut into chunks of kChunkDefault frames that rotate over up to kMaxLanes lanes (stream +
// arenas + pinned staging each): while chunk k runs its kernels, chunk k+1's frames cross PCIe and
// chunk k-1's keypoints/descriptors are scattered into the caller's arrays by the host thread.
#include <algorithm>
#include <cstdlib>
#include <chrono>
#include <cstdio>
#include <cstring>
#include <new>

#include "orbx_host.h"
#include "orbx_stereo.cuh"

using namespace orbx;

namespace {

constexpr int ChunkTrace_kMaxChunks = 256;  // most chunks one orbx_extract_batch call is cut into

// Chunking of a host-buffer batch: fixed chunks of kChunkDefault frames rotate over the lanes. Launches over few
// frames fill the GPU badly (tools/batch_probe.py: 512 frames cost 3.6 ms in one launch set, 3.9 ms as 4 x 128,
// 4.3 ms as 8 x 64, 5.1 ms as 16 x 32), large chunks lengthen the pipeline's fill (first H2D) and drain (last
// D2H); 64 measured best with 4 lanes (profiles/r1_q_e2e_probe.txt; a ramp-up / taper-off schedule of chunk sizes
// was tried and was no better). ORBX_CHUNK=<n> overrides (tuning knob; any chunking gives identical results).
constexpr int kChunkDefault = 64;                         // at 640x480; other shapes keep the bytes per chunk
constexpr long long kChunkBytes = 64LL * 640 * 480;       // ~20 MB of frames per chunk

int chunk_frames(int width, int height) {
  static const int v = [] {
    const char* e = std::getenv("ORBX_CHUNK");
    const int x = e ? std::atoi(e) : 0;
    return x > 0 ? x : 0;
  }();
  if (v) return v;
  return (int)std::max(4LL, std::min((long long)kChunkDefault, kChunkBytes / ((long long)width * height)));
}

struct ChunkPlan {
  int first[ChunkTrace_kMaxChunks], count[ChunkTrace_kMaxChunks];
  int n = 0, largest = 0;
  void add(int f0, int c) { first[n] = f0; count[n] = c; n++; largest = std::max(largest, c); }
};

bool plan_chunks(int total, int chunk, ChunkPlan* p) {
  for (int f0 = 0; f0 < total; f0 += chunk) {
    if (p->n >= ChunkTrace_kMaxChunks) return false;
    p->add(f0, std::min(chunk, total - f0));
  }
  return true;
}

// The blocking call pays the H2D of its first chunk before any kernel runs and the kernels + D2H of its last chunk
// after the last copy: ramp the chunk size up at the head and down at the tail (chunk/4, chunk/2, chunk ... chunk,
// chunk/2, chunk/4) so that both exposed pieces are a quarter chunk. Small chunks use the GPU less well, but they
// run next to full-sized ones on the other lanes.
bool plan_chunks_tapered(int total, int chunk, ChunkPlan* p) {
  const int q = chunk / 4, hq = chunk / 2;
  if (q < 8 || total < 3 * chunk) return plan_chunks(total, chunk, p);
  const int body = total - 2 * (q + hq);
  if (2 + (body + chunk - 1) / chunk + 2 > ChunkTrace_kMaxChunks) return false;
  int f0 = 0;
  p->add(f0, q); f0 += q;
  p->add(f0, hq); f0 += hq;
  for (int left = body; left > 0; left -= chunk) { const int c = std::min(chunk, left); p->add(f0, c); f0 += c; }
  p->add(f0, hq); f0 += hq;
  p->add(f0, q);
  return true;
}

// ORBX_TRACE=1: per-chunk device timeline of orbx_extract_batch (H2D / kernels / D2H begin-end and the host time of
// the submit), printed to stderr after the call. A debugging aid for the host-buffer pipeline; off by default.
struct ChunkTrace {
  static constexpr int kMax = 64;
  bool on = false, init = false;
  cudaEvent_t ev[kMax][4];
  double hostSubmit[kMax], hostDone[kMax];
  int n = 0;
  std::chrono::steady_clock::time_point t0;
  bool enabled() {
    if (!init) {
      init = true;
      const char* e = std::getenv("ORBX_TRACE");
      on = e && std::atoi(e) > 0;
      if (on) for (auto& c : ev) for (auto& x : c) cudaEventCreate(&x);
    }
    return on;
  }
  double now() const { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); }
  void mark(int stage, cudaStream_t s) { if (on && n < kMax) cudaEventRecord(ev[n][stage], s); }
  void dump() {
    if (!on || n == 0) return;
    std::fprintf(stderr, "orbx trace: chunk  host_submit  h2d_begin  h2d_end  kernels_end  d2h_end  host_collected  (ms)\n");
    for (int c = 0; c < std::min(n, (int)kMax); c++) {
      float t[4];
      for (int k = 0; k < 4; k++) cudaEventElapsedTime(&t[k], ev[0][0], ev[c][k]);
      std::fprintf(stderr, "orbx trace: %5d  %11.3f  %9.3f  %7.3f  %11.3f  %7.3f  %14.3f\n", c, hostSubmit[c], t[0], t[1], t[2], t[3], hostDone[c]);
    }
  }
};
thread_local ChunkTrace g_trace;

int lane_count() {  // chunks in flight; ORBX_LANES overrides (1..kMaxLanes), default 4
  static const int v = [] {
    const char* e = std::getenv("ORBX_LANES");
    const int x = e ? std::atoi(e) : 0;
    return x > 0 ? std::min(x, kMaxLanes) : 4;
  }();
  return v;
}

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

int init_lane(Lane& L) {
  if (L.stream) return ORBX_OK;
  ORBX_CUDA_TRY(cudaStreamCreateWithFlags(&L.stream, cudaStreamNonBlocking));
  for (int i = 0; i < 8; i++) ORBX_CUDA_TRY(cudaEventCreate(&L.ev[i]));
  ORBX_CUDA_TRY(cudaEventCreateWithFlags(&L.done, cudaEventDisableTiming));
  return ORBX_OK;
}

// resize coefficient tables for a frame shape: x/y taps and (when they exist) 4-pixel groups per level
int ensure_taps(orbx_extractor* h, const FrameSetGeom& g, int w, int hh, cudaStream_t s) {
  if (h->tapW == w && h->tapH == hh) return ORBX_OK;
  // both lanes may still be reading the old tables
  for (Lane& L : h->lane) if (L.stream) ORBX_CUDA_TRY(cudaStreamSynchronize(L.stream));
  const int nl = g.nlevels;
  size_t tapCount = 0, groupCount = 0;
  for (int l = 1; l < nl; l++) { tapCount += g.lv[l].w + g.lv[l].h; groupCount += (g.lv[l].w + 3) / 4; }
  const size_t groupBytes = groupCount * sizeof(ResizeGroup);
  int rc;
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
    if (build_resize_groups(tx, ty, &xg)) {
      h->taps[l].xg = gbase + groups.size();
      groups.insert(groups.end(), xg.begin(), xg.end());
    }
  }
  if (!groups.empty())
    ORBX_CUDA_TRY(cudaMemcpyAsync(h->tapArena.p, groups.data(), groups.size() * sizeof(ResizeGroup), cudaMemcpyHostToDevice, s));
  if (!all.empty())
    ORBX_CUDA_TRY(cudaMemcpyAsync(static_cast<uint8_t*>(h->tapArena.p) + groupBytes, all.data(), all.size() * sizeof(ResizeTap),
                                  cudaMemcpyHostToDevice, s));
  ORBX_CUDA_TRY(cudaStreamSynchronize(s));  // `all` and `groups` are temporaries; the other lane may read next
  h->tapW = w; h->tapH = hh;
  return ORBX_OK;
}

// (Re)build geometry and arenas of a lane for a frame shape and a capacity of `cap` frames.
int ensure_lane(orbx_extractor* h, Lane& L, int w, int hh, int cap) {
  if (w <= 0 || hh <= 0 || cap <= 0) { set_error("bad shape %dx%d x%d", w, hh, cap); return ORBX_ERR_ARG; }
  if (w > h->params.max_width || hh > h->params.max_height || cap > h->params.max_batch) {
    set_error("frame %dx%d x%d exceeds the handle's limits %dx%d x%d", w, hh, cap, h->params.max_width,
              h->params.max_height, h->params.max_batch);
    return ORBX_ERR_ARG;
  }
  int rc;
  if ((rc = init_lane(L))) return rc;
  if (w == L.w && hh == L.h && cap <= L.batchCap) return ensure_taps(h, L.geom, w, hh, L.stream);
  ORBX_CUDA_TRY(cudaStreamSynchronize(L.stream));
  const int batchCap = (w == L.w && hh == L.h) ? std::max(cap, L.batchCap) : cap;
  FrameSetGeom g;
  std::vector<FastCellDesc> cellTable;
  if ((rc = build_geometry(h->params, h->tables, w, hh, batchCap, &g, &cellTable))) return rc;
  const int nl = g.nlevels;
  long long candTotal = 0, keptTotal = 0;
  for (int l = 0; l < nl; l++) {
    candTotal += (long long)g.lv[l].candCap * batchCap;
    keptTotal += (long long)g.lv[l].keptCap * batchCap;
  }
  if ((rc = L.pyrArena.ensure((size_t)g.pyrFrameStride + 64))) return rc;
  if ((rc = L.blurArena.ensure((size_t)g.pyrFrameStride + 64))) return rc;
  if ((rc = L.candArena.ensure((size_t)candTotal * 4))) return rc;
  if ((rc = L.nodeArena.ensure((size_t)candTotal * 2))) return rc;
  if ((rc = L.keptArena.ensure((size_t)keptTotal * 4))) return rc;
  if ((rc = L.counts.ensure((size_t)(2 * batchCap * nl + 4) * sizeof(int)))) return rc;
  if ((rc = L.dGeom.ensure(sizeof(FrameSetGeom)))) return rc;
  if ((rc = L.dCells.ensure(std::max<size_t>(cellTable.size(), 1) * sizeof(FastCellDesc)))) return rc;
  ORBX_CUDA_TRY(cudaMemcpyAsync(L.dGeom.p, &g, sizeof(g), cudaMemcpyHostToDevice, L.stream));
  if (!cellTable.empty())
    ORBX_CUDA_TRY(cudaMemcpyAsync(L.dCells.p, cellTable.data(), cellTable.size() * sizeof(FastCellDesc), cudaMemcpyHostToDevice, L.stream));
  ORBX_CUDA_TRY(cudaStreamSynchronize(L.stream));  // `g` and the table are stack temporaries
  L.geom = g;
  L.w = w; L.h = hh; L.batchCap = batchCap;
  if (&L == &h->lane[0]) h->haveResult = false;
  return ensure_taps(h, L.geom, w, hh, L.stream);
}

// device + pinned staging for the host-buffer entry points
int ensure_lane_staging(Lane& L) {
  const FrameSetGeom& g = L.geom;
  int rc;
  if ((rc = L.outKps.ensure((size_t)L.batchCap * g.keptPerFrame * sizeof(orbx_keypoint)))) return rc;
  if ((rc = L.outDesc.ensure((size_t)L.batchCap * g.keptPerFrame * 32))) return rc;
  if ((rc = L.outN.ensure((size_t)L.batchCap * sizeof(int)))) return rc;
  const size_t stageBytes = (size_t)L.batchCap * g.keptPerFrame * (sizeof(orbx_keypoint) + 32) + (size_t)L.batchCap * 4 + 64;
  if (stageBytes > L.hostStageBytes) {
    if (L.hostStage) cudaFreeHost(L.hostStage);
    L.hostStage = nullptr;
    L.hostStageBytes = 0;
    ORBX_CUDA_TRY(cudaMallocHost(&L.hostStage, stageBytes));
    L.hostStageBytes = stageBytes;
  }
  return ORBX_OK;
}

// Frames per call up to which the fork / join form of the pipeline is used (ORBX_FORK=0 disables it).
int fork_frames() {
  static const int v = [] {
    const char* e = std::getenv("ORBX_FORK");
    return e ? std::atoi(e) : 4;
  }();
  return v;
}

struct PipelineArgs {
  PyramidPtrs pyr;
  int count, cap;
  orbx_keypoint* dKps;
  uint8_t* dDesc;
  int* dN;
};

// One stream after the other: what batches use (every kernel fills the GPU by itself; the programmatic launches hide
// the launch latencies). Stage events for orbx_last_stage_ms when profiling.
int enqueue_linear(orbx_extractor* h, Lane& L, const PipelineArgs& a, cudaStream_t s) {
  const FrameSetGeom& g = L.geom;
  const int nl = g.nlevels, count = a.count;
  uint8_t* arena = static_cast<uint8_t*>(L.pyrArena.p);
  int* candCount = static_cast<int*>(L.counts.p);
  int* keptCount = candCount + L.batchCap * nl;
  int* overflow = keptCount + L.batchCap * nl;
  const FrameSetGeom* dG = static_cast<const FrameSetGeom*>(L.dGeom.p);
  const bool prof = h->profiling;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[0], s));
  ORBX_CUDA_TRY(cudaMemsetAsync(candCount, 0, (size_t)(2 * L.batchCap * nl + 1) * sizeof(int), s));
  for (int l = 1; l < nl; l++) {
    launch_resize(a.pyr.lvl[l - 1], a.pyr.frameStride[l - 1], a.pyr.pitch[l - 1], g.lv[l - 1].w, g.lv[l - 1].h,
                  arena + g.lv[l].planeOff, g.lv[l].planeBytes, g.lv[l].pitch, g.lv[l].w, g.lv[l].h, h->taps[l],
                  count, s);
    h->launches++;
  }
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[1], s));
  int fe = launch_fast(a.pyr, g, static_cast<const FastCellDesc*>(L.dCells.p), static_cast<uint32_t*>(L.candArena.p), candCount, overflow, count, s);
  if (fe) { set_error("FAST kernel configuration failed: %s", cudaGetErrorString((cudaError_t)fe)); return ORBX_ERR_CUDA; }
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[2], s));
  int oe = launch_octree(dG, g, static_cast<const uint32_t*>(L.candArena.p), candCount,
                         static_cast<uint16_t*>(L.nodeArena.p), static_cast<uint32_t*>(L.keptArena.p), keptCount,
                         count, s);
  if (oe) { set_error("octree kernel configuration failed: %s", cudaGetErrorString((cudaError_t)oe)); return ORBX_ERR_CUDA; }
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[3], s));
  launch_blur(a.pyr, g, static_cast<uint8_t*>(L.blurArena.p), count, s);
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[4], s));
  launch_describe(a.pyr, static_cast<const uint8_t*>(L.blurArena.p), g, static_cast<const uint32_t*>(L.keptArena.p),
                  keptCount, candCount, a.dKps, a.dDesc, a.cap, a.dN, count, s);
  h->launches++;
  if (prof) ORBX_CUDA_TRY(cudaEventRecord(L.ev[5], s));
  ORBX_CUDA_TRY(cudaGetLastError());
  return ORBX_OK;
}

// Fork / join: s carries the memset, the resize chain and the blur; aux[l] carries FAST + octree of level l from the
// moment level l exists (level 0: at once); describe joins everything. One frame is far too small to fill the GPU, so
// its latency is the longest dependent chain: 7 resizes + FAST + octree + blur + describe in a row before (107 us at
// 640x480), max(resizes + blur, l resizes + FAST(l) + octree(l)) + describe now. Plain launches (no programmatic
// edges): the set is meant to be captured. *kernels = number of kernel launches enqueued.
int enqueue_forked(orbx_extractor* h, Lane& L, const PipelineArgs& a, cudaStream_t s, int* kernels) {
  const FrameSetGeom& g = L.geom;
  const int nl = g.nlevels, count = a.count;
  uint8_t* arena = static_cast<uint8_t*>(L.pyrArena.p);
  int* candCount = static_cast<int*>(L.counts.p);
  int* keptCount = candCount + L.batchCap * nl;
  int* overflow = keptCount + L.batchCap * nl;
  const FrameSetGeom* dG = static_cast<const FrameSetGeom*>(L.dGeom.p);
  int n = 0;
  static const bool keepPdl = [] { const char* e = std::getenv("ORBX_FORK_PDL"); return e && std::atoi(e) > 0; }();
  struct PdlOff { PdlOff() { tl_programmatic_launch = keepPdl; } ~PdlOff() { tl_programmatic_launch = true; } } pdlOff;
  ORBX_CUDA_TRY(cudaMemsetAsync(candCount, 0, (size_t)(2 * L.batchCap * nl + 1) * sizeof(int), s));
  for (int l = 0; l < nl; l++) {
    if (l > 0) {
      launch_resize(a.pyr.lvl[l - 1], a.pyr.frameStride[l - 1], a.pyr.pitch[l - 1], g.lv[l - 1].w, g.lv[l - 1].h,
                    arena + g.lv[l].planeOff, g.lv[l].planeBytes, g.lv[l].pitch, g.lv[l].w, g.lv[l].h, h->taps[l],
                    count, s);
      n++;
    }
    ORBX_CUDA_TRY(cudaEventRecord(L.evFork[l], s));
    ORBX_CUDA_TRY(cudaStreamWaitEvent(L.aux[l], L.evFork[l], 0));
    int fe = launch_fast(a.pyr, g, static_cast<const FastCellDesc*>(L.dCells.p), static_cast<uint32_t*>(L.candArena.p), candCount, overflow, count, L.aux[l], l, l + 1);
    if (fe) { set_error("FAST kernel configuration failed: %s", cudaGetErrorString((cudaError_t)fe)); return ORBX_ERR_CUDA; }
    int oe = launch_octree(dG, g, static_cast<const uint32_t*>(L.candArena.p), candCount, static_cast<uint16_t*>(L.nodeArena.p),
                           static_cast<uint32_t*>(L.keptArena.p), keptCount, count, L.aux[l], l, l + 1);
    if (oe) { set_error("octree kernel configuration failed: %s", cudaGetErrorString((cudaError_t)oe)); return ORBX_ERR_CUDA; }
    n += 2;
    ORBX_CUDA_TRY(cudaEventRecord(L.evJoin[l], L.aux[l]));
  }
  launch_blur(a.pyr, g, static_cast<uint8_t*>(L.blurArena.p), count, s);
  n++;
  for (int l = 0; l < nl; l++) ORBX_CUDA_TRY(cudaStreamWaitEvent(s, L.evJoin[l], 0));
  launch_describe(a.pyr, static_cast<const uint8_t*>(L.blurArena.p), g, static_cast<const uint32_t*>(L.keptArena.p),
                  keptCount, candCount, a.dKps, a.dDesc, a.cap, a.dN, count, s);
  n++;
  ORBX_CUDA_TRY(cudaGetLastError());
  *kernels = n;
  return ORBX_OK;
}

int ensure_fork_resources(Lane& L, int nl) {
  for (int l = 0; l < nl; l++) {
    if (!L.aux[l]) ORBX_CUDA_TRY(cudaStreamCreateWithFlags(&L.aux[l], cudaStreamNonBlocking));
    if (!L.evFork[l]) ORBX_CUDA_TRY(cudaEventCreateWithFlags(&L.evFork[l], cudaEventDisableTiming));
    if (!L.evJoin[l]) ORBX_CUDA_TRY(cudaEventCreateWithFlags(&L.evJoin[l], cudaEventDisableTiming));
  }
  return ORBX_OK;
}

// The extraction pipeline on `count` frames whose level 0 is already on the device, on stream `s`.
int run_pipeline(orbx_extractor* h, Lane& L, const uint8_t* dL0, int l0Pitch, long long l0FrameStride, int count,
                 orbx_keypoint* dKps, uint8_t* dDesc, int cap, int* dN, cudaStream_t s) {
  const FrameSetGeom& g = L.geom;
  const int nl = g.nlevels;
  PipelineArgs a;
  for (int l = 0; l < kMaxLevels; l++) { a.pyr.lvl[l] = nullptr; a.pyr.frameStride[l] = 0; a.pyr.pitch[l] = 0; }
  a.pyr.lvl[0] = dL0; a.pyr.frameStride[0] = l0FrameStride; a.pyr.pitch[0] = l0Pitch;
  uint8_t* arena = static_cast<uint8_t*>(L.pyrArena.p);
  for (int l = 1; l < nl; l++) {
    a.pyr.lvl[l] = arena + g.lv[l].planeOff;
    a.pyr.frameStride[l] = g.lv[l].planeBytes;
    a.pyr.pitch[l] = g.lv[l].pitch;
  }
  a.count = count; a.cap = cap; a.dKps = dKps; a.dDesc = dDesc; a.dN = dN;

  int rc;
  cudaStreamCaptureStatus capturing = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(s, &capturing) != cudaSuccess) { cudaGetLastError(); capturing = cudaStreamCaptureStatusNone; }
  if (h->profiling || count > fork_frames() || nl < 2 || capturing != cudaStreamCaptureStatusNone) {
    // (a caller that captures this call into a graph of its own gets the single-stream form)
    if ((rc = enqueue_linear(h, L, a, s))) return rc;
  } else {
    if ((rc = ensure_fork_resources(L, nl))) return rc;
    // everything a captured launch set has baked in
    Lane::GraphKey key = {};
    const void* ptrs[12] = {dL0, dKps, dDesc, dN, L.pyrArena.p, L.blurArena.p, L.candArena.p, L.nodeArena.p, L.keptArena.p,
                            L.counts.p, L.dCells.p, h->tapArena.p};
    for (int i = 0; i < 12; i++) key.p[i] = ptrs[i];
    key.stride = l0FrameStride;
    const int vals[8] = {l0Pitch, count, cap, L.w, L.h, L.batchCap, h->tapW, h->tapH};
    for (int i = 0; i < 8; i++) key.v[i] = vals[i];
    const bool same = L.graphKeySet && std::memcmp(&key, &L.graphKey, sizeof(key)) == 0;
    if (same && L.graphExec) {
      ORBX_CUDA_TRY(cudaGraphLaunch(L.graphExec, s));
      h->launches += L.graphKernels;
    } else if (same) {
      // second call in a row with the same arguments (the drop-in extractor: same arena slot, same result arrays):
      // capture the set once, replay it from now on
      cudaGraph_t graph = nullptr;
      ORBX_CUDA_TRY(cudaStreamBeginCapture(s, cudaStreamCaptureModeThreadLocal));
      int kernels = 0;
      rc = enqueue_forked(h, L, a, s, &kernels);
      const cudaError_t ce = cudaStreamEndCapture(s, &graph);
      if (rc || ce != cudaSuccess || !graph) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        if (!rc) { set_error("graph capture of the extraction pipeline failed: %s", cudaGetErrorString(ce)); rc = ORBX_ERR_CUDA; }
        return rc;
      }
      const cudaError_t ie = cudaGraphInstantiate(&L.graphExec, graph, 0);
      cudaGraphDestroy(graph);
      if (ie != cudaSuccess) { L.graphExec = nullptr; set_error("cudaGraphInstantiate failed: %s", cudaGetErrorString(ie)); return ORBX_ERR_CUDA; }
      L.graphKernels = kernels;
      ORBX_CUDA_TRY(cudaGraphLaunch(L.graphExec, s));
      h->launches += kernels;
    } else {
      if (L.graphExec) { cudaGraphExecDestroy(L.graphExec); L.graphExec = nullptr; }
      L.graphKey = key;
      L.graphKeySet = true;
      int kernels = 0;
      if ((rc = enqueue_forked(h, L, a, s, &kernels))) return rc;
      h->launches += kernels;
    }
  }
  if (&L == &h->lane[0]) {
    h->lastCount = count;
    h->lastLevel0 = dL0; h->lastL0FrameStride = l0FrameStride; h->lastL0Pitch = l0Pitch;
    h->haveResult = true;
  }
  return ORBX_OK;
}

struct HostStage {  // layout of Lane::hostStage
  orbx_keypoint* kps;
  uint8_t* desc;
  int* n;
  int* overflow;
};
HostStage stage_of(const Lane& L) {
  uint8_t* st = static_cast<uint8_t*>(L.hostStage);
  const size_t rows = (size_t)L.batchCap * L.geom.keptPerFrame;
  HostStage s;
  s.kps = reinterpret_cast<orbx_keypoint*>(st);
  s.desc = st + rows * sizeof(orbx_keypoint);
  s.n = reinterpret_cast<int*>(s.desc + rows * 32);
  s.overflow = s.n + L.batchCap;
  return s;
}

// H2D + pipeline + D2H of one chunk, all asynchronous on the lane's stream
// Where a chunk's rows go on the host. Pinned caller arrays take the D2H copies directly (row pitch cap);
// pageable ones go through the lane's pinned staging and are scattered by collect_chunk.
struct HostOut {
  orbx_keypoint* kps;
  uint8_t* desc;
  int cap;
  bool direct;
};

bool is_pinned(const void* p, size_t bytes) {
  cudaPointerAttributes a0, a1;
  if (cudaPointerGetAttributes(&a0, p) != cudaSuccess || cudaPointerGetAttributes(&a1, static_cast<const char*>(p) + bytes - 1) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a0.type == cudaMemoryTypeHost && a1.type == cudaMemoryTypeHost;
}

// Stereo sequences (orbx_stereo_batch_submit): the chunk's frames are L R L R ...; after the extraction the lane runs
// Frame::ComputeStereoMatches for its count/2 pairs against its own pyramid arena and copies the rows out.
struct StereoOut {
  float mbf, mb;
  float* uRight;   // host, [pairs][cap]
  float* depth;
  int* nMatched;   // host, [pairs]
  int cap;
};

int run_chunk_stereo(orbx_extractor* h, Lane& L, const uint8_t* l0, int count, const orbx_keypoint* dK, const uint8_t* dD,
                     const int* dN, const StereoOut& so) {
  const FrameSetGeom& g = L.geom;
  const int nlv = g.nlevels, rows = g.keptPerFrame, pairs = count / 2;
  if (rows > 65535) { set_error("more than 65535 keypoint rows per frame"); return ORBX_ERR_ARG; }
  StereoLevels lv;
  for (int l = 0; l < kMaxLevels; l++) {
    lv.left[l] = lv.right[l] = nullptr;
    lv.strideL[l] = lv.strideR[l] = 0;
    lv.pitchL[l] = lv.pitchR[l] = lv.width[l] = 0;
    lv.scale[l] = lv.invScale[l] = 1.f;
  }
  const uint8_t* arena = static_cast<const uint8_t*>(L.pyrArena.p);
  for (int l = 0; l < nlv; l++) {
    const LevelGeom& G = g.lv[l];
    lv.left[l] = lv.right[l] = l == 0 ? l0 : arena + G.planeOff;
    lv.strideL[l] = lv.strideR[l] = G.planeBytes;
    lv.pitchL[l] = lv.pitchR[l] = G.pitch;
    lv.width[l] = G.w;
    lv.scale[l] = h->tables.scale[l];
    lv.invScale[l] = h->tables.invScale[l];
  }
  lv.rows0 = g.lv[0].h;
  lv.pyrLeft0 = lv.arrLeft0 = 0;
  lv.pyrRight0 = lv.arrRight0 = 1;
  lv.pyrStep = lv.arrStep = 2;
  lv.cap = rows;
  lv.fixedNL = lv.fixedNR = 0;
  const size_t P = (size_t)std::max(L.batchCap / 2, 1);
  const size_t itemCap = (size_t)rows * (size_t)stereo_items_per_keypoint(h->tables.scale[nlv - 1]);
  auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
  const size_t bUR = al(P * rows * 4), bSad = al(P * rows * 4), bN = al(P * 4), bRow = al(P * ((size_t)lv.rows0 + 1) * 4),
               bItems = al(P * itemCap * 4);
  int rc;
  if ((rc = L.stereoBuf.ensure(2 * bUR + bSad + bN + bRow + bItems))) return rc;
  uint8_t* base = static_cast<uint8_t*>(L.stereoBuf.p);
  StereoArrays a;
  a.kpL = a.kpR = dK;
  a.descL = a.descR = reinterpret_cast<const uint4*>(dD);
  a.nL = a.nR = dN;
  a.uRight = reinterpret_cast<float*>(base);
  a.depth = reinterpret_cast<float*>(base + bUR);
  a.sad = reinterpret_cast<int*>(base + 2 * bUR);
  a.nKept = reinterpret_cast<int*>(base + 2 * bUR + bSad);
  a.rowStart = reinterpret_cast<int*>(base + 2 * bUR + bSad + bN);
  a.rowItems = reinterpret_cast<int*>(base + 2 * bUR + bSad + bN + bRow);
  a.itemCap = (int)itemCap;
  cudaStream_t s = L.stream;
  const int le = launch_stereo(lv, a, pairs, rows, so.mbf, so.mb, s);
  if (le) { set_error("stereo kernel configuration failed: %s", cudaGetErrorString((cudaError_t)le)); return ORBX_ERR_CUDA; }
  h->launches += 3;
  const size_t take = (size_t)std::min(rows, so.cap) * 4;
  ORBX_CUDA_TRY(cudaMemcpy2DAsync(so.uRight, (size_t)so.cap * 4, a.uRight, (size_t)rows * 4, take, pairs, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpy2DAsync(so.depth, (size_t)so.cap * 4, a.depth, (size_t)rows * 4, take, pairs, cudaMemcpyDeviceToHost, s));
  ORBX_CUDA_TRY(cudaMemcpyAsync(so.nMatched, a.nKept, (size_t)pairs * 4, cudaMemcpyDeviceToHost, s));
  return ORBX_OK;
}

int submit_chunk(orbx_extractor* h, Lane& L, const uint8_t* const* imgs, int count, int width, int height, size_t stride,
                 const HostOut& out, int* nOut = nullptr, int* overflowOut = nullptr, const StereoOut* stereo = nullptr) {
  const FrameSetGeom& g = L.geom;
  cudaStream_t s = L.stream;
  uint8_t* l0 = static_cast<uint8_t*>(L.pyrArena.p);  // level-0 slot of the arena (planeOff 0)
  bool contiguous = true;
  if (g_trace.on && g_trace.n < ChunkTrace::kMax) g_trace.hostSubmit[g_trace.n] = g_trace.now();
  g_trace.mark(0, s);
  for (int f = 0; f < count; f++) {
    if (!imgs[f]) { set_error("imgs[%d] is null", f); return ORBX_ERR_ARG; }
    contiguous = contiguous && imgs[f] == imgs[0] + (size_t)f * stride * height;
  }
  if (contiguous && g.lv[0].planeBytes == (long long)g.lv[0].pitch * height) {
    // frames back to back on the host: one copy of count*height rows
    if (stride == (size_t)width && g.lv[0].pitch == width) {
      ORBX_CUDA_TRY(cudaMemcpyAsync(l0, imgs[0], (size_t)width * height * count, cudaMemcpyHostToDevice, s));
    } else if (stride == (size_t)width) {
      // tightly packed rows whose width is not the arena pitch (e.g. 1241): a row-by-row DMA over PCIe is slow, so
      // the bytes cross flat into the blur arena (unused until the blur stage) and are re-pitched on the device
      uint8_t* flat = static_cast<uint8_t*>(L.blurArena.p);
      ORBX_CUDA_TRY(cudaMemcpyAsync(flat, imgs[0], (size_t)width * height * count, cudaMemcpyHostToDevice, s));
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0, g.lv[0].pitch, flat, width, width, (size_t)height * count, cudaMemcpyDeviceToDevice, s));
    } else {
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0, g.lv[0].pitch, imgs[0], stride, width, (size_t)height * count, cudaMemcpyHostToDevice, s));
    }
  } else {
    for (int f = 0; f < count; f++)
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0 + (size_t)f * g.lv[0].planeBytes, g.lv[0].pitch, imgs[f], stride, width, height,
                                      cudaMemcpyHostToDevice, s));
  }
  g_trace.mark(1, s);
  const int rows = g.keptPerFrame;
  orbx_keypoint* dK = static_cast<orbx_keypoint*>(L.outKps.p);
  uint8_t* dD = static_cast<uint8_t*>(L.outDesc.p);
  int* dN = static_cast<int*>(L.outN.p);
  HostStage st = stage_of(L);
  // Small chunks (one frame per call is the reference's own pattern): the describe kernel writes keypoints, descriptors
  // and counts straight into the page-locked destination (the caller's arrays, or the lane's staging for pageable
  // ones) through their device aliases -- 75 KB of posted writes that overlap the kernel, instead of four dependent
  // DMA copies behind it (19 us of a 126 us call at 640x480). A truncated frame reports n = -1 (describe_kernel), so
  // the overflow word needs no copy either.
  bool zeroCopy = false;
  if (!stereo && count <= fork_frames()) {
    void *zk = nullptr, *zd = nullptr, *zn = nullptr;
    orbx_keypoint* hk = out.direct ? out.kps : st.kps;
    uint8_t* hd = out.direct ? out.desc : st.desc;
    int* hn = nOut ? nOut : st.n;
    if (cudaHostGetDevicePointer(&zk, hk, 0) == cudaSuccess && cudaHostGetDevicePointer(&zd, hd, 0) == cudaSuccess &&
        cudaHostGetDevicePointer(&zn, hn, 0) == cudaSuccess) {
      zeroCopy = true;
      *(overflowOut ? overflowOut : st.overflow) = 0;
      int rc = run_pipeline(h, L, l0, g.lv[0].pitch, g.lv[0].planeBytes, count, static_cast<orbx_keypoint*>(zk),
                            static_cast<uint8_t*>(zd), out.direct ? out.cap : rows, static_cast<int*>(zn), s);
      if (rc) return rc;
      g_trace.mark(2, s);
    } else {
      cudaGetLastError();
    }
  }
  int rc = ORBX_OK;
  if (!zeroCopy) {
    rc = run_pipeline(h, L, l0, g.lv[0].pitch, g.lv[0].planeBytes, count, dK, dD, rows, dN, s);
    if (rc) return rc;
    g_trace.mark(2, s);
    const int* dOverflow = static_cast<const int*>(L.counts.p) + 2 * L.batchCap * g.nlevels;
    // streaming submissions hand in pinned destinations of their own: the lane's staging may be rewritten by the next
    // chunk on this lane before anybody has looked at it
    ORBX_CUDA_TRY(cudaMemcpyAsync(nOut ? nOut : st.n, dN, (size_t)count * sizeof(int), cudaMemcpyDeviceToHost, s));
    ORBX_CUDA_TRY(cudaMemcpyAsync(overflowOut ? overflowOut : st.overflow, dOverflow, sizeof(int), cudaMemcpyDeviceToHost, s));
    if (out.direct) {
      // one "row" per frame: the first min(rows, cap) keypoints; entries past n[f] are unspecified
      const size_t take = (size_t)std::min(rows, out.cap);
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(out.kps, (size_t)out.cap * sizeof(orbx_keypoint), dK, (size_t)rows * sizeof(orbx_keypoint),
                                      take * sizeof(orbx_keypoint), count, cudaMemcpyDeviceToHost, s));
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(out.desc, (size_t)out.cap * 32, dD, (size_t)rows * 32, take * 32, count,
                                      cudaMemcpyDeviceToHost, s));
    } else {
      ORBX_CUDA_TRY(cudaMemcpyAsync(st.kps, dK, (size_t)count * rows * sizeof(orbx_keypoint), cudaMemcpyDeviceToHost, s));
      ORBX_CUDA_TRY(cudaMemcpyAsync(st.desc, dD, (size_t)count * rows * 32, cudaMemcpyDeviceToHost, s));
    }
  }
  if (stereo && (rc = run_chunk_stereo(h, L, l0, count, dK, dD, dN, *stereo))) return rc;
  g_trace.mark(3, s);
  if (g_trace.on) g_trace.n++;
  ORBX_CUDA_TRY(cudaEventRecord(L.done, s));
  return ORBX_OK;
}

// An error after the first chunk was enqueued: the other lanes' asynchronous copies into the caller's arrays must
// have landed before the call returns (the caller may free or reuse them).
void drain_lanes(orbx_extractor* h) {
  for (Lane& L : h->lane) if (L.stream) cudaStreamSynchronize(L.stream);
}

// wait for a submitted chunk and scatter its rows into the caller's arrays
int collect_chunk(orbx_extractor* h, Lane& L, int count, const HostOut& out, int* n, bool* tooSmall) {
  orbx_keypoint* kps = out.kps;
  uint8_t* desc = out.desc;
  const int cap = out.cap;
  ORBX_CUDA_TRY(cudaEventSynchronize(L.done));
  HostStage st = stage_of(L);
  if (*st.overflow) {
    set_error("FAST candidate list overflowed its capacity (corner density above 1 per 8 pixels)");
    return ORBX_ERR_CAPACITY;
  }
  if (h->profiling) {
    for (int i = 0; i < 5; i++) ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[i], L.ev[i], L.ev[i + 1]));
    ORBX_CUDA_TRY(cudaEventElapsedTime(&h->stageMs[5], L.ev[0], L.ev[5]));
  }
  const int rows = L.geom.keptPerFrame;
  for (int f = 0; f < count; f++) {
    const int nf = st.n[f];
    n[f] = nf;
    if (nf < 0) { set_error("FAST candidate list of frame %d overflowed its capacity", f); return ORBX_ERR_CAPACITY; }
    if (nf > cap) { *tooSmall = true; continue; }
    if (out.direct) continue;
    std::memcpy(kps + (size_t)f * cap, st.kps + (size_t)f * rows, (size_t)nf * sizeof(orbx_keypoint));
    std::memcpy(desc + (size_t)f * cap * 32, st.desc + (size_t)f * rows * 32, (size_t)nf * 32);
  }
  return ORBX_OK;
}

}  // namespace

void orbx::Lane::release() {
  if (stream) cudaStreamSynchronize(stream);
  dGeom.release(); dCells.release(); pyrArena.release(); blurArena.release(); candArena.release(); nodeArena.release();
  keptArena.release(); counts.release(); outKps.release(); outDesc.release(); outN.release(); stereoBuf.release();
  if (hostStage) cudaFreeHost(hostStage);
  hostStage = nullptr; hostStageBytes = 0;
  for (int i = 0; i < 8; i++) if (ev[i]) { cudaEventDestroy(ev[i]); ev[i] = nullptr; }
  if (graphExec) { cudaGraphExecDestroy(graphExec); graphExec = nullptr; }
  graphKeySet = false;
  for (int l = 0; l < kMaxLevels; l++) {
    if (aux[l]) { cudaStreamSynchronize(aux[l]); cudaStreamDestroy(aux[l]); aux[l] = nullptr; }
    if (evFork[l]) { cudaEventDestroy(evFork[l]); evFork[l] = nullptr; }
    if (evJoin[l]) { cudaEventDestroy(evJoin[l]); evJoin[l] = nullptr; }
  }
  if (done) { cudaEventDestroy(done); done = nullptr; }
  if (stream) { cudaStreamDestroy(stream); stream = nullptr; }
  w = h = batchCap = 0;
}

extern "C" {

const char* orbx_last_error(void) { return last_error(); }
const char* orbx_version(void) { return "orbx 0.3 sm_100a"; }

int orbx_host_alloc(int32_t device, size_t bytes, void** out) {
  if (!out || bytes == 0) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(device));
  ORBX_CUDA_TRY(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
  return ORBX_OK;
}

int orbx_host_free(void* p) {
  if (p) ORBX_CUDA_TRY(cudaFreeHost(p));
  return ORBX_OK;
}
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
  if ((rc = init_lane(h->lane[0]))) { orbx_extractor_destroy(h); return rc; }
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
  for (Lane& L : h->lane) L.release();
  h->tapArena.release();
  for (auto& t : h->tickets) {
    for (cudaEvent_t& e : t.done) if (e) cudaEventDestroy(e);
    if (t.overflow) cudaFreeHost(t.overflow);
  }
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
  if (row_stride < (size_t)width) { set_error("row_stride %zu < width %d", row_stride, width); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  Lane& L = h->lane[0];
  int rc = ensure_lane(h, L, width, height, count);
  if (rc) return rc;
  cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : L.stream;
  if ((reinterpret_cast<uintptr_t>(d_imgs) | row_stride | frame_stride) & 3) {
    // the kernels read rows as aligned 32-bit words: stage unaligned frames in the arena's level-0 slot
    const FrameSetGeom& g = L.geom;
    uint8_t* l0 = static_cast<uint8_t*>(L.pyrArena.p);
    if (frame_stride == row_stride * (size_t)height && g.lv[0].planeBytes == (long long)g.lv[0].pitch * height) {
      ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0, g.lv[0].pitch, d_imgs, row_stride, width, (size_t)height * count, cudaMemcpyDeviceToDevice, s));
    } else {
      for (int f = 0; f < count; f++)
        ORBX_CUDA_TRY(cudaMemcpy2DAsync(l0 + (size_t)f * g.lv[0].planeBytes, g.lv[0].pitch, d_imgs + (size_t)f * frame_stride,
                                        row_stride, width, height, cudaMemcpyDeviceToDevice, s));
    }
    return run_pipeline(h, L, l0, g.lv[0].pitch, g.lv[0].planeBytes, count, d_kps, d_desc, cap, d_n, s);
  }
  return run_pipeline(h, L, d_imgs, (int)row_stride, (long long)frame_stride, count, d_kps, d_desc, cap, d_n, s);
}

int orbx_extract_batch(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height,
                       size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n) {
  if (!h || !imgs || !kps || !desc || !n || cap <= 0 || count <= 0) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  if (stride < (size_t)width) { set_error("stride %zu < width %d", stride, width); return ORBX_ERR_ARG; }
  if (count > h->params.max_batch) {
    set_error("batch of %d frames exceeds the handle's max_batch %d", count, h->params.max_batch);
    return ORBX_ERR_ARG;
  }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  ChunkPlan plan;
  if (!plan_chunks_tapered(count, h->chunkFrames > 0 ? h->chunkFrames : chunk_frames(width, height), &plan)) { set_error("batch of %d frames needs more than %d chunks (raise ORBX_CHUNK)", count, ChunkTrace_kMaxChunks); return ORBX_ERR_ARG; }
  const int nchunks = plan.n;
  const int nlanes = std::min(nchunks, h->laneCount > 0 ? h->laneCount : lane_count());
  int rc;
  for (int i = 0; i < nlanes; i++) {
    if ((rc = ensure_lane(h, h->lane[i], width, height, plan.largest))) return rc;
    if ((rc = ensure_lane_staging(h->lane[i]))) return rc;
  }
  bool tooSmall = false;
  if (g_trace.enabled()) { g_trace.n = 0; g_trace.t0 = std::chrono::steady_clock::now(); }
  int collected = 0;
  const bool direct = is_pinned(kps, (size_t)count * cap * sizeof(orbx_keypoint)) && is_pinned(desc, (size_t)count * cap * 32);
  auto out_at = [&](int f0) { return HostOut{kps + (size_t)f0 * cap, desc + (size_t)f0 * cap * 32, cap, direct}; };
  for (int c = 0; c <= nchunks; c++) {
    if (c < nchunks) {
      const int f0 = plan.first[c], cnt = plan.count[c];
      if (c >= nlanes) {  // the lane's staging still holds chunk c-nlanes: scatter it before reusing the lane
        const int g0 = plan.first[c - nlanes], gc = plan.count[c - nlanes];
        if ((rc = collect_chunk(h, h->lane[c % nlanes], gc, out_at(g0), n + g0, &tooSmall))) { drain_lanes(h); return rc; }
        if (g_trace.on && collected < ChunkTrace::kMax) g_trace.hostDone[collected++] = g_trace.now();
      }
      if ((rc = submit_chunk(h, h->lane[c % nlanes], imgs + f0, cnt, width, height, stride, out_at(f0)))) { drain_lanes(h); return rc; }
    }
  }
  for (int c = std::max(nchunks - nlanes, 0); c < nchunks; c++) {
    const int f0 = plan.first[c], cnt = plan.count[c];
    if ((rc = collect_chunk(h, h->lane[c % nlanes], cnt, out_at(f0), n + f0, &tooSmall))) { drain_lanes(h); return rc; }
    if (g_trace.on && collected < ChunkTrace::kMax) g_trace.hostDone[collected++] = g_trace.now();
  }
  g_trace.dump();
  if (nchunks > 1) h->haveResult = false;  // introspection only covers single-chunk batches
  if (tooSmall) { set_error("keypoint capacity %d too small (see n[])", cap); return ORBX_ERR_CAPACITY; }
  return ORBX_OK;
}

}  // extern "C"

namespace {
// Common body of the streaming submissions. stereo != nullptr: the frames are L R L R ... and every chunk also runs
// the stereo matcher on its pairs (stereo->uRight / depth / nMatched are the arrays of the whole submission).
int submit_stream(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height, size_t stride,
                  orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n, const StereoOut* stereo, int32_t* ticket) {
  if (!h || !imgs || !kps || !desc || !n || !ticket || cap <= 0 || count <= 0) { set_error("null/invalid argument"); return ORBX_ERR_ARG; }
  if (stride < (size_t)width) { set_error("stride %zu < width %d", stride, width); return ORBX_ERR_ARG; }
  if (count > h->params.max_batch) {
    set_error("batch of %d frames exceeds the handle's max_batch %d", count, h->params.max_batch);
    return ORBX_ERR_ARG;
  }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  if (!is_pinned(kps, (size_t)count * cap * sizeof(orbx_keypoint)) || !is_pinned(desc, (size_t)count * cap * 32) ||
      !is_pinned(n, (size_t)count * sizeof(int32_t))) {
    set_error("streaming submissions need page-locked result arrays (orbx_host_alloc): kps, desc and n");
    return ORBX_ERR_ARG;
  }
  if (stereo) {
    const size_t pairs = (size_t)count / 2;
    if ((count & 1) || !stereo->uRight || !stereo->depth || !stereo->nMatched) { set_error("stereo submissions take an even number of frames (L R L R ...) and three result arrays"); return ORBX_ERR_ARG; }
    if (!is_pinned(stereo->uRight, pairs * cap * 4) || !is_pinned(stereo->depth, pairs * cap * 4) || !is_pinned(stereo->nMatched, pairs * 4)) {
      set_error("streaming submissions need page-locked result arrays (orbx_host_alloc): u_right, depth and n_matched");
      return ORBX_ERR_ARG;
    }
  }
  ChunkPlan plan;
  // without fill and drain to pay per call, larger chunks win: they fill the GPU better (tools/stream_probe.py:
  // 128 frames x 4 lanes 167.7 k frames/s, 64 x 4 143 k, 256 x 3 167.7 k at 640x480)
  int chunk = h->chunkFrames > 0 ? h->chunkFrames : 2 * chunk_frames(width, height);
  if (stereo) chunk = std::max(2, chunk & ~1);  // a pair never straddles two chunks
  if (!plan_chunks(count, chunk, &plan) || plan.n > orbx_extractor::kTicketChunks) {
    set_error("batch of %d frames needs more than %d chunks (raise the chunk size)", count, orbx_extractor::kTicketChunks);
    return ORBX_ERR_ARG;
  }
  orbx_extractor::Ticket& t = h->tickets[h->nextTicket % orbx_extractor::kTickets];
  if (t.open) { set_error("%d submissions in flight: wait for ticket %d first", orbx_extractor::kTickets, t.id); return ORBX_ERR_STATE; }
  const int nlanes = h->laneCount > 0 ? h->laneCount : lane_count();
  int rc;
  for (int i = 0; i < nlanes; i++) {
    // a change of shape or capacity synchronises the lane before its arenas are rebuilt (ensure_lane)
    if ((rc = ensure_lane(h, h->lane[i], width, height, plan.largest))) return rc;
    if ((rc = ensure_lane_staging(h->lane[i]))) return rc;
  }
  if (!t.overflow) ORBX_CUDA_TRY(cudaHostAlloc(reinterpret_cast<void**>(&t.overflow), orbx_extractor::kTicketChunks * sizeof(int), cudaHostAllocPortable));
  for (int i = 0; i < nlanes; i++)
    if (!t.done[i]) ORBX_CUDA_TRY(cudaEventCreateWithFlags(&t.done[i], cudaEventDisableTiming));
  for (int c = 0; c < plan.n; c++) t.overflow[c] = 0;
  bool used[kMaxLanes] = {};
  rc = ORBX_OK;
  int enqueued = 0;
  for (int c = 0; c < plan.n; c++) {
    const int f0 = plan.first[c], cnt = plan.count[c];
    const int li = (int)(h->laneRotor++ % (unsigned)nlanes);
    const HostOut out{kps + (size_t)f0 * cap, desc + (size_t)f0 * cap * 32, cap, true};
    StereoOut so;
    if (stereo) {
      so = *stereo;
      so.uRight += (size_t)(f0 / 2) * cap; so.depth += (size_t)(f0 / 2) * cap; so.nMatched += f0 / 2; so.cap = cap;
    }
    if ((rc = submit_chunk(h, h->lane[li], imgs + f0, cnt, width, height, stride, out, n + f0, t.overflow + c, stereo ? &so : nullptr))) break;
    used[li] = true;
    enqueued++;
  }
  if (rc) {
    // nothing of a failed submission stays in flight: the caller gets no ticket to wait on
    drain_lanes(h);
    return rc;
  }
  t.lanesUsed = 0;
  for (int i = 0; i < nlanes; i++)
    if (used[i]) { ORBX_CUDA_TRY(cudaEventRecord(t.done[i], h->lane[i].stream)); t.lanesUsed |= 1 << i; }
  t.open = true; t.id = h->nextTicket++; t.chunks = plan.n; t.cap = cap; t.count = count; t.n = n;
  h->haveResult = false;  // introspection covers synchronous single-chunk calls only
  *ticket = t.id;
  return ORBX_OK;
}
}  // namespace

extern "C" {

int orbx_extract_batch_submit(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height,
                              size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n, int32_t* ticket) {
  return submit_stream(h, imgs, count, width, height, stride, kps, desc, cap, n, nullptr, ticket);
}

int orbx_stereo_batch_submit(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height,
                             size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n, float mbf, float mb,
                             float* u_right, float* depth, int32_t* n_matched, int32_t* ticket) {
  const StereoOut so{mbf, mb, u_right, depth, n_matched, cap};
  return submit_stream(h, imgs, count, width, height, stride, kps, desc, cap, n, &so, ticket);
}

int orbx_extract_batch_wait(orbx_extractor* h, int32_t ticket) {
  if (!h) { set_error("null handle"); return ORBX_ERR_ARG; }
  ORBX_CUDA_TRY(cudaSetDevice(h->params.device));
  orbx_extractor::Ticket& t = h->tickets[((ticket % orbx_extractor::kTickets) + orbx_extractor::kTickets) % orbx_extractor::kTickets];
  if (!t.open || t.id != ticket) { set_error("ticket %d is not in flight", ticket); return ORBX_ERR_STATE; }
  for (int i = 0; i < kMaxLanes; i++)
    if (t.lanesUsed >> i & 1) ORBX_CUDA_TRY(cudaEventSynchronize(t.done[i]));
  t.open = false;
  for (int c = 0; c < t.chunks; c++)
    if (t.overflow[c]) {
      set_error("FAST candidate list overflowed its capacity (corner density above 1 per 8 pixels)");
      return ORBX_ERR_CAPACITY;
    }
  for (int f = 0; f < t.count; f++) {
    if (t.n[f] < 0) { set_error("FAST candidate list of frame %d overflowed its capacity", f); return ORBX_ERR_CAPACITY; }
    if (t.n[f] > t.cap) { set_error("keypoint capacity %d too small (see n[])", t.cap); return ORBX_ERR_CAPACITY; }
  }
  return ORBX_OK;
}

int orbx_extract(orbx_extractor* h, const uint8_t* img, int32_t width, int32_t height, size_t stride,
                 orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n) {
  const uint8_t* imgs[1] = {img};
  return orbx_extract_batch(h, imgs, 1, width, height, stride, kps, desc, cap, n);
}

}  // extern "C"
