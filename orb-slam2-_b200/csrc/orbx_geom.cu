// Host-side, shape-dependent precomputation: the ORBextractor constructor tables, the per-level
// geometry (level sizes, FAST cell grid, octree roots) and cv::resize's fixed-point coefficient
// tables. All of it is O(levels + width + height) work done once per shape; the float expressions
// keep the reference's types and operation order because they decide integer sizes downstream.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>

#include "orbx_host.h"

namespace orbx {

static thread_local char tlsError[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(tlsError, sizeof(tlsError), fmt, ap);
  va_end(ap);
}
const char* last_error() { return tlsError; }

static inline int cv_round_f(float v) { return (int)nearbyintf(v); }   // cvRound: ties to even
static inline int cv_round_d(double v) { return (int)nearbyint(v); }
static inline int cv_floor_f(float v) { int i = (int)v; return i - (i > v); }

// R/src/ORBextractor.cpp:435-500. `scaleFactor` is a double member initialised from the float
// ctor argument (R/include/ORBextractor.h:106), which matters for the products below.
void build_ctor_tables(int nfeatures, float scaleFactorArg, int nlevels, CtorTables* t) {
  const double scaleFactor = scaleFactorArg;
  t->scale.assign(nlevels, 1.f);
  t->sigma2.assign(nlevels, 1.f);
  for (int i = 1; i < nlevels; i++) {
    t->scale[i] = (float)(t->scale[i - 1] * scaleFactor);
    t->sigma2[i] = t->scale[i] * t->scale[i];
  }
  t->invScale.resize(nlevels);
  t->invSigma2.resize(nlevels);
  for (int i = 0; i < nlevels; i++) {
    t->invScale[i] = 1.0f / t->scale[i];
    t->invSigma2[i] = 1.0f / t->sigma2[i];
  }
  t->featsPerLevel.assign(nlevels, 0);
  const float factor = (float)(1.0f / scaleFactor);
  float nDesired = nfeatures * (1 - factor) / (1 - (float)pow((double)factor, (double)nlevels));
  int sum = 0;
  for (int level = 0; level < nlevels - 1; level++) {
    t->featsPerLevel[level] = cv_round_f(nDesired);
    sum += t->featsPerLevel[level];
    nDesired *= factor;
  }
  t->featsPerLevel[nlevels - 1] = std::max(nfeatures - sum, 0);

  const int HP = kHalfPatch;
  int v, v0;
  const int vmax = (int)floorf(HP * sqrtf(2.f) / 2 + 1);
  const int vmin = (int)ceilf(HP * sqrtf(2.f) / 2);
  const double hp2 = HP * HP;
  for (v = 0; v <= vmax; ++v) t->umax[v] = cv_round_d(sqrt(hp2 - v * v));
  for (v = HP, v0 = 0; v >= vmin; --v) {
    while (t->umax[v0] == t->umax[v0 + 1]) ++v0;
    t->umax[v] = v0;
    ++v0;
  }
}

// OpenCV imgproc/resize.cpp (INTER_LINEAR, 8U): per destination index the source index pair and
// the 11-bit weights; past the right/bottom edge the neighbour is clamped.
void build_resize_axis(int srcLen, int dstLen, std::vector<ResizeTap>* taps, bool horizontal) {
  const double scale = 1. / ((double)dstLen / srcLen);
  taps->resize(dstLen);
  for (int d = 0; d < dstLen; d++) {
    float f = (float)((d + 0.5) * scale - 0.5);
    int s = cv_floor_f(f);
    f -= s;
    ResizeTap t;
    if (horizontal) {
      // x: index clamped AND fraction zeroed at both ends (xofs/ialpha construction)
      if (s < 0) { f = 0; s = 0; }
      if (s >= srcLen - 1) { f = 0; s = srcLen - 1; }
      t.p0 = (short)s;
      t.p1 = (short)std::min(s + 1, srcLen - 1);
    } else {
      // y: fraction kept, the two row indices clipped independently (resizeGeneric_Invoker)
      t.p0 = (short)std::min(std::max(s, 0), srcLen - 1);
      t.p1 = (short)std::min(std::max(s + 1, 0), srcLen - 1);
    }
    t.c0 = (short)std::min(std::max(cv_round_f((1.f - f) * 2048), -32768), 32767);
    t.c1 = (short)std::min(std::max(cv_round_f(f * 2048), -32768), 32767);
    (*taps)[d] = t;
  }
}
void build_resize_axis(int srcLen, int dstLen, std::vector<ResizeTap>* taps) {
  build_resize_axis(srcLen, dstLen, taps, true);
}
void build_resize_axis_y(int srcLen, int dstLen, std::vector<ResizeTap>* taps) {
  build_resize_axis(srcLen, dstLen, taps, false);
}

bool build_resize_groups(const std::vector<ResizeTap>& tx, const std::vector<ResizeTap>& ty,
                         std::vector<ResizeGroup>* groups) {
  for (const ResizeTap& t : ty)  // resize4_kernel's vertical pass wants non-negative rows and weights
    if (t.p0 < 0 || t.p1 < 0 || t.c0 < 0 || t.c1 < 0) return false;
  const int dw = (int)tx.size();
  groups->assign((dw + 3) / 4, ResizeGroup());
  for (int gi = 0; gi < (int)groups->size(); gi++) {
    ResizeGroup& G = (*groups)[gi];
    const int s0 = tx[4 * gi].p0;  // p0 is non-decreasing along the row
    G.word = s0 >> 2;
    G.shift = 8 * (s0 & 3);
    uint32_t sel[4];
    for (int j = 0; j < 4; j++) {
      const ResizeTap& t = tx[std::min(4 * gi + j, dw - 1)];  // pixels past the row end repeat the last one
      const int o0 = t.p0 - s0, o1 = t.p1 - s0;
      if (o0 < 0 || o1 < 0 || o0 > 7 || o1 > 7 || t.c0 < 0 || t.c1 < 0) return false;
      sel[j] = (uint32_t)o0 | ((uint32_t)o1 << 4);
      G.c[j] = (uint32_t)(uint16_t)t.c0 | ((uint32_t)(uint16_t)t.c1 << 16);
    }
    G.sel01 = sel[0] | (sel[1] << 8);
    G.sel23 = sel[2] | (sel[3] << 8);
  }
  return true;
}

int build_geometry(const orbx_params& p, const CtorTables& t, int w, int h, int batchCap, FrameSetGeom* g,
                   std::vector<FastCellDesc>* cellTable) {
  *g = FrameSetGeom();
  if (cellTable) cellTable->clear();
  g->nlevels = p.nlevels;
  g->iniTh = std::min(std::max(p.ini_th_fast, 0), 255);  // cv::FAST clamps its threshold
  g->minTh = std::min(std::max(p.min_th_fast, 0), 255);
  long long planeOff = 0, candOff = 0, keptOff = 0;
  int cells = 0, tiles = 0;
  for (int l = 0; l < p.nlevels; l++) {
    LevelGeom& L = g->lv[l];
    const float inv = t.invScale[l];
    L.w = cv_round_f((float)w * inv);  // :1203
    L.h = cv_round_f((float)h * inv);
    if (L.w > 4095 || L.h > 4095) {
      set_error("level %d is %dx%d: sizes above 4095 are not supported (12-bit candidate packing)", l, L.w, L.h);
      return ORBX_ERR_ARG;
    }
    L.pitch = (L.w + 15) & ~15;
    L.maxBX = L.w - kEdge + 3;
    L.maxBY = L.h - kEdge + 3;
    const float width = (float)(L.maxBX - kMinBorder), height = (float)(L.maxBY - kMinBorder);
    L.nCols = (int)(width / 30.f);  // :844-847
    L.nRows = (int)(height / 30.f);
    if (L.nCols < 1 || L.nRows < 1) {
      set_error("level %d is %dx%d: too small for the reference's 30-pixel FAST grid "
                "(R/src/ORBextractor.cpp:844-847 divides by zero there)", l, L.w, L.h);
      return ORBX_ERR_ARG;
    }
    L.wCell = (int)ceilf(width / L.nCols);
    L.hCell = (int)ceilf(height / L.nRows);
    if (L.wCell > kMaxCellDim - 4 || L.hCell > kMaxCellDim - 4) {
      set_error("level %d: FAST cell %dx%d exceeds the kernel tile", l, L.wCell, L.hCell);
      return ORBX_ERR_ARG;
    }
    L.nfeat = t.featsPerLevel[l];
    L.nIni = (int)roundf(width / height);  // :577
    if (L.nIni < 1 || L.nIni > 64) {
      set_error("level %d: aspect ratio gives %d octree roots; the reference needs width >= height/2 "
                "(R/src/ORBextractor.cpp:577-580)", l, L.nIni);
      return ORBX_ERR_ARG;
    }
    L.hX = width / L.nIni;
    L.keptCap = std::max(std::max(L.nfeat + 3, 4 * L.nIni), 64) + 1;
    if (L.keptCap > kMaxNodes) {
      set_error("level %d: %d features exceed the octree kernel's node table (%d)", l, L.nfeat, kMaxNodes);
      return ORBX_ERR_ARG;
    }
    L.candCap = (int)(((long long)L.w * L.h) / 8) + 256;
    L.cellBase = cells;
    for (int ci = 0; ci < L.nRows; ci++) {
      for (int cj = 0; cj < L.nCols; cj++) {
        // the cell rectangle and the reference's skip rules (:851-870); cv::FAST scans [3,cw-3) x [3,ch-3) of it
        const int iniY = kMinBorder + ci * L.hCell, iniX = kMinBorder + cj * L.wCell;
        if (iniY >= L.maxBY - 3 || iniX >= L.maxBX - 6) continue;
        const int maxY = std::min(iniY + L.hCell + 6, L.maxBY), maxX = std::min(iniX + L.wCell + 6, L.maxBX);
        const int ch = maxY - iniY, sw = maxX - iniX - 6;
        if (sw <= 0 || ch - 6 <= 0) continue;
        cells++;
        if (!cellTable) continue;
        FastCellDesc d = {};
        const int gx0 = ((iniX + 3) & ~3) - 4, a = iniX + 3 - gx0;
        d.level = (uint8_t)l; d.ch = (uint8_t)ch; d.a = (uint8_t)a; d.b = (uint8_t)(a + sw);
        d.iniY = (uint16_t)iniY; d.gx0 = (uint16_t)gx0;
        d.outX0 = (int16_t)(cj * L.wCell + 3 - a);
        d.outY0 = (int16_t)(ci * L.hCell + 3);
        cellTable->push_back(d);
      }
    }
    blur_tiles_for(L.w, L.h, &L.tilesX, &L.tilesY);
    L.tileBase = tiles;
    tiles += L.tilesX * L.tilesY;
    L.scale = t.scale[l];
    L.patchSize = (float)(int)(kPatch * t.scale[l]);  // :905
    L.planeBytes = (long long)L.pitch * L.h;
    L.planeOff = planeOff;
    planeOff += L.planeBytes * batchCap;
    L.candOff = candOff;
    candOff += (long long)L.candCap * batchCap;
    L.keptOff = keptOff;
    keptOff += (long long)L.keptCap * batchCap;
    g->candPerFrame += L.candCap;
    g->keptPerFrame += L.keptCap;
  }
  g->totalCells = cells;
  g->totalTiles = tiles;
  g->pyrFrameStride = planeOff;  // total arena bytes for batchCap frames
  return ORBX_OK;
}

int DeviceBuf::ensure(size_t need) {
  if (need <= bytes) return ORBX_OK;
  if (p) { cudaFree(p); p = nullptr; bytes = 0; }
  ORBX_CUDA_TRY(cudaMalloc(&p, need));
  bytes = need;
  return ORBX_OK;
}
void DeviceBuf::release() {
  if (p) cudaFree(p);
  p = nullptr;
  bytes = 0;
}

}  // namespace orbx
