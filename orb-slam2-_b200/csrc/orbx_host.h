// Host-side state behind the C ABI (not installed; include/orbx.h is the public surface).
#pragma once
#include <string>
#include <vector>

#include "orbx_common.cuh"

namespace orbx {

// ORBextractor ctor tables (R/src/ORBextractor.cpp:435-500), computed with the reference's types.
struct CtorTables {
  std::vector<float> scale, invScale, sigma2, invSigma2;
  std::vector<int> featsPerLevel;
  int umax[16];
};
void build_ctor_tables(int nfeatures, float scaleFactor, int nlevels, CtorTables* t);

// cv::resize INTER_LINEAR coefficient tables for one axis (OpenCV imgproc resize.cpp semantics).
void build_resize_axis(int srcLen, int dstLen, std::vector<ResizeTap>* taps);    // x: clamp + zero fraction
void build_resize_axis_y(int srcLen, int dstLen, std::vector<ResizeTap>* taps);  // y: rows clipped, fraction kept
// 4-pixel groups for resize4_kernel from the x taps; false when some group spans more than 8 source bytes or a
// weight is negative on either axis.
bool build_resize_groups(const std::vector<ResizeTap>& tx, const std::vector<ResizeTap>& ty,
                         std::vector<ResizeGroup>* groups);
const char* last_error();

// Fill the per-level geometry for a (w,h) frame and `batchCap` frames per arena.
// Returns 0, or ORBX_ERR_ARG with set_error() when a level is too small for the reference's grid.
// cellTable (optional): the FAST cell table of the shape, g->totalCells entries.
int build_geometry(const orbx_params& p, const CtorTables& t, int w, int h, int batchCap, FrameSetGeom* g,
                   std::vector<FastCellDesc>* cellTable = nullptr);

struct DeviceBuf {
  void* p = nullptr;
  size_t bytes = 0;
  int ensure(size_t need);  // grow-only; returns ORBX_OK / ORBX_ERR_CUDA
  void release();
};

}  // namespace orbx

namespace orbx {

constexpr int kMaxLanes = 8;

// One in-flight batch: a stream, its device arenas and pinned staging. The host-buffer entry points
// rotate over the lanes so that the H2D copy of chunk k+1, the kernels of chunk k and the D2H +
// host scatter of chunk k-1 overlap; the device-resident entry point uses lane 0 only.
struct Lane {
  cudaStream_t stream = nullptr;
  cudaEvent_t ev[8] = {};             // stage timing (profiling)
  cudaEvent_t done = nullptr;         // results of the chunk are in hostStage
  int w = 0, h = 0, batchCap = 0;     // shape the arenas are laid out for
  FrameSetGeom geom;                  // host copy (plane/candidate offsets depend on batchCap)
  DeviceBuf dGeom;                    // FrameSetGeom on device (octree kernel)
  DeviceBuf dCells;                   // FastCellDesc table of the shape (fast_cells_kernel)
  DeviceBuf pyrArena, blurArena, candArena, nodeArena, keptArena, counts;
  DeviceBuf outKps, outDesc, outN;    // device staging for the host-buffer entry points
  DeviceBuf stereoBuf;                // stereo sequences: uRight | depth | sad | nKept | row table of the chunk's pairs
  void* hostStage = nullptr;          // pinned: kps | desc | n | overflow flag
  size_t hostStageBytes = 0;
  // Small batches (one frame per call is what the reference's Frame constructor does): the levels are independent after
  // the pyramid, so FAST + octree of level l run on aux[l] as soon as level l exists, next to the rest of the resize
  // chain and the blur; the whole fork / join set is captured into a CUDA graph once its arguments repeat.
  cudaStream_t aux[kMaxLevels] = {};
  cudaEvent_t evFork[kMaxLevels] = {}, evJoin[kMaxLevels] = {};
  cudaGraphExec_t graphExec = nullptr;
  struct GraphKey {
    const void* p[12];
    long long stride;
    int v[8];
  } graphKey = {};
  bool graphKeySet = false;
  int graphKernels = 0;
  void release();
};

}  // namespace orbx

struct orbx_extractor {
  orbx_params params;
  orbx::CtorTables tables;
  orbx::Lane lane[orbx::kMaxLanes];
  bool profiling = false;
  int chunkFrames = 0, laneCount = 0;  // host-buffer pipeline shape; 0 = library default / environment
  float stageMs[6] = {};
  long long launches = 0;

  // resize tables of the current shape (shared by the lanes)
  int tapW = 0, tapH = 0;
  orbx::DeviceBuf tapArena;
  orbx::ResizeTables taps[orbx::kMaxLevels];

  // last batch run on lane 0 (introspection / mvImagePyramid read-back)
  int lastCount = 0;                  // frames in the last batch
  const uint8_t* lastLevel0 = nullptr;  // device pointer of level 0 of the last batch
  long long lastL0FrameStride = 0;
  int lastL0Pitch = 0;
  bool haveResult = false;

  // streaming submissions (orbx_extract_batch_submit / _wait): a ring of tickets, each with one event per lane it
  // touched, the result ranges to validate at wait time and pinned overflow words the chunks copy into
  static constexpr int kTickets = 8, kTicketChunks = 64;
  struct Ticket {
    bool open = false;
    int id = 0, lanesUsed = 0, chunks = 0, cap = 0, count = 0;
    const int* n = nullptr;
    cudaEvent_t done[orbx::kMaxLanes] = {};
    int* overflow = nullptr;   // pinned, kTicketChunks words
  };
  Ticket tickets[kTickets];
  int nextTicket = 1;
  unsigned laneRotor = 0;
};
