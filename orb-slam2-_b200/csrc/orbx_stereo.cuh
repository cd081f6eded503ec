// Parameter blocks of the stereo kernels (orbx_stereo.cu), shared with the C ABI layer (orbx_api_match.cu).
#pragma once
#include "orbx_common.cuh"

namespace orbx {

// Pyramids of both sides (frame 0 of each; frame f is `stride` bytes further) and where pair p finds its frames
// and its keypoint rows: pyramid frame = pyrLeft0 + p*pyrStep, keypoint rows start at (arrLeft0 + p*arrStep)*cap.
struct StereoLevels {
  const uint8_t* left[kMaxLevels];
  const uint8_t* right[kMaxLevels];
  long long strideL[kMaxLevels], strideR[kMaxLevels];
  int pitchL[kMaxLevels], pitchR[kMaxLevels];
  int width[kMaxLevels];
  float scale[kMaxLevels], invScale[kMaxLevels];
  int rows0;  // mvImagePyramid[0].rows
  int pyrLeft0, pyrRight0, pyrStep;
  int arrLeft0, arrRight0, arrStep, cap;
  int fixedNL, fixedNR;  // keypoint counts when no per-frame count array is given
};

struct StereoArrays {
  const orbx_keypoint* kpL; const uint4* descL; const int* nL;  // nL / nR: per-frame counts (device) or null
  const orbx_keypoint* kpR; const uint4* descR; const int* nR;
  float* uRight; float* depth; int* sad; int* nKept;            // [pair][cap], nKept[pair]
  int* rowStart; int* rowItems; int itemCap;                    // [pair][rows0 + 1], [pair][itemCap]
};

// rows of the row table one right keypoint can occupy: 2*ceil(2*maxScale) + 3
int stereo_items_per_keypoint(float maxScale);
// rows -> match -> median for `pairs` pairs; maxNL bounds the left keypoint count of any pair. Returns a cudaError.
int launch_stereo(const StereoLevels& lv, const StereoArrays& a, int pairs, int maxNL, float mbf, float mb, cudaStream_t s);

}  // namespace orbx
