// Drop-in replacement of ONE member of R/src/Frame.cpp: Frame::ComputeStereoMatches (:551-770), over liborbx.so.
//
// The stereo constructor (R/src/Frame.cpp:67-128) extracts both images -- through host/ORBextractor.cc their pyramids
// stay on the device -- and then calls this member. The reference walks the left keypoints on the host: row-band
// candidates, best Hamming distance, 11x11 SAD over 11 offsets on the pyramid level of the keypoint, parabola fit,
// depth = mbf / disparity, median-SAD filter. orbx_stereo_match does all of that in three kernels on the two
// extractors' device pyramids and returns mvuRight / mvDepth (bit-identical: tests/test_gpu_vs_ref_front.py,
// tests/test_gpu_dropin_matcher.py).
//
// How it is linked: a maintainer deletes the member's body from Frame.cc and adds this file. The test build keeps
// R/src/Frame.cpp untouched instead: its object is passed through `objcopy --weaken-symbol` for this one member, so
// that the definition below wins at link time (oracle/Makefile, libdropin_front.so).
#include "Frame.h"

#include <cstdio>
#include <cstdlib>

#include "orbx.h"
#include "orbx_thread.h"

namespace ORB_SLAM2 {

static_assert(sizeof(cv::KeyPoint) == sizeof(orbx_keypoint), "cv::KeyPoint must be the 28-byte record of include/orbx.h");

void Frame::ComputeStereoMatches() {
  mvuRight = std::vector<float>(N, -1.0f);
  mvDepth = std::vector<float>(N, -1.0f);
  const int nRight = (int)mvKeysRight.size();
  if (N == 0 || nRight == 0) return;
  // `mb` is read as the reference reads it here: the constructor assigns mb = mbf/fx only after this call (:126), so
  // the reference's minZ / maxD (:592-594) see the mb(0) of the initialiser list (:69), i.e. maxD = +inf
  int nMatched = 0;
  const int rc = orbx_stereo_match(orbx_host::ThreadMatcher(), mpORBextractorLeft->handle(), mpORBextractorRight->handle(), 0, 0,
                                   reinterpret_cast<const orbx_keypoint*>(&mvKeys[0]), mDescriptors.ptr<unsigned char>(0), N,
                                   reinterpret_cast<const orbx_keypoint*>(&mvKeysRight[0]),
                                   mDescriptorsRight.ptr<unsigned char>(0), nRight, mbf, mb, &mvuRight[0], &mvDepth[0], &nMatched);
  if (rc != ORBX_OK) {
    std::fprintf(stderr, "Frame::ComputeStereoMatches (orbx): %s\n", orbx_last_error());
    std::abort();   // no CPU fallback
  }
}

}  // namespace ORB_SLAM2
