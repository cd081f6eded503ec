// ORB_SLAM2::ORBextractor over liborbx.so -- the class the reference declares in R/include/ORBextractor.h:45-123,
// re-declared with the same name, namespace, include guard and public surface (constructor, operator(), the six
// getters, the public mvImagePyramid member, the score-type enum), so that Frame.cpp / Tracking.cpp compile against
// this header unchanged (R/src/Frame.cpp:86-89,149,204,262-268,558-695; R/src/Tracking.cpp:130-149). Everything
// behind that surface is different: there is no pyramid / FAST / octree / BRIEF code on the host, only a handle to the
// CUDA library of include/orbx.h. There is no CPU path either: a missing device or a failing call aborts with the
// library's message (the reference's interface has no error channel: no exceptions, no return codes).
#ifndef ORBEXTRACTOR_H
#define ORBEXTRACTOR_H

#include <cstddef>
#include <vector>
#include <opencv/cv.h>

struct orbx_extractor;

namespace ORB_SLAM2 {

class ORBextractor {
 public:
  enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };

  ORBextractor(int nfeatures, float scaleFactor, int nlevels, int iniThFAST, int minThFAST);
  ~ORBextractor();

  // Keypoints (dispersed per level by the reference's octree rule) and 256-bit descriptors of one 8-bit image.
  // The mask is ignored, as in the reference.
  void operator()(cv::InputArray image, cv::InputArray mask, std::vector<cv::KeyPoint>& keypoints,
                  cv::OutputArray descriptors);

  // Scale tables, computed by the library with the reference's types and operation order
  // (R/src/ORBextractor.cpp:435-500) when the extractor is constructed.
  int GetLevels() const { return nlevels; }
  float GetScaleFactor() const { return (float)scaleFactor; }
  std::vector<float> GetScaleFactors() const { return mvScaleFactor; }
  std::vector<float> GetInverseScaleFactors() const { return mvInvScaleFactor; }
  std::vector<float> GetScaleSigmaSquares() const { return mvLevelSigma2; }
  std::vector<float> GetInverseScaleSigmaSquares() const { return mvInvLevelSigma2; }

  // Host copies of the pyramid levels of the last image. Level 0 aliases the caller's image
  // (R/src/ORBextractor.cpp:1223); the others alias a page-locked buffer of this extractor and are valid until its
  // next operator() call. Only the reference's own Frame::ComputeStereoMatches reads them: monocular / RGB-D
  // callers, and stereo callers that link host/FrameStereo.cc, switch the read-back off
  // (SetPyramidReadback(false) or ORBX_PYRAMID_READBACK=0) and save the device->host copy.
  std::vector<cv::Mat> mvImagePyramid;
  void SetPyramidReadback(bool on) { mbPyramidReadback = on; }

  // CUDA device the extractor runs on (default: ORBX_DEVICE from the environment, else 0); takes effect at the
  // next operator() call.
  void SetDevice(int device);

  // The liborbx handle behind this extractor: orbx_stereo_match reads the device pyramids of the last operator()
  // call through it (the device-side Frame::ComputeStereoMatches, host/FrameStereo.cc).
  orbx_extractor* handle() const { return mpHandle; }

 private:
  ORBextractor(const ORBextractor&);   // owns device and page-locked memory: not copyable
  ORBextractor& operator=(const ORBextractor&);

  void EnsureHandle(int width, int height);
  void EnsureOutputs(int cap);

  // constructor arguments
  int nfeatures, nlevels, iniThFAST, minThFAST;
  double scaleFactor;

  // per-level tables behind the getters (+ the two the reference keeps as protected members)
  std::vector<float> mvScaleFactor, mvInvScaleFactor, mvLevelSigma2, mvInvLevelSigma2;
  std::vector<int> mnFeaturesPerLevel, umax;

  // device side
  orbx_extractor* mpHandle;
  int mnDevice, mnMaxWidth, mnMaxHeight;

  // page-locked host side: backing store of mvImagePyramid[1..], result rows of the last call
  // (orbx_keypoint and cv::KeyPoint are the same 28-byte record)
  bool mbPyramidReadback;
  unsigned char* mpPyramidHost;
  std::size_t mnPyramidBytes;
  void* mpOutKeys;
  unsigned char* mpOutDesc;
  int mnOutCap;
};

}  // namespace ORB_SLAM2

#endif
