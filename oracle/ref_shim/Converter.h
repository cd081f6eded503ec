// TEST INFRASTRUCTURE (oracle/_ref). Shadows R/include/Converter.h, whose Eigen and g2o includes cannot be
// satisfied here (Eigen is not installed): the translation units compiled into oracle/_ref
// (R/src/Frame.cpp) use exactly one member of it, declared below. It is never called by the functions the
// parity tests drive (Frame::ComputeBoW is out of the hot path), so it has no definition.
#pragma once
#include <vector>

#include <opencv2/core/core.hpp>

namespace ORB_SLAM2 {
class Converter {
 public:
  static std::vector<cv::Mat> toDescriptorVector(const cv::Mat& Descriptors);
};
}  // namespace ORB_SLAM2
