// TEST INFRASTRUCTURE (oracle/_ref). Shadows the vendored DBoW2 FORB.h: the bag-of-words vocabulary is out
// of the hot path; R/include/ORBVocabulary.h only needs the two names to form its typedef.
#pragma once
#include <opencv2/core/core.hpp>

namespace DBoW2 {
class FORB {
 public:
  typedef cv::Mat TDescriptor;
};
}  // namespace DBoW2
