// TEST INFRASTRUCTURE (oracle/_ref). Shadows the vendored DBoW2 TemplatedVocabulary.h (needs cv::FileStorage
// and DUtils; out of the hot path). Declares the one member R/src/Frame.cpp names (ComputeBoW, never called
// by the parity harness), without a definition.
#pragma once
#include <algorithm>
#include <cmath>
#include <list>
#include <map>
#include <set>
#include <string>
#include <vector>

#include "Thirdparty/DBoW2/DBoW2/BowVector.h"
#include "Thirdparty/DBoW2/DBoW2/FeatureVector.h"

// the vendored header leaks this into every includer (its line 38) and the reference's headers rely on it
using namespace std;

namespace DBoW2 {
template <class TDescriptor, class F>
class TemplatedVocabulary {
 public:
  void transform(const std::vector<TDescriptor>& features, BowVector& v, FeatureVector& fv, int levelsup) const;
};
}  // namespace DBoW2
