// TEST INFRASTRUCTURE (oracle/_ref/libref_voc.so) -- C entry points around the REFERENCE's own vendored DBoW2:
// Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h (the real header, not the declaration-only stub libref_front.so is
// built with), FORB.cpp, ScoringObject.cpp, BowVector.cpp, FeatureVector.cpp and DUtils, compiled where they lie.
// Used only by tests/ to pin the vocabulary transform:
//   TemplatedVocabulary::loadFromTextFile                       TemplatedVocabulary.h:1362-1436
//   TemplatedVocabulary::transform(features, BowVector&, FeatureVector&, levelsup)   :1151-1228, :1242-1296
// which is what Frame::ComputeBoW / KeyFrame::ComputeBoW call (R/src/Frame.cpp:262-274, R/src/KeyFrame.cpp:66-75).
#include REAL_VOC_H
#include REAL_FORB_H

#include <cstdint>
#include <cstring>

typedef DBoW2::TemplatedVocabulary<DBoW2::FORB::TDescriptor, DBoW2::FORB> ORBVocabulary;   // R/include/ORBVocabulary.h:31

extern "C" {

void* ref_voc_load(const char* path) {
  ORBVocabulary* v = new ORBVocabulary();
  if (!v->loadFromTextFile(path)) { delete v; return nullptr; }
  return v;
}
void ref_voc_destroy(void* h) { delete static_cast<ORBVocabulary*>(h); }
int ref_voc_words(void* h) { return (int)static_cast<ORBVocabulary*>(h)->size(); }

// transform(features, v, fv, levelsup) on n 32-byte rows (Converter::toDescriptorVector = one 1x32 Mat per row).
// Outputs flattened in map order: BowVector (word id, value) pairs; FeatureVector node ids, CSR offsets, indices.
// Returns the number of words; counts of the feature vector through nNodes / nIdx. Capacities: n each.
int ref_voc_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values, uint32_t* nodes,
                      int32_t* off, uint32_t* idx, int* nNodes, int* nIdx) {
  ORBVocabulary& voc = *static_cast<ORBVocabulary*>(h);
  std::vector<cv::Mat> features(n);
  for (int i = 0; i < n; i++) {
    features[i].create(1, 32, CV_8U);
    std::memcpy(features[i].ptr(0), desc + (size_t)i * 32, 32);
  }
  DBoW2::BowVector v;
  DBoW2::FeatureVector fv;
  voc.transform(features, v, fv, levelsup);
  int k = 0;
  for (DBoW2::BowVector::const_iterator it = v.begin(); it != v.end(); ++it, ++k) { words[k] = it->first; values[k] = it->second; }
  int a = 0, b = 0;
  off[0] = 0;
  for (DBoW2::FeatureVector::const_iterator it = fv.begin(); it != fv.end(); ++it, ++a) {
    nodes[a] = it->first;
    for (size_t j = 0; j < it->second.size(); j++) idx[b++] = it->second[j];
    off[a + 1] = b;
  }
  *nNodes = a; *nIdx = b;
  return k;
}

}  // extern "C"
