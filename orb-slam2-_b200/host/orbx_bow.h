// Host half of the vocabulary transform over liborbx.so: what Frame::ComputeBoW and
// KeyFrame::ComputeBoW obtain from mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4)
// (R/src/Frame.cpp:262-274, R/src/KeyFrame.cpp:66-75).
//
// The per-descriptor descent of the tree runs on the device (orbx_vocabulary_descend). What remains here is the
// loader of the text format (R/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1362-1436) and the assembly of the two
// maps in feature order -- BowVector::addWeight / addIfNotExist / normalize (BowVector.cpp:34-84),
// FeatureVector::addFeature (FeatureVector.cpp:31-45) -- with the reference's control flow (:1151-1228). The maps
// have the element types of DBoW2::BowVector / DBoW2::FeatureVector, so a replacement ComputeBoW swaps them in.
#ifndef ORBX_BOW_H_
#define ORBX_BOW_H_

#include <map>
#include <vector>

#include "orbx.h"

namespace orbx_host {

struct Vocabulary {
  int k = 0, L = 0;
  int scoring = 0;     // DBoW2::ScoringType: L1_NORM, L2_NORM, CHI_SQUARE, KL, BHATTACHARYYA, DOT_PRODUCT
  int weighting = 0;   // DBoW2::WeightingType: TF_IDF, TF, IDF, BINARY
  std::vector<int> parent;            // per node; node 0 is the root
  std::vector<unsigned char> desc;    // nodes x 32
  std::vector<double> weight;         // m_nodes[i].weight
  std::vector<int> wordId;            // m_nodes[i].word_id, -1 for inner nodes
  orbx_vocabulary* device = nullptr;
  int words = 0;

  ~Vocabulary();
  // loadFromTextFile + upload of the tree; false when the file is not a vocabulary (like the reference)
  bool loadFromTextFile(orbx_matcher* m, const char* filename);
  // transform(features, v, fv, levelsup): features = n rows of 32 bytes (mDescriptors)
  void transform(const unsigned char* features, int n, std::map<unsigned, double>* v,
                 std::map<unsigned, std::vector<unsigned> >* fv, int levelsup) const;
};

}  // namespace orbx_host

#endif
