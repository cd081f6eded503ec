// TEST INFRASTRUCTURE (oracle) -- CPU restatement of the vocabulary transform of the reference's vendored DBoW2
// (SURVEY.md 8f row 4): TemplatedVocabulary::loadFromTextFile (R/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1362-1436),
// transform(features, BowVector&, FeatureVector&, levelsup) (:1151-1228) over transform(feature, word, weight, nid,
// levelsup) (:1242-1296), FORB::distance (FORB.cpp:74-92), BowVector::addWeight / addIfNotExist / normalize
// (BowVector.cpp:34-84), the scoring classes' mustNormalize (ScoringObject.h:76-91).
// PINNED: tests/test_oracle_vs_ref_voc.py runs the reference's own DBoW2 (oracle/_ref/libref_voc.so) on the same
// vocabulary files and descriptors and requires identical words, values (bit for bit) and feature vectors.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <cstdio>
#include <cstdlib>
#include <map>
#include <string>
#include <vector>

namespace {

struct Node {
  int parent = 0;
  std::vector<int> children;
  uint8_t desc[32] = {};
  double weight = 0;
  int wordId = -1;
  bool isLeaf() const { return children.empty(); }
};

struct Voc {
  int k = 0, L = 0, scoring = 0, weighting = 0;
  std::vector<Node> nodes;
  int nWords = 0;
};

int distance(const uint8_t* a, const uint8_t* b) {   // FORB.cpp:74-92
  int dist = 0;
  for (int i = 0; i < 8; i++) {
    uint32_t x, y;
    std::memcpy(&x, a + 4 * i, 4);
    std::memcpy(&y, b + 4 * i, 4);
    uint32_t v = x ^ y;
    v = v - ((v >> 1) & 0x55555555);
    v = (v & 0x33333333) + ((v >> 2) & 0x33333333);
    dist += (((v + (v >> 4)) & 0xF0F0F0F) * 0x1010101) >> 24;
  }
  return dist;
}

}  // namespace

extern "C" {

// Parsing is done with the C library (strtol / strtod read the same values as the reference's stringstream
// extractions on a well-formed file): liborb_oracle.so carries a statically linked libstdc++ whose iostream locale
// facets do not survive next to a second copy of libstdc++ in the same process.
void* orb_oracle_voc_load(const char* path) {
  FILE* fp = std::fopen(path, "rb");
  if (!fp) return nullptr;
  std::string text;
  char buf[1 << 16];
  size_t got;
  while ((got = std::fread(buf, 1, sizeof buf, fp)) > 0) text.append(buf, got);
  std::fclose(fp);
  // std::getline semantics: lines end at '\n'; a trailing newline leaves one more, empty, line before eof
  std::vector<std::string> lines;
  size_t pos = 0;
  for (;;) {
    const size_t e = text.find('\n', pos);
    if (e == std::string::npos) { lines.push_back(text.substr(pos)); break; }
    lines.push_back(text.substr(pos, e - pos));
    pos = e + 1;
  }
  Voc* v = new Voc();
  int n1 = -1, n2 = -1;
  {
    char* q = const_cast<char*>(lines[0].c_str());
    v->k = (int)std::strtol(q, &q, 10); v->L = (int)std::strtol(q, &q, 10);
    n1 = (int)std::strtol(q, &q, 10); n2 = (int)std::strtol(q, &q, 10);
  }
  if (v->k < 0 || v->k > 20 || v->L < 1 || v->L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) { delete v; return nullptr; }
  v->scoring = n1; v->weighting = n2;
  v->nodes.resize(1);
  for (size_t li = 1; li < lines.size(); li++) {   // like the reference's while(!f.eof()) loop: every remaining line is a node
    char* q = const_cast<char*>(lines[li].c_str());
    const int nid = (int)v->nodes.size();
    v->nodes.resize(nid + 1);
    const int pid = (int)std::strtol(q, &q, 10);          // a failed extraction leaves 0
    v->nodes[nid].parent = pid;
    v->nodes[pid].children.push_back(nid);
    const int nIsLeaf = (int)std::strtol(q, &q, 10);
    for (int iD = 0; iD < 32; iD++) v->nodes[nid].desc[iD] = (uint8_t)std::strtol(q, &q, 10);   // FORB::fromString
    v->nodes[nid].weight = std::strtod(q, &q);
    if (nIsLeaf > 0) v->nodes[nid].wordId = v->nWords++;
  }
  return v;
}
void orb_oracle_voc_destroy(void* h) { delete static_cast<Voc*>(h); }
int orb_oracle_voc_nodes(void* h) { return (int)static_cast<Voc*>(h)->nodes.size(); }
int orb_oracle_voc_words(void* h) { return static_cast<Voc*>(h)->nWords; }
/* flattened tree for the product's constructor: parent (nnodes), desc (nnodes x 32), weight, word id (-1: inner node) */
void orb_oracle_voc_export(void* h, int32_t* parent, uint8_t* desc, double* weight, int32_t* word, int32_t* kLsw4) {
  const Voc& v = *static_cast<Voc*>(h);
  for (size_t i = 0; i < v.nodes.size(); i++) {
    parent[i] = v.nodes[i].parent;
    std::memcpy(desc + i * 32, v.nodes[i].desc, 32);
    weight[i] = v.nodes[i].weight;
    word[i] = v.nodes[i].wordId;
  }
  kLsw4[0] = v.k; kLsw4[1] = v.L; kLsw4[2] = v.scoring; kLsw4[3] = v.weighting;
}

/* same flattened outputs as ref_voc_transform (oracle/ref_shim/ref_voc_capi.cpp) */
int orb_oracle_voc_transform(void* h, const uint8_t* desc, int n, int levelsup, uint32_t* words, double* values,
                             uint32_t* nodes, int32_t* off, uint32_t* idx, int* nNodes, int* nIdx) {
  const Voc& voc = *static_cast<Voc*>(h);
  std::map<unsigned, double> v;
  std::map<unsigned, std::vector<unsigned>> fv;
  *nNodes = *nIdx = 0;
  off[0] = 0;
  if (voc.nodes.size() <= 1) return 0;   // empty()
  const bool sum = voc.weighting == 0 || voc.weighting == 1;            // TF_IDF, TF: addWeight; IDF, BINARY: addIfNotExist
  const bool must = voc.scoring != 5;                                   // DotProductScoring does not normalise
  const bool l2 = voc.scoring == 1;
  const int nid_level = voc.L - levelsup;
  for (int i = 0; i < n; i++) {
    const uint8_t* feature = desc + (size_t)i * 32;
    unsigned nid = 0;   // the reference leaves it unset when a leaf lies above nid_level; vocabularies here have none
    int final_id = 0, current_level = 0;
    do {
      ++current_level;
      const std::vector<int>& ch = voc.nodes[final_id].children;
      final_id = ch[0];
      double best_d = distance(feature, voc.nodes[final_id].desc);
      for (size_t c = 1; c < ch.size(); c++) {
        const double d = distance(feature, voc.nodes[ch[c]].desc);
        if (d < best_d) { best_d = d; final_id = ch[c]; }
      }
      if (current_level == nid_level) nid = (unsigned)final_id;
    } while (!voc.nodes[final_id].isLeaf());
    const double w = voc.nodes[final_id].weight;
    const unsigned id = (unsigned)voc.nodes[final_id].wordId;
    if (w > 0) {
      if (sum) v[id] += w;                       // addWeight: insert or accumulate
      else v.insert(std::make_pair(id, w));      // addIfNotExist
      fv[nid].push_back((unsigned)i);
    }
  }
  if (sum && !v.empty() && !must) {
    const double nd = (double)v.size();
    for (auto& e : v) e.second /= nd;
  }
  if (must) {
    double norm = 0.0;
    if (!l2) { for (auto& e : v) norm += std::fabs(e.second); }
    else { for (auto& e : v) norm = std::fma(e.second, e.second, norm); norm = std::sqrt(norm); }
    if (norm > 0.0) for (auto& e : v) e.second /= norm;
  }
  int k = 0;
  for (auto& e : v) { words[k] = e.first; values[k] = e.second; k++; }
  int a = 0, b = 0;
  for (auto& e : fv) {
    nodes[a] = e.first;
    for (unsigned j : e.second) idx[b++] = j;
    off[++a] = b;
  }
  *nNodes = a; *nIdx = b;
  return k;
}

}  // extern "C"
