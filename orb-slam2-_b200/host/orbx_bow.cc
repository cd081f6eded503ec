// See orbx_bow.h.
#include "orbx_bow.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <string>

namespace orbx_host {

Vocabulary::~Vocabulary() { if (device) orbx_vocabulary_destroy(device); }

bool Vocabulary::loadFromTextFile(orbx_matcher* m, const char* filename) {
  FILE* fp = std::fopen(filename, "rb");
  if (!fp) return false;
  std::string text;
  char buf[1 << 16];
  size_t got;
  while ((got = std::fread(buf, 1, sizeof buf, fp)) > 0) text.append(buf, got);
  std::fclose(fp);
  // the reference reads with getline until eof: every '\n'-terminated piece is a line, and so is what follows the
  // last '\n' (an empty piece after a trailing newline becomes one more node there, and here)
  size_t pos = 0;
  auto next_line = [&](std::string* line) {
    if (pos == std::string::npos) return false;
    const size_t e = text.find('\n', pos);
    if (e == std::string::npos) { *line = text.substr(pos); pos = std::string::npos; }
    else { *line = text.substr(pos, e - pos); pos = e + 1; }
    return true;
  };
  std::string line;
  if (!next_line(&line)) return false;
  char* q = const_cast<char*>(line.c_str());
  k = (int)std::strtol(q, &q, 10);
  L = (int)std::strtol(q, &q, 10);
  const int n1 = (int)std::strtol(q, &q, 10), n2 = (int)std::strtol(q, &q, 10);
  if (k < 0 || k > 20 || L < 1 || L > 10 || n1 < 0 || n1 > 5 || n2 < 0 || n2 > 3) {   // :1383-1388
    std::fprintf(stderr, "Vocabulary loading failure: This is not a correct text file!\n");
    return false;
  }
  scoring = n1; weighting = n2;
  parent.assign(1, 0); desc.assign(32, 0); weight.assign(1, 0.0); wordId.assign(1, -1);
  words = 0;
  while (next_line(&line)) {
    q = const_cast<char*>(line.c_str());
    const int pid = (int)std::strtol(q, &q, 10);
    const int nIsLeaf = (int)std::strtol(q, &q, 10);
    if (pid < 0 || pid >= (int)parent.size()) { std::fprintf(stderr, "Vocabulary: node %zu names parent %d\n", parent.size(), pid); return false; }
    parent.push_back(pid);
    for (int i = 0; i < 32; i++) desc.push_back((unsigned char)std::strtol(q, &q, 10));   // FORB::fromString
    weight.push_back(std::strtod(q, &q));
    wordId.push_back(nIsLeaf > 0 ? words++ : -1);
  }
  if (device) { orbx_vocabulary_destroy(device); device = nullptr; }
  if (orbx_vocabulary_create(m, (int)parent.size(), &parent[0], &desc[0], L, &device) != ORBX_OK) {
    std::fprintf(stderr, "orbx_host: orbx_vocabulary_create: %s\n", orbx_last_error());
    std::abort();   // no CPU fallback
  }
  return true;
}

void Vocabulary::transform(const unsigned char* features, int n, std::map<unsigned, double>* v,
                           std::map<unsigned, std::vector<unsigned> >* fv, int levelsup) const {
  v->clear();
  fv->clear();
  if (parent.size() <= 1) return;   // empty()
  std::vector<int> leaf((size_t)(n > 0 ? n : 1)), level((size_t)(n > 0 ? n : 1));
  if (orbx_vocabulary_descend(device, features, n, levelsup, &leaf[0], &level[0]) != ORBX_OK) {
    std::fprintf(stderr, "orbx_host: orbx_vocabulary_descend: %s\n", orbx_last_error());
    std::abort();
  }
  const bool accumulate = weighting == 0 || weighting == 1;   // TF_IDF, TF (:1166)
  const bool must = scoring != 5;                              // every scoring class but DotProductScoring normalises
  const bool l2 = scoring == 1;
  for (int i = 0; i < n; i++) {
    const double w = weight[leaf[i]];
    const unsigned id = (unsigned)wordId[leaf[i]];
    if (!(w > 0)) continue;
    if (level[i] < 0) {   // a word above level L - levelsup: the reference reads an unset NodeId here
      std::fprintf(stderr, "orbx_host: vocabulary has a leaf above level %d\n", L - levelsup);
      std::abort();
    }
    if (accumulate) {
      std::map<unsigned, double>::iterator it = v->lower_bound(id);           // BowVector::addWeight
      if (it != v->end() && !(v->key_comp()(id, it->first))) it->second += w;
      else v->insert(it, std::make_pair(id, w));
    } else {
      std::map<unsigned, double>::iterator it = v->lower_bound(id);           // BowVector::addIfNotExist
      if (it == v->end() || v->key_comp()(id, it->first)) v->insert(it, std::make_pair(id, w));
    }
    (*fv)[(unsigned)level[i]].push_back((unsigned)i);                          // FeatureVector::addFeature
  }
  if (accumulate && !v->empty() && !must) {
    const double nd = (double)v->size();
    for (std::map<unsigned, double>::iterator it = v->begin(); it != v->end(); ++it) it->second /= nd;
  }
  if (must) {   // BowVector::normalize
    double norm = 0.0;
    if (!l2) { for (std::map<unsigned, double>::iterator it = v->begin(); it != v->end(); ++it) norm += std::fabs(it->second); }
    else {
      // the reference's -O3 build contracts norm += x*x into a fused multiply-add
      for (std::map<unsigned, double>::iterator it = v->begin(); it != v->end(); ++it) norm = std::fma(it->second, it->second, norm);
      norm = std::sqrt(norm);
    }
    if (norm > 0.0) for (std::map<unsigned, double>::iterator it = v->begin(); it != v->end(); ++it) it->second /= norm;
  }
}

}  // namespace orbx_host
