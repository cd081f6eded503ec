// TEST INFRASTRUCTURE (oracle) -- CPU restatement of the ORBmatcher / Frame hot-path pieces.
// Follows R/src/ORBmatcher.cpp (DescriptorDistance :1901-1917, top-2 + ratio loops :266-300,
// SearchForInitialization :499-617, SearchByProjection :63-163 / :1564-1718, ComputeThreeMaxima
// :1854-1895) and R/src/Frame.cpp (grid :244-260, :387-452; ComputeStereoMatches :551-770).
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "orb_oracle.h"

namespace orb_oracle {

// R/src/ORBmatcher.cpp:1901-1917 -- 8 x 32-bit SWAR population count of a XOR b.
static inline int descriptor_distance(const uint8_t* a, const uint8_t* b) {
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

}  // namespace orb_oracle

using namespace orb_oracle;

extern "C" {

int orb_oracle_descriptor_distance(const uint8_t* a, const uint8_t* b) { return descriptor_distance(a, b); }

// Brute-force top-2 + ratio: the inner-loop pattern of R/src/ORBmatcher.cpp:266-300 / :526-556
// applied to every database row in index order.
void orb_oracle_match_bf2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, float ratio, int th,
                          int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted, int nthreads) {
  if (nthreads < 1) nthreads = 1;
  auto work = [&](int q0, int q1) {
    for (int i = q0; i < q1; i++) {
      const uint8_t* dq = q + (size_t)i * 32;
      int best = INT_MAX, best2 = INT_MAX, bidx = -1;
      for (int64_t j = 0; j < ndb; j++) {
        const int dist = descriptor_distance(dq, db + (size_t)j * 32);
        if (dist < best) { best2 = best; best = dist; bidx = (int)j; }
        else if (dist < best2) { best2 = dist; }
      }
      idx[i] = bidx; d1[i] = best; d2[i] = best2;
      if (accepted) accepted[i] = (best <= th && (float)best < ratio * (float)best2) ? 1 : 0;
    }
  };
  std::vector<std::thread> ts;
  for (int t = 0; t < nthreads; t++) ts.emplace_back(work, (int)((int64_t)nq * t / nthreads), (int)((int64_t)nq * (t + 1) / nthreads));
  for (auto& t : ts) t.join();
}

}  // extern "C"
