// TEST INFRASTRUCTURE (oracle) -- exhaustive sweeps used to pin float restatements.
#include <atomic>
#include <cmath>
#include <cstring>
#include <thread>
#include <vector>

#include "cv_prims.h"
#include "orb_oracle.h"

extern "C" int64_t orb_oracle_sincosf_sweep(uint32_t lo_bits, uint32_t hi_bits, int nthreads, uint32_t* first_bad) {
  if (nthreads < 1) nthreads = 1;
  std::atomic<int64_t> bad{0};
  std::atomic<uint32_t> first{0xffffffffu};
  std::vector<std::thread> th;
  const uint64_t total = (uint64_t)hi_bits - lo_bits;
  for (int t = 0; t < nthreads; t++) {
    th.emplace_back([&, t]() {
      uint64_t b = lo_bits + total * t / nthreads, e = lo_bits + total * (t + 1) / nthreads;
      int64_t local = 0;
      for (uint64_t u = b; u < e; u++) {
        uint32_t bits = (uint32_t)u;
        float a;
        std::memcpy(&a, &bits, 4);
        float s0, c0, s1, c1;
        ::sincosf(a, &s0, &c0);
        orb_oracle::sincosf_port(a, &s1, &c1);
        if (std::memcmp(&s0, &s1, 4) != 0 || std::memcmp(&c0, &c1, 4) != 0) {
          local++;
          uint32_t cur = first.load();
          while (bits < cur && !first.compare_exchange_weak(cur, bits)) {}
        }
      }
      bad += local;
    });
  }
  for (auto& x : th) x.join();
  if (first_bad) *first_bad = first.load();
  return bad.load();
}
