// One orbx_matcher handle (stream + device scratch) per host thread, shared by the drop-in translation units
// (ORBmatcher.cc, FrameStereo.cc): ORBmatcher objects are stack locals used from the Tracking, LocalMapping and
// LoopClosing threads at the same time (threading), and a handle serves one thread at a time.
#ifndef ORBX_THREAD_H_
#define ORBX_THREAD_H_

#include <cstdio>
#include <cstdlib>

#include "orbx.h"

namespace orbx_host {

struct ThreadHandle {
  orbx_matcher* h;
  ThreadHandle() : h(nullptr) {}
  ~ThreadHandle() { if (h) orbx_matcher_destroy(h); }   // a thread that ends gives its stream and scratch back
};

// Device: ORBX_DEVICE from the environment, else 0. There is no CPU path: a missing device aborts with the
// library's message (the reference's interface has no error channel).
inline orbx_matcher* ThreadMatcher() {
  static thread_local ThreadHandle t;
  if (!t.h) {
    const char* dev = std::getenv("ORBX_DEVICE");
    if (orbx_matcher_create(dev ? std::atoi(dev) : 0, &t.h) != ORBX_OK) {
      std::fprintf(stderr, "orbx: %s\n", orbx_last_error());
      std::abort();
    }
  }
  return t.h;
}

}  // namespace orbx_host

#endif
