// C++ face of the multi-GPU part of liborbx.so (include/orbx.h, orbx_shard_*; implementation: csrc/orbx_shard.cu).
//
// What a host program that used to loop over a sequence with ORB_SLAM2::ORBextractor, or over a descriptor database
// with ORBmatcher::DescriptorDistance, uses to spread that work over the GPUs of one box:
//   * SequenceExtractor: contiguous, pair-aligned blocks of frames per GPU, results in sequence order, no collective
//     (a stereo pair stays on one GPU, like the two extraction threads of R/src/Frame.cpp:86-89);
//   * ShardedDatabase: row ranges of the database per GPU, queries replicated, per-GPU top-2, ncclAllGather + merge
//     with the sequential rule of R/src/ORBmatcher.cpp:282-291.
// Header-only and C++11 like the rest of host/; errors throw std::runtime_error with the library's message.
#ifndef ORBX_SHARD_H_
#define ORBX_SHARD_H_

#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>

#include "orbx.h"

namespace orbx_host {

class ShardGroup {
 public:
  // every rank of one process: one host thread per device inside the calls
  explicit ShardGroup(const std::vector<int>& devices) : g_(NULL) {
    std::vector<int32_t> d(devices.begin(), devices.end());
    Check(orbx_shard_group_create(d.empty() ? NULL : &d[0], (int32_t)d.size(), &g_));
  }
  // one rank of a multi-process job (id: UniqueId() of rank 0, broadcast by the launcher)
  ShardGroup(int device, int rank, int nranks, const std::vector<uint8_t>& id) : g_(NULL) {
    Check(orbx_shard_group_create_rank(device, rank, nranks, id.empty() ? NULL : &id[0], &g_));
  }
  ~ShardGroup() { orbx_shard_group_destroy(g_); }

  static std::vector<uint8_t> UniqueId() {
    std::vector<uint8_t> id(ORBX_SHARD_ID_BYTES);
    Check(orbx_shard_unique_id(&id[0]));
    return id;
  }
  static void BlockRange(int64_t total, int rank, int nranks, int group, int64_t* lo, int64_t* hi) {
    Check(orbx_shard_block_range(total, rank, nranks, group, lo, hi));
  }

  int Ranks() const { int32_t n = 0; Check(orbx_shard_group_info(g_, &n, NULL, NULL, NULL)); return n; }
  int LocalRanks() const { int32_t n = 0; Check(orbx_shard_group_info(g_, NULL, &n, NULL, NULL)); return n; }
  orbx_shard_group* get() const { return g_; }

  // ---- config 5: brute-force top-2 + ratio over a sharded database ----
  void SetDatabase(const uint8_t* rows, int64_t n) { Check(orbx_shard_set_database(g_, rows, n)); }
  struct Matches {
    std::vector<int32_t> idx, d1, d2;      // global row of the best match, best and second-best distance
    std::vector<uint8_t> accepted;         // d1 <= th && d1 < ratio * d2
  };
  Matches MatchBf2(const uint8_t* queries, int nq, float ratio, int th) const {
    Matches m;
    m.idx.resize(nq); m.d1.resize(nq); m.d2.resize(nq); m.accepted.resize(nq);
    if (nq > 0) Check(orbx_shard_match_bf2(g_, queries, nq, ratio, th, &m.idx[0], &m.d1[0], &m.d2[0], &m.accepted[0]));
    return m;
  }

  // ---- configs 1-4: a sequence of same-sized frames; result arrays page-locked (orbx_host_alloc), count x cap rows ----
  void ExtractSequence(const orbx_params& params, const uint8_t* const* imgs, int64_t count, int width, int height,
                       size_t stride, orbx_keypoint* kps, uint8_t* desc, int cap, int32_t* n) const {
    Check(orbx_shard_extract_sequence(g_, &params, imgs, count, 1, width, height, stride, kps, desc, cap, n, 0.f, 0.f, NULL,
                                      NULL, NULL));
  }
  // L R L R ...: both extractions + Frame::ComputeStereoMatches per pair (mb == 0 reproduces the reference's
  // constructor, see orbx_stereo_match); uRight / depth: [count/2][cap], nMatched: [count/2]
  void ExtractStereoSequence(const orbx_params& params, const uint8_t* const* imgs, int64_t count, int width, int height,
                             size_t stride, orbx_keypoint* kps, uint8_t* desc, int cap, int32_t* n, float mbf, float mb,
                             float* uRight, float* depth, int32_t* nMatched) const {
    Check(orbx_shard_extract_sequence(g_, &params, imgs, count, 2, width, height, stride, kps, desc, cap, n, mbf, mb, uRight,
                                      depth, nMatched));
  }

 private:
  ShardGroup(const ShardGroup&);
  ShardGroup& operator=(const ShardGroup&);
  static void Check(int rc) {
    if (rc != ORBX_OK) throw std::runtime_error(std::string("orbx: ") + orbx_last_error());
  }
  orbx_shard_group* g_;
};

}  // namespace orbx_host

#endif  // ORBX_SHARD_H_
