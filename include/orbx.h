/* orbx.h -- C ABI of liborbx.so: the B200-native (sm_100a) ORB front-end hot path.
 *
 * This is the drop-in boundary for the ONE data-parallel path of ORB-SLAM2 that this repository
 * replaces: ORBextractor::operator() and ORBmatcher's Hamming searches. The reference has no FFI;
 * its boundary is two C++ classes (R/include/ORBextractor.h:45-123, R/include/ORBmatcher.h:37-143).
 * The C++ mirror classes under orb-slam2-_b200/host/ keep those declarations and call only the
 * functions below. Plain pointers and sizes, no CUDA/torch types in any signature. There is no CPU
 * fallback: every entry point fails with ORBX_ERR_CUDA when no sm_100 device is usable.
 *
 * Conventions: return 0 (ORBX_OK) on success, <0 on error; orbx_last_error() returns a
 * thread-local message for the last failing call on this thread. Handles are thread-compatible:
 * one thread per handle at a time, different handles concurrently (the reference runs two
 * extractor instances on two std::threads per stereo frame, R/src/Frame.cpp:86-89).
 * "host" pointers may be pageable or pinned; "device" pointers must be device memory on the
 * handle's device. `stream` arguments are a cudaStream_t passed as void* (NULL = the handle's own
 * stream).
 */
#ifndef ORBX_H_
#define ORBX_H_
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ORBX_OK 0
#define ORBX_ERR_ARG (-1)      /* bad argument / unsupported size */
#define ORBX_ERR_CUDA (-2)     /* CUDA runtime/driver failure, or no usable device */
#define ORBX_ERR_CAPACITY (-3) /* caller buffer too small; *n holds the needed count */
#define ORBX_ERR_STATE (-4)    /* call order (e.g. orbx_get_level before any extract) */

#define ORBX_MAX_LEVELS 16
#define ORBX_DESC_BYTES 32

/* Same 28-byte layout as cv::KeyPoint {Point2f pt; float size, angle, response; int octave, class_id}
 * so the C++ wrapper can memcpy rows straight into std::vector<cv::KeyPoint>
 * (fields written by R/src/ORBextractor.cpp:905-915, 1178-1184; FAST sets class_id = -1). */
typedef struct orbx_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orbx_keypoint;

/* ORBextractor ctor arguments (R/src/ORBextractor.cpp:418-421; YAML keys ORBextractor.nFeatures,
 * .scaleFactor, .nLevels, .iniThFAST, .minThFAST read at R/src/Tracking.cpp:130-138) plus the
 * sizing the device buffers need. */
typedef struct orbx_params {
  int32_t nfeatures;
  float scale_factor;
  int32_t nlevels;
  int32_t ini_th_fast;
  int32_t min_th_fast;
  int32_t device;      /* CUDA device ordinal */
  int32_t max_width;   /* largest frame this handle will see */
  int32_t max_height;
  int32_t max_batch;   /* frames per orbx_extract_batch* call (>= 1) */
} orbx_params;

typedef struct orbx_extractor orbx_extractor;

const char* orbx_last_error(void);
/* "orbx <version> sm_100a" */
const char* orbx_version(void);
/* number of usable CUDA devices (0 when none: the library does not fall back to the CPU) */
int orbx_device_count(void);

/* ---- ORBextractor (replaces R/src/ORBextractor.cpp:418-502 ctor, :1120-1188 operator()) ---- */
int orbx_extractor_create(const orbx_params* params, orbx_extractor** out);
void orbx_extractor_destroy(orbx_extractor* h);

/* The six getters' tables (R/include/ORBextractor.h:66-86) + per-level quotas and umax;
 * any output pointer may be NULL. Arrays hold nlevels entries (umax16: 16). */
int orbx_extractor_tables(const orbx_extractor* h, float* scale_factors, float* inv_scale_factors,
                          float* level_sigma2, float* inv_level_sigma2, int32_t* features_per_level,
                          int32_t* umax16);

/* Shape of the host-buffer pipeline of orbx_extract_batch: frames per chunk and chunks in flight (1..8).
 * 0 keeps the library default (64 frames, 4 chunks; environment ORBX_CHUNK / ORBX_LANES). Results do not
 * depend on it. */
int orbx_extractor_set_pipeline(orbx_extractor* h, int32_t chunk_frames, int32_t lanes);

/* One frame, host buffers: the body of operator() (R/src/ORBextractor.cpp:1120-1188).
 * img: CV_8UC1 rows of `stride` bytes. kps/desc: capacity `cap` rows. *n = keypoints found.
 * Synchronous. ORBX_ERR_CAPACITY if cap is too small (nothing written but *n). */
int orbx_extract(orbx_extractor* h, const uint8_t* img, int32_t width, int32_t height, size_t stride,
                 orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n);

/* Page-locked host memory for frames and results (cudaHostAlloc on `device`'s context). Optional: every
 * host-buffer entry point accepts pageable memory too, pinned buffers make its copies asynchronous. */
int orbx_host_alloc(int32_t device, size_t bytes, void** out);
int orbx_host_free(void* p);

/* `count` same-sized frames, host buffers (frame f at imgs[f]); outputs are count x cap rows,
 * n[f] per frame. Used for image sequences / stereo pairs (BASELINE.json configs 2-4).
 * Frames are processed in chunks that rotate over several streams, so H2D, kernels and D2H overlap.
 * Pinned kps/desc arrays receive the device rows directly (entries past n[f] are then unspecified);
 * pageable ones are filled from a pinned staging buffer, rows [0, n[f]) only. */
int orbx_extract_batch(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width,
                       int32_t height, size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap,
                       int32_t* n);

/* Streaming form of orbx_extract_batch for continuous sequences: returns once the chunks are enqueued, so that the
 * next submission's host->device copies overlap this one's kernels and device->host copies (a synchronous call pays
 * one chunk of pipeline fill and one of drain). kps / desc / n must be page-locked (orbx_host_alloc) and, like the
 * frames, stay untouched until orbx_extract_batch_wait(h, *ticket) returns; up to 8 submissions may be in flight,
 * and they complete in order. Results and error conditions are those of orbx_extract_batch, reported by the wait. */
int orbx_extract_batch_submit(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width,
                              int32_t height, size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap,
                              int32_t* n, int32_t* ticket);
int orbx_extract_batch_wait(orbx_extractor* h, int32_t ticket);

/* Stereo sequences, streaming (BASELINE.json config 3; the per-frame work of the stereo constructor
 * R/src/Frame.cpp:68-129: two extractions :86-89, then ComputeStereoMatches :101). imgs are L R L R ... (count even);
 * kps / desc / n as in orbx_extract_batch_submit. Every chunk of the submission also runs the row matcher of
 * orbx_stereo_match on its pairs, against the pyramids the chunk has just built: u_right / depth are [count/2][cap]
 * (row i of pair p belongs to keypoint i of frame 2p; -1 where unmatched; rows >= n[2p] unspecified), n_matched is
 * [count/2]; all three page-locked. Wait with orbx_extract_batch_wait. Remember that the reference runs
 * ComputeStereoMatches with mb == 0 (see orbx_stereo_match). */
int orbx_stereo_batch_submit(orbx_extractor* h, const uint8_t* const* imgs, int32_t count, int32_t width, int32_t height,
                             size_t stride, orbx_keypoint* kps, uint8_t* desc, int32_t cap, int32_t* n, float mbf, float mb,
                             float* u_right, float* depth, int32_t* n_matched, int32_t* ticket);

/* Same, frames and outputs resident in device memory; asynchronous on `stream`.
 * Frame f starts at d_imgs + f*frame_stride. d_n[f] may exceed cap: rows beyond cap are dropped.
 * d_n[f] == -1: the FAST candidate list of frame f overflowed its capacity (corner density above 1 per 8 pixels on
 * some level; which candidates were dropped depends on atomics) -- the frame's rows are meaningless. The host-buffer
 * entry points report the same condition as ORBX_ERR_CAPACITY. */
int orbx_extract_batch_device(orbx_extractor* h, const uint8_t* d_imgs, int32_t count, int32_t width,
                              int32_t height, size_t row_stride, size_t frame_stride,
                              orbx_keypoint* d_kps, uint8_t* d_desc, int32_t cap, int32_t* d_n,
                              void* stream);

/* mvImagePyramid (public member, R/include/ORBextractor.h:88; read by Frame::ComputeStereoMatches,
 * R/src/Frame.cpp:558,675-695): size of level `level` for the last extracted shape, and a lazy
 * device->host copy of level `level` of frame `frame` of the last batch. */
int orbx_level_size(const orbx_extractor* h, int32_t level, int32_t* width, int32_t* height);
int orbx_get_level(orbx_extractor* h, int32_t frame, int32_t level, uint8_t* dst, size_t dst_stride);
/* Levels 1 .. nlevels-1 of frame `frame` of the last batch in one go (what the C++ ORBextractor mirror fills
 * mvImagePyramid from after every operator() call): level l lands at dst + offsets[l] with row pitch pitches[l] (the
 * device planes as they lie, padding included; offsets[0] / pitches[0] are not written: level 0 is the caller's
 * image). The copies are queued back to back and waited for once; with page-locked dst (orbx_host_alloc) they are
 * DMA transfers with no staging. *bytes = size dst must have; dst == NULL only reports offsets, pitches and *bytes.
 * offsets / pitches: ORBX_MAX_LEVELS entries. */
int orbx_get_pyramid(orbx_extractor* h, int32_t frame, uint8_t* dst, size_t dst_bytes, size_t* offsets,
                     int32_t* pitches, size_t* bytes);

/* Introspection for the parity tests (not used by the C++ mirror classes). */
int orbx_get_blurred_level(orbx_extractor* h, int32_t frame, int32_t level, uint8_t* dst, size_t dst_stride);
/* FAST candidates handed to the octree, (x, y, score) triples relative to minBorder; unordered. */
int orbx_get_level_candidates(orbx_extractor* h, int32_t frame, int32_t level, int32_t* xys, int32_t cap,
                              int32_t* n);
/* number of kernels launched by this handle since creation (bench.py's gpu_launches) */
int64_t orbx_launch_count(const orbx_extractor* h);
/* device time in ms of the stages of the last extract call (single or batch) when profiling is enabled
 * with orbx_set_profiling(h, 1): [0]=pyramid [1]=fast [2]=octree [3]=blur [4]=orient+describe [5]=total */
int orbx_set_profiling(orbx_extractor* h, int32_t enable);
int orbx_last_stage_ms(orbx_extractor* h, float* ms6);

/* ---- stand-alone primitives on host buffers (parity tests of each kernel; synchronous) ---- */
int orbx_resize_u8(int32_t device, const uint8_t* src, int32_t sw, int32_t sh, size_t sstride,
                   uint8_t* dst, int32_t dw, int32_t dh, size_t dstride);
int orbx_blur7_u8(int32_t device, const uint8_t* src, int32_t w, int32_t h, size_t sstride, uint8_t* dst,
                  size_t dstride);
/* cv::fastAtan2 and the glibc-exact sincosf on arrays */
int orbx_fast_atan2(int32_t device, const float* y, const float* x, float* out_deg, int64_t n);
int orbx_sincosf(int32_t device, const float* a, float* out_sin, float* out_cos, int64_t n);
/* sin/cos of every float whose bit pattern lies in [lo_bits, hi_bits) (exhaustive device-vs-libm
 * comparison); out arrays hold hi_bits-lo_bits floats */
int orbx_sincosf_sweep(int32_t device, uint32_t lo_bits, uint32_t hi_bits, float* out_sin, float* out_cos);
/* DistributeOctTree (R/src/ORBextractor.cpp:571-817) on arbitrary candidates: (x,y,score) triples
 * relative to (min_x,min_y). cell_w/cell_h = the FAST cell size the candidates came from; it defines
 * cv::FAST's emission order (cell row, cell col, y, x), the tie-break between equal responses in a
 * node. out_xys receives the kept triples in the reference's list order. */
int orbx_distribute(int32_t device, const int32_t* xys, int32_t n, int32_t min_x, int32_t max_x,
                    int32_t min_y, int32_t max_y, int32_t n_features, int32_t cell_w, int32_t cell_h,
                    int32_t* out_xys, int32_t cap, int32_t* n_out);

/* ---- ORBmatcher (replaces R/src/ORBmatcher.cpp DescriptorDistance :1901-1917 and the search loops) ---- */
typedef struct orbx_matcher orbx_matcher;
int orbx_matcher_create(int32_t device, orbx_matcher** out);
void orbx_matcher_destroy(orbx_matcher* h);
int64_t orbx_matcher_launch_count(const orbx_matcher* h);

/* DescriptorDistance for n pairs of 32-byte rows (a[i] vs b[i]) */
int orbx_descriptor_distance(orbx_matcher* h, const uint8_t* a, const uint8_t* b, int32_t* dist, int64_t n);

/* Brute-force top-2 + ratio test (inner loop of R/src/ORBmatcher.cpp:266-300, :526-556 over the
 * whole database in index order): idx/d1 = first-minimum index and distance, d2 = second-smallest
 * distance of the multiset, accepted = d1 <= th && d1 < ratio*d2. Host buffers, synchronous. */
int orbx_match_bf2(orbx_matcher* h, const uint8_t* q, int32_t nq, const uint8_t* db, int64_t ndb,
                   float ratio, int32_t th, int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted);
/* Device-resident, asynchronous; idx_base is added to every index (database shards). */
int orbx_match_bf2_device(orbx_matcher* h, const uint8_t* d_q, int32_t nq, const uint8_t* d_db, int64_t ndb,
                          int64_t idx_base, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2, void* stream);
/* Merge `nshards` per-shard results laid out [shard][nq] (e.g. after an all-gather) with the
 * sequential rule of :282-291 (equal minima in two shards => d2 = d1; lowest index wins) and apply
 * threshold + ratio. Device-resident, asynchronous. */
int orbx_match_merge_device(orbx_matcher* h, const int32_t* d_idx, const int32_t* d_d1, const int32_t* d_d2,
                            int32_t nshards, int32_t nq, float ratio, int32_t th, int32_t* d_out_idx,
                            int32_t* d_out_d1, int32_t* d_out_d2, uint8_t* d_accepted, void* stream);

/* Allocation-free brute-force top-2 + ratio test. NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it). Same results as orbx_match_bf2_device followed by orbx_match_merge_device (idx / d1 / d2 /
 * accepted of orbx_match_bf2), all pointers device memory on `device`, asynchronous on `stream` (a cudaStream_t;
 * NULL = the default stream). No handle: the per-chunk partial results live in ONE file-scope device array of
 * 12 MiB that every GPU loading the library holds until the process exits; the call itself performs no cudaMalloc,
 * cudaFree or other allocation. Limits: nq <= ORBX_NOALLOC_MAX_QUERIES, and (database chunks) x nq <= 2^20 partial
 * triples -- beyond them the call fails with ORBX_ERR_ARG and a message (use orbx_match_bf2_device). All calls on a
 * device share that array: calls that overlap on the GPU (different streams, or graphs replayed concurrently) race
 * and must be serialised by the caller. Nothing in the array needs clearing between calls (every slot a call reads
 * it has written first). The first call per device resolves the array's address (cudaGetSymbolAddress): make it
 * outside stream capture. d_accepted may be NULL. */
#define ORBX_NOALLOC_MAX_QUERIES 16384
int orbx_match_bf2_noalloc(int32_t device, const uint8_t* d_q, int32_t nq, const uint8_t* d_db, int64_t ndb, float ratio,
                           int32_t th, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2, uint8_t* d_accepted, void* stream);

/* Windowed search on the Frame grid: the data-parallel half of SearchForInitialization / SearchByProjection
 * (R/src/ORBmatcher.cpp:499-617, :63-163, :1564-1718). The target frame is (kps_un, desc, n) with the image
 * bounds mnMinX..mnMaxY that define its 64x48 grid (R/src/Frame.cpp:168-169, :244-260). For query i the
 * candidates are exactly Frame::GetFeaturesInArea(q_xy[2i], q_xy[2i+1], q_r[i], q_min_level[i],
 * q_max_level[i]) (R/src/Frame.cpp:387-440) in the reference's order: cand_idx[cand_off[i] .. cand_off[i+1])
 * with cand_dist = DescriptorDistance(q_desc[i], desc[cand_idx]). Host buffers, synchronous.
 * ORBX_ERR_CAPACITY when cand_cap is too small (*n_cand = needed, cand_off is valid). The order-dependent
 * resolve runs on the host (orb-slam2-_b200/host/orbx_search.h). */
int orbx_match_window(orbx_matcher* h, const orbx_keypoint* kps_un, const uint8_t* desc, int32_t n, float min_x,
                      float max_x, float min_y, float max_y, const float* q_xy, const float* q_r,
                      const int32_t* q_min_level, const int32_t* q_max_level, const uint8_t* q_desc, int32_t nq,
                      int32_t* cand_off, int32_t* cand_idx, int32_t* cand_dist, int32_t cand_cap, int32_t* n_cand);

/* Frame::ComputeStereoMatches (R/src/Frame.cpp:551-770) for the last frames extracted by `left` and `right`
 * (single-chunk calls; their pyramids are still on the device): row-band Hamming match, 11x11 SAD refinement
 * over 11 offsets, sub-pixel parabola, depth = mbf/disparity, median-SAD filter. kps/desc are the two
 * extractors' outputs (host). u_right/depth: nl floats, -1 where unmatched; *n_matched = surviving matches. */
int orbx_stereo_match(orbx_matcher* h, orbx_extractor* left, orbx_extractor* right, int32_t frame_left,
                      int32_t frame_right, const orbx_keypoint* kps_l, const uint8_t* desc_l, int32_t nl,
                      const orbx_keypoint* kps_r, const uint8_t* desc_r, int32_t nr, float mbf, float mb,
                      float* u_right, float* depth, int32_t* n_matched);

/* The same for `pairs` stereo pairs of ONE device-resident batch (BASELINE config 2: image sequences of stereo
 * pairs): `ex` processed the frames with orbx_extract_batch_device; pair p is left frame left_frame0 + p*frame_step
 * and right frame right_frame0 + p*frame_step of that batch (0, 1, 2 = interleaved L R L R ...). d_kps / d_desc /
 * d_n are that call's device outputs (cap rows per frame). Outputs, device: d_u_right / d_depth [pairs][cap]
 * (rows >= n of the left frame are not written), d_n_matched [pairs]. Asynchronous on `stream` (NULL: the
 * matcher's own stream, after waiting for the extractor's). Remember that the reference runs this with mb == 0. */
int orbx_stereo_match_batch_device(orbx_matcher* h, orbx_extractor* ex, int32_t pairs, int32_t left_frame0,
                                   int32_t right_frame0, int32_t frame_step, const orbx_keypoint* d_kps,
                                   const uint8_t* d_desc, const int32_t* d_n, int32_t cap, float mbf, float mb,
                                   float* d_u_right, float* d_depth, int32_t* d_n_matched, void* stream);

/* Node-restricted distance blocks: the data-parallel half of ORBmatcher::SearchByBoW (R/src/ORBmatcher.cpp:220-365
 * key frame -> frame, :632-772 key frame -> key frame) and SearchForTriangulation (:785-983). Block k is one
 * vocabulary node both DBoW2::FeatureVectors hold (the merge-join of the two maps, :229-231 / :328-340, is the
 * caller's): it pairs rows a_idx[a_off[k] .. a_off[k+1]) of desc_a with rows b_idx[b_off[k] .. b_off[k+1]) of
 * desc_b. dist[dist_off[k] + ia*nb_k + ib] = DescriptorDistance(desc_a[a_idx[a_off[k]+ia]], desc_b[b_idx[b_off[k]+ib]]).
 * dist_off (nblocks+1 entries) is written by the call; *n_dist = dist_off[nblocks]. Host buffers, synchronous.
 * ORBX_ERR_CAPACITY when dist_cap < *n_dist (dist_off and *n_dist are valid). The order-dependent resolve runs on
 * the host (orb-slam2-_b200/host/orbx_search.h). */
int orbx_match_blocks(orbx_matcher* h, const uint8_t* desc_a, int32_t na, const uint8_t* desc_b, int32_t nb,
                      const int32_t* a_off, const int32_t* a_idx, const int32_t* b_off, const int32_t* b_idx,
                      int32_t nblocks, int64_t* dist_off, uint16_t* dist, int64_t dist_cap, int64_t* n_dist);

/* MapPoint::ComputeDistinctiveDescriptors (R/src/MapPoint.cpp:306-385) for `npoints` map points in one launch
 * (LocalMapping calls it per created / fused point, R/src/LocalMapping.cpp). Point p observes the descriptor rows
 * desc[obs_off[p] .. obs_off[p+1]) -- gathered by the caller in the iteration order of its mObservations map,
 * bad key frames skipped (:326-333). best[p] = position within the group of the first row whose sorted distance
 * row has the least median vDists[0.5*(N-1)] (:357-375), best_median[p] that median; -1 / -1 for a point without
 * rows (the reference returns without touching mDescriptor, :335-336). Host buffers, synchronous. */
int orbx_distinctive(orbx_matcher* h, const uint8_t* desc, const int32_t* obs_off, int32_t npoints, int32_t* best,
                     int32_t* best_median);
/* Device-resident, asynchronous on `stream` (NULL: the matcher's own stream). */
int orbx_distinctive_device(orbx_matcher* h, const uint8_t* d_desc, const int32_t* d_obs_off, int32_t npoints,
                            int32_t* d_best, int32_t* d_best_median, void* stream);

/* ---- frame epilogue on the device ----
 * cv::undistortPoints(src, dst, K, distCoef, cv::Mat(), K) as Frame::UndistortKeyPoints (R/src/Frame.cpp:470-504)
 * and ComputeImageBounds (:506-540) call it: n (x, y) float pairs, K4 = fx fy cx cy, dist = k1 k2 p1 p2 [k3 ...]
 * (n_dist <= 12, OpenCV order). Bit-exact with OpenCV 4.13 (double arithmetic, five iterations). Host buffers. */
int orbx_undistort_points(int32_t device, const float* xy, int64_t n, const float* K4, const float* dist, int32_t n_dist,
                          float* out_xy);

/* A frame whose keypoints, descriptors and 64x48 grid stay on the device between extraction and the windowed
 * searches (Frame::Frame's epilogue: UndistortKeyPoints, AssignFeaturesToGrid R/src/Frame.cpp:244-260). Lives on the
 * matcher's device and stream; one frame handle serves one frame at a time and is refilled with orbx_frame_set. */
typedef struct orbx_frame orbx_frame;
int orbx_frame_create(orbx_matcher* m, orbx_frame** out);
void orbx_frame_destroy(orbx_frame* f);
/* kps / desc: n rows, host (on_device = 0) or device pointers (on_device = 1, e.g. the outputs of
 * orbx_extract_batch_device). dist[0] == 0 or n_dist == 0: mvKeysUn = mvKeys (:472-476). bounds4 = mnMinX mnMaxX mnMinY
 * mnMaxY (computed once per camera, :506-540; orbx_undistort_points serves the four corners). Asynchronous. */
int orbx_frame_set(orbx_frame* f, const orbx_keypoint* kps, const uint8_t* desc, int32_t n, int32_t on_device,
                   const float* K4, const float* dist, int32_t n_dist, const float* bounds4);
/* mvKeysUn back on the host (pose optimisation reads it): n rows */
int orbx_frame_get_keys_un(orbx_frame* f, orbx_keypoint* kps_un);
/* orbx_match_window against the resident frame: no keypoint / descriptor upload, no grid build, one packed upload of
 * the queries. Same outputs and error behaviour. */
int orbx_frame_match_window(orbx_frame* f, const float* q_xy, const float* q_r, const int32_t* q_min_level,
                            const int32_t* q_max_level, const uint8_t* q_desc, int32_t nq, int32_t* cand_off,
                            int32_t* cand_idx, int32_t* cand_dist, int32_t cand_cap, int32_t* n_cand);

/* ---- vocabulary descent ----
 * The tree of a DBoW2::TemplatedVocabulary<cv::Mat, FORB> resident on the device: node 0 is the root, parent[i] < i,
 * the children of a node keep their id order (the order loadFromTextFile appends them in,
 * R/Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1362-1436), desc = nnodes x 32 bytes, depth_L = m_L. */
typedef struct orbx_vocabulary orbx_vocabulary;
int orbx_vocabulary_create(orbx_matcher* m, int32_t nnodes, const int32_t* parent, const uint8_t* desc, int32_t depth_L,
                           orbx_vocabulary** out);
void orbx_vocabulary_destroy(orbx_vocabulary* v);
/* transform(feature, word_id, weight, nid, levelsup) (:1242-1296) for n descriptors: leaf_node[i] = the node the
 * descent ends in (its word id and weight are the caller's m_nodes[leaf].word_id / .weight), level_node[i] = the node
 * passed at level depth_L - levelsup (0 when that level is <= 0; -1 when the leaf lies above it, where the reference
 * leaves *nid unset). Host buffers, synchronous. BowVector / FeatureVector assembly: host/orbx_bow.h. */
int orbx_vocabulary_descend(orbx_vocabulary* v, const uint8_t* desc, int32_t n, int32_t levelsup, int32_t* leaf_node,
                            int32_t* level_node);

/* ---- multi-GPU ----
 * The path shards two ways over the GPUs of one box: extraction by frames (contiguous, pair-aligned blocks of a
 * sequence per GPU -- a stereo pair stays on one GPU like the two extraction threads of R/src/Frame.cpp:86-89 -- no
 * collective) and brute-force Hamming by database rows (queries replicated, per-GPU top-2, ONE ncclAllGather of
 * 12 bytes per query and rank, merge with the sequential rule of R/src/ORBmatcher.cpp:282-291 + ratio test on every
 * rank). A group is either every rank of ONE process (orbx_shard_group_create: the calls below run one host thread
 * per device) or ONE rank of a multi-process job (orbx_shard_group_create_rank; rank 0 obtains the id with
 * orbx_shard_unique_id and the launcher broadcasts it). NCCL is bound at run time (libnccl.so.2); a single-rank
 * group never touches it. */
#define ORBX_SHARD_ID_BYTES 128
typedef struct orbx_shard_group orbx_shard_group;
int orbx_shard_unique_id(uint8_t* id /* ORBX_SHARD_ID_BYTES */);
int orbx_shard_group_create(const int32_t* devices, int32_t ndev, orbx_shard_group** out);
int orbx_shard_group_create_rank(int32_t device, int32_t rank, int32_t nranks, const uint8_t* id, orbx_shard_group** out);
void orbx_shard_group_destroy(orbx_shard_group* g);
/* any output may be NULL; nccl_version = ncclGetVersion() of the library bound, 0 when NCCL was never needed */
int orbx_shard_group_info(const orbx_shard_group* g, int32_t* nranks, int32_t* nlocal, int32_t* first_rank,
                          int32_t* nccl_version);
int64_t orbx_shard_launch_count(const orbx_shard_group* g);
/* rank's block [lo, hi) of `total` units cut into nranks contiguous blocks of whole groups (group = 2: stereo pairs) */
int orbx_shard_block_range(int64_t total, int32_t rank, int32_t nranks, int32_t group, int64_t* lo, int64_t* hi);

/* The database of config 5: ndb rows of 32 bytes on the host; every local rank uploads its own row range
 * (orbx_shard_block_range(ndb, rank, nranks, 1)). Synchronous. */
int orbx_shard_set_database(orbx_shard_group* g, const uint8_t* db, int64_t ndb);
/* ... or adopts rows [row_lo, row_hi) that already lie in its device memory (not copied, not owned) */
int orbx_shard_set_database_device(orbx_shard_group* g, int32_t local_index, const uint8_t* d_rows, int64_t row_lo,
                                   int64_t row_hi, int64_t ndb);
/* orbx_match_bf2 over the sharded database: host queries in, host results out (global row indices), identical on
 * every rank and to orbx_match_bf2 on the whole database. Synchronous; in a multi-process job every rank calls it. */
int orbx_shard_match_bf2(orbx_shard_group* g, const uint8_t* q, int32_t nq, float ratio, int32_t th, int32_t* idx,
                         int32_t* d1, int32_t* d2, uint8_t* accepted);
/* One rank's part, device-resident and asynchronous on `stream` (NULL: the rank's own): top-2 on its shard,
 * ncclAllGather, merge + ratio test. For one-rank-per-process groups (every rank calls it once per step). */
int orbx_shard_match_bf2_device(orbx_shard_group* g, int32_t local_index, const uint8_t* d_q, int32_t nq, float ratio,
                                int32_t th, int32_t* d_idx, int32_t* d_d1, int32_t* d_d2, uint8_t* d_accepted, void* stream);

/* A sequence of `count` same-sized frames (imgs[f], page-locked or pageable; results page-locked, count x cap rows as
 * in orbx_extract_batch_submit) extracted by the group: rank r streams block orbx_shard_block_range(count, r, nranks,
 * group) through its own extractor (created from *params with .device replaced; steps of params->max_batch frames,
 * two in flight) and the results land in sequence order. group = 2 with u_right / depth / n_matched non-NULL: a
 * stereo sequence L R L R ... (orbx_stereo_batch_submit; outputs [count/2][cap], [count/2]). In a multi-process job
 * every process fills only its own block. Synchronous. */
int orbx_shard_extract_sequence(orbx_shard_group* g, const orbx_params* params, const uint8_t* const* imgs, int64_t count,
                                int32_t group, int32_t width, int32_t height, size_t stride, orbx_keypoint* kps,
                                uint8_t* desc, int32_t cap, int32_t* n, float mbf, float mb, float* u_right, float* depth,
                                int32_t* n_matched);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H_ */
