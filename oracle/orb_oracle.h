/* TEST INFRASTRUCTURE (oracle) -- C API of liborb_oracle.so.
 *
 * CPU restatement of the reference's ORB front-end hot path (SURVEY.md section 8a rows E0-E9,
 * M0-M6, S1). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library, and only as the checker or the timed CPU baseline -- never as part
 * of the product path (the product is liborbx.so, CUDA only, no CPU fallback).
 *
 * Parity status: the OpenCV primitives are pinned against cv2 4.13.0 (tests/test_oracle_prims.py,
 * tests/golden/); the extractor control flow is pinned against the reference's own ORBextractor.cpp
 * compiled against a cv shim (oracle/_ref, tests/test_oracle_vs_ref.py) on tie-insensitive inputs.
 * DistributeOctTree's node-size ties are broken by heap address in the reference
 * (R/src/ORBextractor.cpp:633,736); this oracle uses the canonical (size, creation sequence) key.
 */
#ifndef ORB_ORACLE_H_
#define ORB_ORACLE_H_
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same 28-byte layout as cv::KeyPoint {Point2f pt; float size, angle, response; int octave, class_id}. */
typedef struct orb_oracle_keypoint {
  float x, y, size, angle, response;
  int32_t octave, class_id;
} orb_oracle_keypoint;

/* ---- primitives (cv_prims.cpp) ---- */
void orb_oracle_resize(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orb_oracle_blur7(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
/* returns the number of corners found (may exceed cap; only cap are written), xys = x,y,score triples */
int orb_oracle_fast(const uint8_t* img, int w, int h, int stride, int threshold, int nms, int32_t* xys, int cap);
float orb_oracle_fast_atan2(float y, float x);
void orb_oracle_sincosf_port(float a, float* s, float* c);
void orb_oracle_sincosf_libm(float a, float* s, float* c);
int orb_oracle_cv_round(double v);
/* exhaustive check of sincosf_port against the host libm over float bit patterns [lo, hi); returns mismatches */
int64_t orb_oracle_sincosf_sweep(uint32_t lo_bits, uint32_t hi_bits, int nthreads, uint32_t* first_bad);

/* ---- extractor (orb_extractor_oracle.cpp): ORBextractor ctor + operator() ---- */
typedef struct orb_oracle_extractor orb_oracle_extractor;
orb_oracle_extractor* orb_oracle_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
void orb_oracle_extractor_destroy(orb_oracle_extractor*);
/* tie_mode: 0 = canonical (size, creation seq) [default]; 1 = (size, reverse creation seq) -- the
 * other extreme the reference's pointer order can produce; used to measure tie sensitivity. */
void orb_oracle_extractor_set_tie_mode(orb_oracle_extractor*, int tie_mode);
/* returns number of keypoints (<= cap or -needed if cap too small); desc is n x 32 bytes */
int orb_oracle_extract(orb_oracle_extractor*, const uint8_t* img, int w, int h, int stride,
                       orb_oracle_keypoint* kps, uint8_t* desc, int cap);
/* tables / introspection after create() or extract() */
void orb_oracle_extractor_tables(orb_oracle_extractor*, float* scale, float* inv_scale, float* sigma2,
                                 float* inv_sigma2, int32_t* feats_per_level, int32_t* umax16);
void orb_oracle_level_size(orb_oracle_extractor*, int level, int* w, int* h);
void orb_oracle_get_level(orb_oracle_extractor*, int level, uint8_t* dst, int dstride);
void orb_oracle_get_blurred_level(orb_oracle_extractor*, int level, uint8_t* dst, int dstride);
/* FAST candidates handed to DistributeOctTree at `level` (x,y relative to minBorder, score), in
 * the reference's cell-major raster order; returns count */
int orb_oracle_level_candidates(orb_oracle_extractor*, int level, int32_t* xys, int cap);
int orb_oracle_level_kept(orb_oracle_extractor*, int level);
/* number of phase-2 sort ties that decided a split/no-split in the last extract (tie sensitivity) */
int orb_oracle_level_tie_sensitive(orb_oracle_extractor*, int level);
/* stand-alone DistributeOctTree on arbitrary candidates (x,y,score triples); returns kept count,
 * out_idx = indices into the candidate array in output (list) order */
int orb_oracle_distribute(const int32_t* xys, int n, int minX, int maxX, int minY, int maxY, int N,
                          int tie_mode, int32_t* out_idx, int cap);

/* ---- matcher (orb_matcher_oracle.cpp) ---- */
int orb_oracle_descriptor_distance(const uint8_t* a, const uint8_t* b);
/* brute-force top-2 + ratio (M1): per query best idx/dist, second dist, accepted flag */
void orb_oracle_match_bf2(const uint8_t* q, int nq, const uint8_t* db, int64_t ndb, float ratio, int th,
                          int32_t* idx, int32_t* d1, int32_t* d2, uint8_t* accepted, int nthreads);

/* Frame grid (Frame.cpp:244-260, 387-452): build + query */
typedef struct orb_oracle_grid orb_oracle_grid;
orb_oracle_grid* orb_oracle_grid_create(const orb_oracle_keypoint* kps_un, int n, float min_x, float max_x,
                                        float min_y, float max_y);
void orb_oracle_grid_destroy(orb_oracle_grid*);
int orb_oracle_grid_query(orb_oracle_grid*, float x, float y, float r, int min_level, int max_level,
                          int32_t* out_idx, int cap);
/* SearchForInitialization (ORBmatcher.cpp:499-617) */
int orb_oracle_search_for_initialization(const orb_oracle_keypoint* kps1, const uint8_t* desc1, int n1,
                                         const orb_oracle_keypoint* kps2, const uint8_t* desc2, int n2,
                                         float min_x, float max_x, float min_y, float max_y,
                                         float* prev_matched_xy /* n1 x 2, in/out */, int32_t* matches12,
                                         int window, float nnratio, int check_orientation);
/* SearchByProjection(Frame&, vector<MapPoint*>, th) (ORBmatcher.cpp:63-163) with the host-side
 * geometric prologue already applied: per map point (u, v, radius, predicted level, descriptor, ur or <=0) */
int orb_oracle_search_by_projection_mp(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n,
                                       const float* uright /* n, or NULL */, const uint8_t* has_mp /* n */,
                                       float min_x, float max_x, float min_y, float max_y,
                                       const float* mp_uv, const float* mp_r, const int32_t* mp_level,
                                       const float* mp_ur, const uint8_t* mp_desc, int nmp,
                                       const float* scale_factors, float nnratio, int32_t* mp_match /* nmp */,
                                       const uint8_t* mp_has_obs /* nmp: Observations()>0, or NULL = all */);
/* SearchByProjection(Frame& cur, const Frame& last, th, mono) (ORBmatcher.cpp:1564-1718) after projection */
int orb_oracle_search_by_projection_frame(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n,
                                          const float* uright, const uint8_t* has_mp,
                                          float min_x, float max_x, float min_y, float max_y,
                                          const float* q_uv, const float* q_r, const int32_t* q_min_level,
                                          const int32_t* q_max_level, const float* q_ur, const float* q_angle,
                                          const uint8_t* q_desc, int nq, int th_dist, int check_orientation,
                                          int32_t* cur_match /* n: query index or -1 */,
                                          const uint8_t* q_has_obs /* nq: Observations()>0, or NULL = all */);
/* Frame::ComputeStereoMatches (Frame.cpp:551-770) */
int orb_oracle_stereo_matches(const orb_oracle_keypoint* kps_l, const uint8_t* desc_l, int nl,
                              const orb_oracle_keypoint* kps_r, const uint8_t* desc_r, int nr,
                              const uint8_t* const* pyr_l, const uint8_t* const* pyr_r,
                              const int32_t* lvl_w, const int32_t* lvl_h, const int32_t* lvl_stride, int nlevels,
                              const float* scale_factors, const float* inv_scale_factors,
                              float mbf, float max_d, float* u_right, float* depth);

/* ---- SURVEY.md 8f rows 1-2 (orb_matcher_oracle.cpp, second half). A DBoW2::FeatureVector is passed flattened:
 * node ids ascending, CSR offsets, feature indices in insertion order. ---- */
void orb_oracle_block_distances(const uint8_t* desc_a, const uint8_t* desc_b, const int32_t* a_off, const int32_t* a_idx,
                                const int32_t* b_off, const int32_t* b_idx, int nblocks, int64_t* dist_off, uint16_t* dist);
/* SearchByBoW(KeyFrame*, Frame&, ...) ORBmatcher.cpp:220-365 */
int orb_oracle_search_by_bow_kf_frame(const uint8_t* desc_kf, const float* angle_kf, const uint8_t* kf_usable,
                                      const int32_t* kf_nodes, const int32_t* kf_off, const int32_t* kf_idx, int kf_nn,
                                      const uint8_t* desc_f, const float* angle_f, int n_f, const int32_t* f_nodes,
                                      const int32_t* f_off, const int32_t* f_idx, int f_nn, float nnratio,
                                      int check_orientation, int32_t* match_f);
/* SearchByBoW(KeyFrame*, KeyFrame*, ...) ORBmatcher.cpp:632-772 */
int orb_oracle_search_by_bow_kf_kf(const uint8_t* desc1, const float* angle1, const uint8_t* usable1, int n1,
                                   const int32_t* nodes1, const int32_t* off1, const int32_t* idx1, int nn1,
                                   const uint8_t* desc2, const float* angle2, const uint8_t* usable2, int n2,
                                   const int32_t* nodes2, const int32_t* off2, const int32_t* idx2, int nn2, float nnratio,
                                   int check_orientation, int32_t* matches12);
/* SearchForTriangulation ORBmatcher.cpp:785-983 (+ CheckDistEpipolarLine :175-203) */
int orb_oracle_search_for_triangulation(const orb_oracle_keypoint* kps1, const uint8_t* desc1, const uint8_t* has_mp1,
                                        const float* uright1, int n1, const int32_t* nodes1, const int32_t* off1,
                                        const int32_t* idx1, int nn1, const orb_oracle_keypoint* kps2, const uint8_t* desc2,
                                        const uint8_t* has_mp2, const float* uright2, int n2, const int32_t* nodes2,
                                        const int32_t* off2, const int32_t* idx2, int nn2, const float* F12,
                                        float ex, float ey, const float* scale_factors2, const float* level_sigma2_2,
                                        int only_stereo, int check_orientation, int32_t* matches12);
/* MapPoint::ComputeDistinctiveDescriptors MapPoint.cpp:306-385, many points */
void orb_oracle_distinctive(const uint8_t* desc, const int32_t* obs_off, int npoints, int32_t* best, int32_t* best_median);

/* windowed best-match core of Fuse x2 / SearchByProjection(KeyFrame*, Scw) / SearchBySim3 (ORBmatcher.cpp:995-1296,
 * :370-497, :1305-1547) after their projection prologues; see the definition */
void orb_oracle_search_best_in_window(const orb_oracle_keypoint* kps_un, const uint8_t* desc, int n, const float* uright,
                                      float min_x, float max_x, float min_y, float max_y, const float* inv_level_sigma2,
                                      const float* q_uv, const float* q_ur, const float* q_radius, const int32_t* q_level,
                                      const uint8_t* q_desc, int nq, uint8_t* taken, int th, int32_t* best_idx,
                                      int32_t* best_dist);

/* cv::undistortPoints(src, dst, K, dist, noArray(), K) as OpenCV 4.13 computes it (cv_prims.cpp) */
void orb_oracle_undistort_points(const float* xy, long long n, const float* K4, const float* dist, int ndist, float* out);

#ifdef __cplusplus
}
#endif
#endif /* ORB_ORACLE_H_ */
