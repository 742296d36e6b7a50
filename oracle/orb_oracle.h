/* orb_oracle.h — C interface of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This library is a CPU restatement of the reference's ORB front end
 * (carry4985/ORB_SLAM2_MAP: src/ORBextractor.cc, src/ORBmatcher.cc, src/Frame.cc) with the OpenCV
 * primitives it calls (resize, GaussianBlur, FAST, fastAtan2) replaced by integer/float models
 * pinned bit-for-bit to cv2 4.13.0 (tests/golden/).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load it, and only as the checker or the
 * timed CPU baseline — the product (liborbx.so) never links or calls it.
 */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>
#include <stddef.h>
#include "../include/orbx.h"   /* POD types only: orbx_keypoint, orbx_params, orbm_* structs */

#ifdef __cplusplus
extern "C" {
#endif

/* --- OpenCV primitive models --- */
void oracle_resize_linear_u8(const uint8_t* src, int sw, int sh, size_t sstride, uint8_t* dst,
                             int dw, int dh, size_t dstride);
void oracle_gaussian_blur7_u8(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst,
                              size_t dstride);
void oracle_copy_make_border101(const uint8_t* src, int w, int h, size_t sstride, uint8_t* dst,
                                size_t dstride, int border);
/* cv::FAST(img, kps, th, nms) TYPE_9_16; out xyr[3*i] = x, y, response. returns count (may
 * exceed cap; only cap written). */
int oracle_fast9_16(const uint8_t* img, int w, int h, size_t stride, int th, int nms, float* xyr,
                    int cap);
float oracle_fast_atan2(float y, float x);
void oracle_sincos(float x, float* s, float* c, int use_libm);
int oracle_cv_round(double v);

/* --- ORBextractor --- */
typedef struct oracle_extractor oracle_extractor;
oracle_extractor* oracle_extractor_create(const orbx_params* p);
void oracle_extractor_destroy(oracle_extractor* e);
void oracle_extractor_tables(const oracle_extractor* e, float* scale, float* inv_scale,
                             float* sigma2, float* inv_sigma2, int32_t* features_per_level,
                             int32_t* umax16);
/* ORBextractor::operator(); returns keypoint count (or -1 if cap too small). */
int oracle_extract(oracle_extractor* e, const uint8_t* img, int w, int h, size_t stride,
                   orbx_keypoint* kps, uint8_t* desc, int cap, int use_libm_sincos);
/* state of the last oracle_extract call */
int oracle_level_size(const oracle_extractor* e, int level, int* w, int* h);
int oracle_get_level(const oracle_extractor* e, int level, int blurred, uint8_t* dst,
                     size_t dstride);
int oracle_get_candidates(const oracle_extractor* e, int level, float* xyr, int cap);
int oracle_get_level_keypoints(const oracle_extractor* e, int level, orbx_keypoint* kps, int cap);
/* DistributeOctTree alone: in xyr[3*n]; out index list (into the input) in result order. */
int oracle_distribute_octtree(const float* xyr, int n, int minX, int maxX, int minY, int maxY,
                              int N, int32_t* out_idx, int cap);

/* --- ORBmatcher / Frame --- */
int oracle_descriptor_distance(const uint8_t* a, const uint8_t* b);
/* Frame::GetFeaturesInArea on a flat frame; returns count, indices in reference order. */
int oracle_features_in_area(const orbm_frame* f, float x, float y, float r, int min_level,
                            int max_level, int32_t* out, int cap);
int oracle_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2,
                                     float* prev_matched, int32_t* matches12, int window_size,
                                     float nnratio, int check_orientation, uint64_t* ncompares);
int oracle_search_by_projection(int variant, const orbm_frame* frame,
                                const orbm_proj_query* queries, const uint8_t* query_desc, int nq,
                                const uint8_t* query_blocks, uint8_t* occupied, int32_t* assigned,
                                int32_t* best_idx, int32_t* best_dist, int th_dist, float nnratio,
                                int check_orientation);
int oracle_compute_stereo_matches(const orbm_stereo_input* in, float* u_right, float* depth,
                                  uint64_t* ncompares);
void oracle_knn2(const uint8_t* queries, int64_t nq, const uint8_t* database, int64_t ndb,
                 int32_t db_index_offset, uint16_t* out_dist, int32_t* out_idx, int nthreads);

/* multi-threaded driver for the CPU baseline: extract `nframes` frames with one extractor per
 * thread ("one frame per core"); returns total keypoints.  */
int64_t oracle_extract_many(const orbx_params* p, const uint8_t* imgs, size_t frame_stride, int w,
                            int h, size_t stride, int nframes, int nthreads);

/* config-2 CPU baseline: extraction of all frames + SearchForInitialization on consecutive pairs;
 * returns total keypoints. */
int64_t oracle_extract_match_many(const orbx_params* p, const uint8_t* imgs, size_t frame_stride,
                                  int w, int h, size_t stride, int nframes, int nthreads,
                                  int window_size, float nnratio, int check_orientation,
                                  int64_t* total_matches, uint64_t* ncompares);

#ifdef __cplusplus
}
#endif
#endif
