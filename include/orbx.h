/* orbx.h — C ABI of the B200-native ORB front end (liborbx.so).
 *
 * This is the drop-in boundary for ONE hot path of carry4985/ORB_SLAM2_MAP: the per-frame ORB
 * front end.  The reference has no FFI of its own (its boundary is the C++ class API), so every
 * entry point below names the reference function it replaces; the C++ classes in
 * orb_slam2_map_b200/host/ re-create the reference's signatures on top of these calls.
 *
 * Conventions
 *  - plain pointers and sizes only; no C++/torch types; never throws, never aborts.
 *  - every function returns ORBX_OK (0) or a negative orbx_status; orbx_last_error() gives the
 *    message of the calling thread's last failure.
 *  - there is NO CPU fallback: without a CUDA device every compute entry point fails with
 *    ORBX_E_CUDA.
 *  - a handle owns one CUDA stream (or borrows the caller's), a device workspace and a result
 *    store; it must be used by one thread at a time (same rule as one ORBextractor object,
 *    reference src/Frame.cc:78-79 uses two objects for two threads).
 *  - "slot" = index of a frame inside the handle's device-resident result store.
 *
 * Environment variables the library reads (none is needed; results are identical under all of them):
 *   ORBX_NO_GRAPH=1         orbx_extract / orbx_frame_rgbd launch their kernels plainly instead of replaying the captured graph
 *   ORBX_LAT_NOFORK=1       single-frame calls: one stream instead of one forked branch per pyramid level
 *   ORBX_LAT_RS=<rows>, ORBX_LAT_BLUR=<groups>, ORBX_LAT_OT=<128|256|512>
 *                           single-frame calls: resize strip height, blur launch groups, quadtree CTA size (tuning sweeps)
 *   ORBX_RS_TARGET=<rows>   batch calls: resize strip height (tuning sweep; disables the graph)
 *   ORBX_OCTREE_KEYCAP=<n>  force DistributeOctTree onto its global-memory key path above n candidates (test hook)
 *   ORBX_RGBD_SLICES=<n>    orbx_frame_rgbd: row slices the pageable colour image is staged in (default 4)
 *   ORBX_NO_SCRATCH_POOL=1  stateless orbm_* calls: cudaMalloc / cudaFree per call instead of the per-thread block cache
 *   ORBX_HOST_CHUNK=<n>     batch calls from host memory: frames per pipelined chunk (default 256, at most the handle's chunk_frames)
 *   ORBX_HOST_THREADS=<n>   batch calls from PAGEABLE host memory: host threads that copy to / from the page-locked pools (default min(8, cores))
 *   ORBX_NO_HOST_POOL=1     batch calls from PAGEABLE host memory: let the driver stage the copies (no pools)
 *   ORBX_PIPE_TRACE=1       batch calls from host memory: chunk timeline (copy / kernels / result copy) to stderr
 *   ORBX_LAT_TRACE=1        single-frame and stateless matcher calls: phase times (averaged over 50-200 calls) to stderr
 */
#ifndef ORBX_H
#define ORBX_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum orbx_status {
    ORBX_OK = 0,
    ORBX_E_INVALID = -1,   /* bad argument */
    ORBX_E_CUDA = -2,      /* CUDA runtime failure (incl. no device) */
    ORBX_E_CAPACITY = -3,  /* caller buffer / handle capacity too small */
    ORBX_E_STATE = -4      /* call sequence error (e.g. slot not extracted yet) */
} orbx_status;

/* Same 28-byte layout as cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id),
 * which is what ORBextractor::operator() fills (reference src/ORBextractor.cc:1043-1105). */
typedef struct orbx_keypoint {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orbx_keypoint;

/* ORBextractor constructor arguments (reference include/ORBextractor.h:51-52). */
typedef struct orbx_params {
    int32_t nfeatures;
    float scale_factor;
    int32_t nlevels;
    int32_t ini_th_fast;
    int32_t min_th_fast;
} orbx_params;

typedef struct orbx_handle orbx_handle;

const char* orbx_last_error(void);
/* number of CUDA devices visible; <0 on failure (lets a caller fail loudly up front). */
int orbx_device_count(void);

/* ---- extractor: replaces ORBextractor::ORBextractor / ~ORBextractor ------------------------
 * max_width/max_height: largest image the handle will see; chunk_frames: frames in flight per
 * kernel wave (workspace size); result_slots: frames whose keypoints/descriptors stay resident
 * on the device; stream: a cudaStream_t to borrow, or NULL to create one. */
int orbx_create(const orbx_params* params, int max_width, int max_height, int chunk_frames,
                int result_slots, int device, void* stream, orbx_handle** out);
int orbx_destroy(orbx_handle* h);
int orbx_sync(orbx_handle* h);

/* ---- measurement hooks (bench.py): per-stage device time via CUDA events recorded on the
 * handle's stream around each stage, and the number of kernels launched per stage. */
enum { ORBX_STAGE_RESIZE = 0, ORBX_STAGE_BLUR, ORBX_STAGE_FAST, ORBX_STAGE_OCTREE, ORBX_STAGE_DESCRIBE,
       ORBX_STAGE_GRID, ORBX_STAGE_SEARCH_INIT, ORBX_STAGE_STEREO, ORBX_NSTAGES };
int orbx_set_profiling(orbx_handle* h, int enable);
/* accumulated since the previous call (synchronises); either pointer may be NULL */
int orbx_get_profile(orbx_handle* h, double* stage_ms, int64_t* stage_launches);

/* scale tables computed exactly as the reference ctor does (src/ORBextractor.cc:410-446);
 * arrays of nlevels entries, any pointer may be NULL.  Replaces GetScaleFactors & friends. */
int orbx_get_tables(const orbx_handle* h, float* scale, float* inv_scale, float* sigma2,
                    float* inv_sigma2, int32_t* features_per_level);
/* upper bound of keypoints per frame (sum over levels of max(N_l+3, 4*nIni)) for the shapes seen so far.
 * The value set at creation covers the handle's maximum shape; a smaller image with a wider aspect ratio
 * (more quadtree roots per level, src/ORBextractor.cc:543-546) can need more: see orbx_reserve_shape. */
int orbx_keypoint_capacity(const orbx_handle* h);
/* Prepare the handle for width x height images (geometry, workspace, tensor maps) ahead of the first
 * extraction of that shape and return the per-frame keypoint capacity output buffers must have (>= 0), or
 * a negative orbx_status.  When the shape needs more than the current capacity the result store grows and
 * the slots stored so far are discarded.  Every extraction entry point does the same on a shape change;
 * callers that size their buffers once call this first (the host ORBextractor does). */
int orbx_reserve_shape(orbx_handle* h, int width, int height);

/* ---- ORBextractor::operator() (src/ORBextractor.cc:1043), single frame, host buffers --------
 * image: 8UC1 host pointer.  kps/desc: host buffers with room for `capacity` keypoints
 * (desc = capacity*32 bytes).  *n receives the count.  width*height==0 -> *n untouched (the
 * reference returns silently on an empty image).
 * This is the latency path (one call per tracked frame, Frame::ExtractORB src/Frame.cc:247-253): the image
 * crosses in one copy through pinned staging, the kernels of the frame are replayed from a CUDA graph captured
 * on the first call of an image shape (handles created on a caller-supplied stream launch them one by one),
 * and count, keypoints and descriptors return behind a single stream synchronisation. */
int orbx_extract(orbx_handle* h, const uint8_t* image, int width, int height, size_t row_stride,
                 orbx_keypoint* kps, uint8_t* desc, int capacity, int* n);

/* Batched form of the same operator: nframes images of one shape, `images` either a host
 * pointer (pinned or pageable) or a device pointer (images_on_device!=0); frame i starts at
 * images + i*frame_stride.  Results land in slots first_slot .. first_slot+nframes-1 and stay
 * on the device; asynchronous on the handle's stream when images are on the device. */
int orbx_extract_batch(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                       size_t row_stride, int width, int height, int nframes,
                       int images_on_device, int first_slot);

/* Same, with the results also delivered to HOST buffers (kps[nframes][capacity],
 * desc[nframes][capacity][32], counts[nframes]; any may be NULL): the host->device copy of chunk
 * i+1 and the device->host copy of chunk i-1 overlap the kernels of chunk i.  PAGE-LOCKED buffers (cudaHostAlloc, or
 * cudaHostRegister once on buffers that are reused) are copied directly: 159 k frames/s at 640x480.  PAGEABLE buffers go
 * through page-locked pools of the handle (a copy thread fills the next chunk's pool while this one is queued, results are
 * handed over one chunk later): 82 k frames/s (25 k when the driver stages the copies itself, ORBX_NO_HOST_POOL=1).  Records beyond counts[f] of a frame
 * are unspecified.  Synchronous when any output pointer is given. */
int orbx_extract_batch_to_host(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                               size_t row_stride, int width, int height, int nframes,
                               int images_on_device, int first_slot, orbx_keypoint* kps,
                               uint8_t* desc, int32_t* counts, int capacity);

/* orbx_extract_batch_to_host with ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) of every
 * consecutive pair (frame f, frame f+1) folded into the same chunk pipeline (BASELINE.json configs[1]): the pairs of a
 * chunk are matched right after its extraction, so the matcher and its result copy overlap the next chunk's input
 * copy too.  matches12: host [nframes-1][capacity] (vnMatches12 of pair f at row f), nmatches: host [nframes-1];
 * the result is what orbx_search_init_pairs(first_slot, nframes-1, ...) returns after a plain batch extraction. */
int orbx_extract_match_to_host(orbx_handle* h, const uint8_t* images, size_t frame_stride,
                               size_t row_stride, int width, int height, int nframes,
                               int images_on_device, int first_slot, orbx_keypoint* kps, uint8_t* desc,
                               int32_t* counts, int capacity, int window_size, float nnratio,
                               int check_orientation, int32_t* matches12, int32_t* nmatches);

/* Copy results of slots [first_slot, first_slot+nframes) to host: kps[nframes][capacity],
 * desc[nframes][capacity][32], counts[nframes].  Synchronises. Any of kps/desc may be NULL. */
int orbx_download(orbx_handle* h, int first_slot, int nframes, orbx_keypoint* kps, uint8_t* desc,
                  int32_t* counts, int capacity);

/* ORBextractor::mvImagePyramid[level] of the LAST single-frame/batch chunk (src/ORBextractor.h:85,
 * read by Frame::ComputeStereoMatches).  frame = index inside the last chunk.  blurred=0 gives the
 * raw level, 1 the 7x7 Gaussian-blurred working copy (src/ORBextractor.cc:1085-1086).
 * border>0 adds a BORDER_REFLECT_101 margin of that many pixels (the reference keeps 19). */
int orbx_get_level(orbx_handle* h, int frame, int level, int blurred, int border, uint8_t* dst,
                   size_t dst_stride, int* width, int* height);

/* test/diagnostic taps (device -> host), same chunk addressing as orbx_get_level:
 * FAST candidates handed to DistributeOctTree for one level, in the reference's order
 * (src/ORBextractor.cc:789-829): xyr[3*i] = x, y (window-relative), response. */
int orbx_get_candidates(orbx_handle* h, int frame, int level, float* xyr, int capacity, int* n);

/* ---- Frame grid + ORBmatcher on device-resident slots ----------------------------------------
 * Consecutive-frame ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520) for pairs
 * (first_slot+i, first_slot+i+1), i<npairs, with vbPrevMatched = F1's keypoint positions
 * (Tracking.cc:862-866) and the frame grid of Frame::AssignFeaturesToGrid (src/Frame.cc:230) on
 * bounds [0,width]x[0,height].  matches12: host [npairs][capacity] (may be NULL), nmatches: host
 * [npairs] (may be NULL); with both NULL the call is asynchronous and results stay on device.
 * ncompares (host, may be NULL): number of DescriptorDistance evaluations performed. */
int orbx_search_init_pairs(orbx_handle* h, int first_slot, int npairs, int window_size,
                           float nnratio, int check_orientation, int32_t* matches12,
                           int32_t* nmatches, int capacity, uint64_t* ncompares);

/* Stereo: extract left+right (two ORBextractor objects in the reference, src/Frame.cc:78-81) and
 * run Frame::ComputeStereoMatches (src/Frame.cc:466-638) per pair while the pyramids are still on
 * the device.  Left frame i -> slot first_slot+2*i, right -> first_slot+2*i+1.
 * u_right/depth: host [npairs][capacity] (mvuRight / mvDepth), may be NULL (async). */
int orbx_extract_stereo_batch(orbx_handle* h, const uint8_t* left, const uint8_t* right,
                              size_t frame_stride, size_t row_stride, int width, int height,
                              int npairs, int images_on_device, int first_slot, float bf, float fx,
                              float* u_right, float* depth, int capacity, uint64_t* ncompares);

/* ---- stateless matchers on caller-provided (host) arrays --------------------------------------
 * A frame as the matchers see it: undistorted keypoints (Frame::mvKeysUn), descriptors
 * (Frame::mDescriptors, n x 32), image bounds (mnMinX..mnMaxY) from which the 64x48 grid is
 * built exactly as Frame.cc:97-98,230-245,382-392 do. */
typedef struct orbm_frame {
    int32_t n;
    const orbx_keypoint* keys_un;
    const uint8_t* descriptors;
    float min_x, max_x, min_y, max_y;
    const float* u_right;     /* Frame::mvuRight (n) or NULL (= all -1) */
    const float* scale_factors; /* Frame::mvScaleFactors (nlevels) or NULL */
    int32_t nlevels;
} orbm_frame;

/* ORBmatcher::DescriptorDistance (src/ORBmatcher.cc:1647-1663) for `count` row pairs on the
 * device: dist[i] = hamming(a[i], b[i]).  Host pointers. */
int orbm_descriptor_distance(const uint8_t* a, const uint8_t* b, int count, int32_t* dist);

/* ORBmatcher::SearchForInitialization (src/ORBmatcher.cc:405-520).  prev_matched: in/out
 * [f1->n][2] floats (vbPrevMatched); matches12: out [f1->n].  Returns match count in *nmatches. */
int orbm_search_for_initialization(const orbm_frame* f1, const orbm_frame* f2, float* prev_matched,
                                   int32_t* matches12, int window_size, float nnratio,
                                   int check_orientation, int32_t* nmatches);

/* The four ORBmatcher::SearchByProjection variants share one device skeleton.  The C++ wrapper
 * flattens MapPoints into queries (projection stays on the host, src/ORBmatcher.cc:1340-1376):
 * per query: window centre (u,v), radius, level window [min_level,max_level] passed to
 * GetFeaturesInArea (src/Frame.cc:327-380), descriptor, and per-variant extras. */
typedef struct orbm_proj_query {
    float u, v, radius;
    int32_t min_level, max_level;
    float u_right;        /* predicted right coordinate (mTrackProjXR / u - mbf*invz)          */
    float max_right_err;  /* stereo gate for keypoints with mvuRight>0:
                             |u_right - mvuRight[idx]| > max_right_err -> skip; <0 = gate off   */
    float angle;          /* keypoint angle of the source observation (rotation histogram)     */
    int32_t level_lo, level_hi; /* extra per-candidate octave filter applied after the area query
                                   (KeyFrame/Sim3 variant, src/ORBmatcher.cc:374-377): kpLevel in [lo, hi].
                                   level_lo < 0 = filter OFF, so a predicted level 0 is passed as [0, 0] */
} orbm_proj_query;

enum {
    ORBM_PROJ_LOCALMAP = 0,  /* (Frame&, vector<MapPoint*>&, th)      src/ORBmatcher.cc:45-129   */
    ORBM_PROJ_LASTFRAME = 1, /* (Frame&, const Frame&, th, bMono)     src/ORBmatcher.cc:1328-1470 */
    ORBM_PROJ_KEYFRAME = 2,  /* (Frame&, KeyFrame*, set&, th, ORBdist) src/ORBmatcher.cc:1472-1599 */
    ORBM_PROJ_SIM3 = 3       /* (KeyFrame*, Scw, vpPoints, vpMatched, th) src/ORBmatcher.cc:290-403 */
};

/* Flattening rules (host/orbx.hpp ProjectionFlatten spells them as code; tests/matcher_cases.py checks them against the
 * reference's own functions): LOCALMAP radius = RadiusByViewingCos(cos) [* th if th != 1] * scale[level] where the float cosine
 * is compared with the DOUBLE literal 0.998; levels [level-1, level]; max_right_err = radius.  LASTFRAME radius = th *
 * scale[last octave]; levels [oct, -1] forward, [0, oct] backward, else [oct-1, oct+1]; u_right = u - mbf*invzc.  KEYFRAME levels
 * [level-1, level+1].  SIM3 / Fuse: min_level = max_level = -1, level_lo = max(level-1, 0), level_hi = level.  Points the
 * reference skips before its search (bad, already found, outside the image bounds / IsInImage, out of the distance range, not in
 * view) are simply not in the query list.
 * occupied: in/out [frame->n] bytes. in: 1 if the keypoint already holds a map point that blocks
 * new matches (Observations()>0 for LOCALMAP/LASTFRAME, non-NULL for KEYFRAME/SIM3);
 * assigned: out [frame->n], index of the query that finally owns keypoint i, or -1 (the
 * reference overwrites F.mvpMapPoints[bestIdx] in query order; rotation-culled ones are -1).
 * query_blocks: [nq] bytes, 1 if a keypoint taken by this query blocks later queries
 * (pMP->Observations()>0); ignored for KEYFRAME/SIM3 (always blocks).
 * best_idx/best_dist: out [nq] per-query chosen keypoint (-1 if none) and distance.
 * th_dist: TH_HIGH(100) / ORBdist / TH_LOW(50) per variant. */
int orbm_search_by_projection(int variant, const orbm_frame* frame, const orbm_proj_query* queries,
                              const uint8_t* query_desc, int nq, const uint8_t* query_blocks,
                              uint8_t* occupied, int32_t* assigned, int32_t* best_idx,
                              int32_t* best_dist, int th_dist, float nnratio,
                              int check_orientation, int32_t* nmatches);

/* Frame::ComputeStereoMatches (src/Frame.cc:466-638) on caller-provided data: keypoints and
 * descriptors of both eyes plus both raw pyramids (mvImagePyramid of the two extractors; level l
 * of eye e at pyr[e][l] with row stride pyr_stride[e][l], WITHOUT the 19-px border). */
typedef struct orbm_stereo_input {
    int32_t n_left, n_right;
    const orbx_keypoint* keys_left;
    const orbx_keypoint* keys_right;
    const uint8_t* desc_left;
    const uint8_t* desc_right;
    int32_t nlevels;
    const float* scale_factors;
    const float* inv_scale_factors;
    const uint8_t* const* pyr_left;   /* nlevels pointers */
    const uint8_t* const* pyr_right;
    const int32_t* level_width;       /* nlevels */
    const int32_t* level_height;
    const int64_t* pyr_stride_left;   /* nlevels */
    const int64_t* pyr_stride_right;
    float bf, fx;                     /* mbf and fx: mb = mbf/fx, maxD = mbf/mb */
} orbm_stereo_input;
int orbm_compute_stereo_matches(const orbm_stereo_input* in, float* u_right, float* depth,
                                uint64_t* ncompares);

/* Brute-force 2-nearest-neighbour Hamming search (BASELINE.json config 5): for every query the
 * best and second-best database rows under DescriptorDistance; ties -> lower database index.
 * Pointers are DEVICE pointers (the bench keeps 1M x 16M resident); out_dist: [nq][2] uint16,
 * out_idx: [nq][2] int32 (global index = db_index_offset + local row).  Asynchronous on
 * `stream` (cudaStream_t or NULL = default stream). */
int orbm_knn2_device(const uint8_t* queries, int64_t nq, const uint8_t* database, int64_t ndb,
                     int32_t db_index_offset, uint16_t* out_dist, int32_t* out_idx, void* stream);
/* orbm_knn2_device for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.  Same arguments, same
 * results bit for bit; calls no cudaMalloc / cudaFree / allocating routine.  The per-split partial results live in an
 * 8 MiB file-scope device array (ORBM_KNN2_NOALLOC_SCRATCH_BYTES) that every GPU loading the library holds until exit.
 * Limits: at most 8 database splits and splits * nq * 12 bytes <= 8 MiB -- i.e. any nq when ndb <= 8388608 rows (the call
 * then runs unsplit if it must), nq <= 87381 at 67108864 rows; beyond that ORBX_E_CAPACITY with the sizes in
 * orbx_last_error().  All calls on a device SHARE the array: calls that overlap on the GPU (two streams, two threads)
 * race; calls on one stream are safe.  Stale contents of an earlier call are never read. */
#define ORBM_KNN2_NOALLOC_SCRATCH_BYTES (8u << 20)
int orbm_knn2_device_noalloc(const uint8_t* queries, int64_t nq, const uint8_t* database, int64_t ndb,
                             int32_t db_index_offset, uint16_t* out_dist, int32_t* out_idx, void* stream);
/* Tuning hook: how the 256-bit popcount is evaluated (0 = 8 POPC, 1 = carry-save 5 POPC [default],
 * 2 = Harley-Seal 4 POPC).  Results are identical; only the instruction mix changes. */
int orbm_knn2_set_variant(int variant);
/* Measurement hook (bench.py): register-only rate probes of the integer pipes K10 lives on.  rates[0..2] = popc.b32 /
 * lop3.b32 / add.u32 results per second over all SMs, rates[3] = Hamming compares per second of K10's inner loop on
 * register operands (current popcount variant): the measured roofline of orbm_knn2_device.  Synchronous, ~10 ms. */
int orbm_measure_int_rates(double rates[4]);
/* Merge `nparts` per-shard results (each [nq][2]) into one, same tie rule; device pointers. */
int orbm_knn2_merge_device(const uint16_t* part_dist, const int32_t* part_idx, int nparts,
                           int64_t nq, uint16_t* out_dist, int32_t* out_idx, void* stream);

/* ---- config 5 across GPUs: database sharded by rows, one rank (process or thread) per GPU ------------------
 * NCCL is bound at run time (dlopen of libnccl.so.2), so only callers of these four entry points need it.
 * A communicator is either the caller's own ncclComm_t (passed as void*) or one made here: rank 0 obtains an id,
 * ships its 128 bytes to the other ranks by any means, every rank calls orbm_nccl_comm_create on its device. */
#define ORBM_NCCL_ID_BYTES 128
int orbm_nccl_unique_id(uint8_t id[ORBM_NCCL_ID_BYTES]);
int orbm_nccl_comm_create(const uint8_t id[ORBM_NCCL_ID_BYTES], int rank, int nranks, void** comm);
int orbm_nccl_comm_destroy(void* comm);
/* orbm_knn2_device over a row-sharded database: queries (replicated on every rank) against this rank's shard
 * [db_index_offset, db_index_offset + ndb_local), ncclAllGather of the per-query (best, second) records, merge under
 * the (distance, global index) order: every rank ends with the result of the whole database in out_dist / out_idx.
 * Queries are processed in chunks of query_chunk (0 = half of nq, at least 2 x SMs x 1024; rounded to 1024): the gather and
 * merge of a chunk run on an internal stream while the next chunk is scanned.  All ranks must call with the same nq and
 * query_chunk.  comm == NULL (or a communicator of one rank with query_chunk == 0) degenerates to orbm_knn2_device.  Device pointers;
 * asynchronous on `stream`. */
int orbm_knn2_sharded(void* comm, const uint8_t* queries, int64_t nq, const uint8_t* db_shard, int64_t ndb_local,
                      int32_t db_index_offset, int64_t query_chunk, uint16_t* out_dist, int32_t* out_idx, void* stream);

/* ================================================================================================
 * RGB-D frame front: the steps of Tracking::GrabImageRGBD / Frame::Frame(RGB-D)
 * either side of ORBextractor::operator().  Arithmetic = OpenCV 4.13 (oracle/frame_oracle.cpp).
 * ================================================================================================ */

/* mK (fx, fy, cx, cy) and mDistCoef (k1, k2, p1, p2, k3) as CV_32F (src/Tracking.cc:60-80), mbf */
typedef struct orbx_camera {
    float fx, fy, cx, cy;
    float k1, k2, p1, p2, k3;
    float bf;
} orbx_camera;

/* cv::cvtColor(RGB|BGR|RGBA|BGRA -> GRAY) of src/Tracking.cc:308-321 (and :265-290), 8-bit, batched.
 * channels 3 or 4; rgb_order != 0 for RGB(A) (mbRGB), 0 for BGR(A).  color/gray are host pointers, or device
 * pointers when on_device != 0.  Strides in bytes. */
int orbx_cvt_gray(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t frame_stride,
                  size_t row_stride, int width, int height, int nframes, int on_device, uint8_t* gray,
                  size_t gray_frame_stride, size_t gray_row_stride);

/* imDepth.convertTo(imDepth, CV_32F, mDepthMapFactor) of src/Tracking.cc:324-327 for a 16-bit depth map
 * (strides in ELEMENTS); host pointers, or device pointers when on_device != 0. */
int orbx_depth_to_float(orbx_handle* h, const uint16_t* depth, size_t row_stride, int width, int height, float factor,
                        int on_device, float* out, size_t out_row_stride);

/* cv::undistortPoints(mat, mat, mK, mDistCoef, cv::Mat(), mK) as Frame::UndistortKeyPoints calls it
 * (src/Frame.cc:404-434): n points (x, y) in, n points out; host arrays, computed on the device. */
int orbx_undistort_points(orbx_handle* h, const float* xy, int n, const orbx_camera* cam, float* xy_out);

/* Frame::ComputeImageBounds (src/Frame.cc:436-464): bounds = mnMinX, mnMaxX, mnMinY, mnMaxY */
int orbx_image_bounds(orbx_handle* h, int cols, int rows, const orbx_camera* cam, float bounds[4]);

/* Batched Frame::Frame(imGray, imDepth, ...) front (src/Frame.cc:136-189) with GrabImageRGBD's conversions:
 * colour -> gray (channels 1 = already gray), ExtractORB, UndistortKeyPoints, ComputeStereoFromRGBD (depth sampled at
 * the distorted keypoint, scaled by depth_factor; depth_is_u16 selects 16-bit or float input), ComputeImageBounds,
 * AssignFeaturesToGrid.  Results stay in slots first_slot ..: mvKeys/descriptors as after orbx_extract_batch, plus
 * mvKeysUn, mvuRight, mvDepth and the grid; the matchers of this library then run on mvKeysUn and these bounds.
 * Images are host pointers, or device pointers when on_device != 0.  Strides in bytes. */
int orbx_frames_rgbd(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t color_frame_stride,
                     size_t color_row_stride, const void* depth, int depth_is_u16, float depth_factor,
                     size_t depth_frame_stride, size_t depth_row_stride, int width, int height, int nframes,
                     int on_device, const orbx_camera* cam, int first_slot);

/* ONE RGB-D frame, host images in, the whole Frame front out — Tracking::GrabImageRGBD + Frame::Frame(RGB-D) for the image pair of
 * one tracked frame (src/Tracking.cc:300-339, src/Frame.cc:136-189): the fork's main mode, built for latency like orbx_extract
 * (pinned staging or in-place DMA of page-locked images, cvtColor straight into the extractor's level 0, the frame's captured
 * CUDA graph, the depth map copied by the host while the GPU extracts, one batch of result copies, one synchronisation).
 * Outputs (host, room for `capacity` keypoints, each may be NULL): kps = mvKeys, keys_un = mvKeysUn, desc = mDescriptors,
 * u_right = mvuRight, depth_out = mvDepth, *n = N, bounds = mnMinX, mnMaxX, mnMinY, mnMaxY.  The frame also stays in slot 0 with
 * its grid, so the matchers of this library can run on it.  Empty image -> outputs untouched. */
int orbx_frame_rgbd(orbx_handle* h, const uint8_t* color, int channels, int rgb_order, size_t color_row_stride, const void* depth,
                    int depth_is_u16, float depth_factor, size_t depth_row_stride, int width, int height, const orbx_camera* cam,
                    orbx_keypoint* kps, orbx_keypoint* keys_un, uint8_t* desc, float* u_right, float* depth_out, int capacity, int* n,
                    float bounds[4]);

/* mvKeysUn (full KeyPoint records), mvuRight, mvDepth of slots first_slot ..; each output may be NULL.
 * keys_un: nframes*capacity records, u_right/depth: nframes*capacity floats, bounds: 4 floats. */
int orbx_download_frame_front(orbx_handle* h, int first_slot, int nframes, orbx_keypoint* keys_un, float* u_right,
                              float* depth, int capacity, float bounds[4]);

/* ================================================================================================
 * Bag-of-words matcher path: DBoW2's per-feature transform and both
 * ORBmatcher::SearchByBoW variants, flattened to plain arrays.
 * ================================================================================================ */

/* A DBoW2 vocabulary tree (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h) flattened: node 0 = root, the children of
 * node i are children[child_begin[i] .. child_begin[i] + child_count[i]), child_count == 0 marks a leaf (word). */
typedef struct orbm_vocabulary {
    int nnodes;
    int depth;                      /* m_L */
    const int32_t* child_begin;     /* [nnodes] */
    const int32_t* child_count;     /* [nnodes] */
    const int32_t* children;        /* node ids */
    const uint8_t* descriptors;     /* [nnodes][32] */
    const uint32_t* word_id;        /* [nnodes], meaningful for leaves */
    const double* weight;           /* [nnodes], WordValue of the leaf (0 = stopped word) */
} orbm_vocabulary;
typedef struct orbm_vocab_handle orbm_vocab_handle;

int orbm_vocabulary_upload(const orbm_vocabulary* voc, int device, orbm_vocab_handle** out);
int orbm_vocabulary_free(orbm_vocab_handle* v);

/* TemplatedVocabulary::transform(feature, word_id, weight, &nid, levelsup) (TemplatedVocabulary.h:1231-1271, distance =
 * FORB::distance FORB.cpp:81-101) for n features: what ::transform(features, BowVector&, FeatureVector&, levelsup)
 * (:1140-1218, called from Frame::ComputeBoW src/Frame.cc:394-402) evaluates per feature.  The caller accumulates
 * BowVector[word_id] += weight and FeatureVector[node_id].push_back(i) for the features with weight > 0.
 * features: host, n x 32 bytes.  Outputs: host arrays of n. */
int orbm_bow_transform(orbm_vocab_handle* v, const uint8_t* features, int n, int levelsup, uint32_t* word_id,
                       double* weight, uint32_t* node_id);

/* One side of SearchByBoW.  flags: bit 0 = the feature is in the FeatureVector (its word weight is > 0),
 * bit 1 = the feature has a MapPoint that is not bad. */
typedef struct orbm_bow_frame {
    int n;
    const uint8_t* descriptors;     /* n x 32 */
    const float* angles;            /* mvKeysUn[i].angle */
    const uint32_t* node_id;        /* FeatureVector node of feature i (orbm_bow_transform) */
    const uint8_t* flags;
} orbm_bow_frame;

/* ORBmatcher::SearchByBoW, npairs independent frame pairs per call (one CTA each).
 * variant 0 = (KeyFrame* pKF, Frame& F, vpMapPointMatches)  src/ORBmatcher.cc:159-288: a = keyframe, b = frame;
 *             match_b[j] = keyframe feature whose MapPoint goes into vpMapPointMatches[j] (-1 = NULL).
 * variant 1 = (KeyFrame* pKF1, KeyFrame* pKF2, vpMatches12) src/ORBmatcher.cc:522-655: match_a[i] = feature of KF2
 *             whose MapPoint goes into vpMatches12[i].
 * Both arrays are filled in both variants; pair p's entries start at the sum of the earlier pairs' a[].n / b[].n.
 * nmatches: npairs counts (the function's return value in the reference).  All pointers are host pointers. */
int orbm_search_by_bow(int variant, const orbm_bow_frame* a, const orbm_bow_frame* b, int npairs, float nnratio,
                       int check_orientation, int device, int32_t* match_a, int32_t* match_b, int32_t* nmatches);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:242-305) for npoints map points
 * per call: point p owns descriptors offsets[p] .. offsets[p+1] of `descriptors` (its observations in good keyframes,
 * in the order the caller walked mObservations).  best_idx[p] = index inside the group of the descriptor with the
 * least median Hamming distance to the others (first minimum; -1 for an empty group); best_median (may be NULL)
 * receives that median.  Host pointers. */
int orbm_distinctive_descriptors(const uint8_t* descriptors, const int32_t* offsets, int npoints, int device,
                                 int32_t* best_idx, int32_t* best_median);

/* ================================================================================================
 * Keyframe-rate matchers: SearchForTriangulation, Fuse x2, SearchBySim3.
 * Poses, projections and the MapPoint graph stay on the host (as for SearchByProjection above): the
 * wrapper flattens what the search loops read, the kernels return indices, the wrapper replays the
 * side effects (AddObservation / Replace / vpMatches12) in the reference's order.
 * ================================================================================================ */

/* A KeyFrame as these matchers read it. */
typedef struct orbm_keyframe {
    orbm_frame frame;               /* mvKeysUn, mDescriptors, mnMinX.., mvuRight, mvScaleFactors, nlevels */
    const float* level_sigma2;      /* mvLevelSigma2 [nlevels]    (SearchForTriangulation)              */
    const float* inv_level_sigma2;  /* mvInvLevelSigma2 [nlevels] (Fuse variant 0)                      */
    const uint32_t* node_id;        /* FeatureVector node of feature i (SearchForTriangulation), else NULL */
    const uint8_t* flags;           /* bit 0 = feature is in mFeatVec, bit 1 = GetMapPoint(i) != NULL; NULL = all 0 */
} orbm_keyframe;

/* ORBmatcher::SearchForTriangulation (src/ORBmatcher.cc:657-823; CheckDistEpipolarLine :140-157) for npairs
 * (KF1, KF2) pairs (LocalMapping::CreateNewMapPoints walks the neighbours of one keyframe, src/LocalMapping.cc:240-268).
 * f12: npairs x 9 floats (row-major F12); epipole: npairs x 2 floats (ex, ey of :664-670, computed by the caller).
 * matches12: out, vMatches12 of every pair, pair p starts at the sum of the earlier kf1[].frame.n; the caller builds
 * vMatchedPairs from it in index order (:812-820).  nmatches: npairs counts.  Host pointers. */
int orbm_search_for_triangulation(const orbm_keyframe* kf1, const orbm_keyframe* kf2, int npairs, const float* f12,
                                  const float* epipole, int only_stereo, int check_orientation, int device,
                                  int32_t* matches12, int32_t* nmatches);

enum {
    ORBM_FUSE_MAPPOINTS = 0, /* Fuse(pKF, vpMapPoints, th)             src/ORBmatcher.cc:825-973: chi2 gate, <= TH_LOW   */
    ORBM_FUSE_SIM3 = 1,      /* Fuse(pKF, Scw, vpPoints, th, vpReplace) src/ORBmatcher.cc:975-1100: no gate, <= TH_LOW   */
    ORBM_FUSE_SIM3_LEG = 2   /* one direction of SearchBySim3            src/ORBmatcher.cc:1146-1220: no gate, <= TH_HIGH */
};

/* The per-MapPoint search of ORBmatcher::Fuse.  One query per MapPoint that passed the caller's projection tests
 * (:844-892): (u, v), radius = th * mvScaleFactors[nPredictedLevel], u_right = u - bf*invz, level_lo = nPredictedLevel-1,
 * level_hi = nPredictedLevel; the other orbm_proj_query fields are ignored.  best_idx[q] = the keypoint of
 * KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:607-646) order with the first least distance that passes the variant's
 * gate and threshold, else -1; best_dist[q] = its distance (256 when no candidate).  Queries are independent. */
int orbm_fuse(int variant, const orbm_keyframe* kf, const orbm_proj_query* queries, const uint8_t* query_desc, int nq,
              int device, int32_t* best_idx, int32_t* best_dist);

/* ORBmatcher::SearchBySim3 (src/ORBmatcher.cc:1102-1326): queries12 = MapPoints of KF1 (feature index index1[q])
 * projected into KF2, queries21 the other way round, both prepared like orbm_fuse queries.  match12: out [kf1->frame.n],
 * the KF2 feature whose MapPoint goes into vpMatches12[i1] after the agreement check (:1307-1322), or -1. */
int orbm_search_by_sim3(const orbm_keyframe* kf1, const orbm_keyframe* kf2, const orbm_proj_query* queries12,
                        const uint8_t* desc12, const int32_t* index1, int nq12, const orbm_proj_query* queries21,
                        const uint8_t* desc21, const int32_t* index2, int nq21, int device, int32_t* match12,
                        int32_t* nfound);

/* ================================================================================================
 * This fork's dense mapping step: per-keyframe depth back-projection.
 * ================================================================================================ */

/* x, y, z and the colour bytes of a pcl::PointXYZRGBA (b, g, r, a = 255). */
typedef struct orbx_cloud_point {
    float x, y, z;
    uint8_t b, g, r, a;
} orbx_cloud_point;

/* PointCloudMapping::generatePointCloud (src/PointCloudMap.cc:78-110) / convertToPointCloud (:112-139) for nframes
 * keyframes per call: every `step`-th row and column (the reference uses 3) of the CV_32F depth map with
 * 0.01 <= d <= 10 becomes a point z = d, x = (n - cx) * z / fx, y = (m - cy) * z / fy with the BGR bytes of mImRGB, in
 * row-major scan order (the push_back order); then pcl::transformPointCloud(tmp, cloud, T.inverse().matrix()) in
 * double: out = (float)(t00*x + t01*y + t02*z + t03), ...
 * depth: nframes maps, row stride depth_row_stride floats, frame stride depth_frame_stride floats; bgr: 3-channel
 * images with byte strides, or NULL (the PointXYZ overload :148-176: colour bytes 0, a = 255).
 * pose_inv: nframes x 16 doubles, row-major T.inverse().matrix(), or NULL for convertToPointCloud (no transform).
 * points: out, frame f's points start at f * capacity; counts: out [nframes] (a frame that yields more than
 * `capacity` points returns ORBX_E_CAPACITY; ceil(w/step)*ceil(h/step) always suffices).  Host pointers. */
int orbm_generate_point_clouds(const float* depth, size_t depth_row_stride, size_t depth_frame_stride, const uint8_t* bgr,
                               size_t bgr_row_stride, size_t bgr_frame_stride, int width, int height, int nframes, int step,
                               const orbx_camera* cam, const double* pose_inv, int device, orbx_cloud_point* points,
                               int capacity, int32_t* counts);

/* pcl::VoxelGrid<pcl::PointXYZRGBA> down-sampling as PointCloudMapping uses it (src/PointCloudMap.cc:41, 240-241, 259-260,
 * 277-280: leaf = resolution on all three axes, default parameters): one point per occupied voxel, in ascending voxel-index
 * order, at the centroid of the voxel's points (colour = truncated mean, a = 255).  A voxel's points are summed in input order
 * (PCL's own order follows an unstable sort: its float sums are not bit-defined; the set of occupied voxels and the colours are).
 * points: n records (host); out: room for `capacity` records (NULL: only count); *nout = voxels.  ORBX_E_CAPACITY when the grid
 * would overflow PCL's int voxel index (PCL then returns the cloud unfiltered) or when capacity is too small. */
int orbm_voxel_grid(const orbx_cloud_point* points, int n, float leaf, int device, orbx_cloud_point* out, int capacity, int32_t* nout);

/* pcl::StatisticalOutlierRemoval<pcl::PointXYZRGBA> as PointCloudMapping applies it to the global map before saving it
 * (src/PointCloudMap.cc:46-47 setMeanK(50) / setStddevMulThresh(1.0), :284-285 sor.setInputCloud(globalMapRGBD); sor.filter(*tmp)):
 * every finite point's mean distance to its mean_k nearest neighbours (exact search; squared distances accumulated in float as
 * FLANN's L2_Simple does, square roots summed in double in ascending order), mean and standard deviation of those distances,
 * points with distance > mean + stddev_mul * stddev removed, input order kept.  Non-finite points get distance 0 and stay; with
 * fewer than mean_k + 1 finite points everything stays.  1 <= mean_k <= 127.
 * points: n records (host); out: room for `capacity` records (NULL: only count); *nout = points kept; mean_dist: NULL or n floats;
 * stats: NULL or 5 doubles (threshold, mean, stddev, points with a valid distance, fraction of tile pairs the search visited). */
int orbm_statistical_outlier_removal(const orbx_cloud_point* points, int64_t n, int mean_k, double stddev_mul, int device,
                                     orbx_cloud_point* out, int64_t capacity, int64_t* nout, float* mean_dist, double* stats);

/* The per-feature record of Map::Save's keyframe block (src/Map.cc:160-180): 64 bytes = pt.x, pt.y, size, angle,
 * response (float), octave (int), the 32 descriptor bytes, and the unsigned long index of the feature's MapPoint
 * (ULONG_MAX = none).  Packs the features of result slot `slot` (mvKeys, mDescriptors as the extractor left them on
 * the device) into `out` (host, capacity records); mp_index: host [n] or NULL (all ULONG_MAX).  *n = record count. */
int orbx_pack_map_records(orbx_handle* h, int slot, const uint64_t* mp_index, uint8_t* out, int capacity, int32_t* n);

#ifdef __cplusplus
}
#endif
#endif /* ORBX_H */
