/* orb_b200.h — C ABI of the B200-native ORB-SLAM2 front-end hot path.
 *
 * Every entry point replaces one piece of the reference's C++ front end (paths are relative to
 * the reference tree yg838457845/ORB_SLAM2-Chinese-annotation):
 *
 *   orb_extractor_create / destroy      ORBextractor::ORBextractor            src/ORBextractor.cc:428-489
 *   orb_extractor_tables                GetScaleFactors() & friends           include/ORBextractor.h:63-83
 *   orb_extract / orb_extract_batch     ORBextractor::operator()              src/ORBextractor.cc:1091-1169
 *   orb_extract_device                  same, inputs already resident in HBM (bench "value" path)
 *   orb_get_level                       public member mvImagePyramid          include/ORBextractor.h:85
 *   orb_descriptor_distance             ORBmatcher::DescriptorDistance        src/ORBmatcher.cc:1814-1830
 *   orb_match_bruteforce                all-pairs best / second-best Hamming  (inner loop of every Search*)
 *   orb_assign_features_to_grid         Frame::AssignFeaturesToGrid           src/Frame.cc:261-276, 426-436
 *   orb_search_by_projection_mappoints  ORBmatcher::SearchByProjection(F,MPs) src/ORBmatcher.cc:47-133
 *   orb_search_by_projection_lastframe  ORBmatcher::SearchByProjection(C,L)   src/ORBmatcher.cc:1460-1619
 *   orb_search_by_bow                   ORBmatcher::SearchByBoW(KF,F)         src/ORBmatcher.cc:164-306
 *   orb_search_for_triangulation        ORBmatcher::SearchForTriangulation    src/ORBmatcher.cc:718-900
 *   orb_stereo_match                    Frame::ComputeStereoMatches           src/Frame.cc:516-704
 *
 * Conventions: plain pointers and sizes only; the library never allocates caller-visible memory
 * (the caller passes capacities); every function returns an orb_status (0 = ok).  All pointers are
 * HOST pointers unless the parameter name starts with d_.  A handle owns its device buffers and
 * one CUDA stream; distinct handles may be used concurrently from different threads (the
 * reference runs the left and right extractor on two threads, src/Frame.cc:81-84), one handle is
 * not re-entrant.  There is no CPU fallback: without a usable CUDA device every call fails with
 * ORB_ERR_CUDA.
 */
#ifndef ORB_B200_H
#define ORB_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    ORB_OK = 0,
    ORB_ERR_BAD_ARG = 1,   /* null pointer, non-positive size, unsupported geometry */
    ORB_ERR_CAPACITY = 2,  /* caller buffer or handle capacity too small */
    ORB_ERR_CUDA = 3,      /* CUDA runtime error (orb_last_error() has the text) */
    ORB_ERR_INTERNAL = 4   /* device-side overflow of an internal bound */
} orb_status;

/* Same 28-byte layout as cv::KeyPoint (pt.x, pt.y, size, angle, response, octave, class_id). */
typedef struct {
    float x, y, size, angle, response;
    int32_t octave, class_id;
} orb_keypoint;

typedef struct orb_extractor orb_extractor;

const char* orb_last_error(void);
int orb_device_count(void);

/* ---------------------------------------------------------------- extractor */
/* nfeatures, scale_factor, nlevels, ini_th_fast, min_th_fast: the reference constructor's
 * arguments (include/ORBextractor.h:51-52).  max_batch images of at most max_w x max_h can be
 * processed per call on CUDA device `device`. */
int orb_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th_fast, int min_th_fast,
                         int max_w, int max_h, int max_batch, int device, orb_extractor** out);
void orb_extractor_destroy(orb_extractor* e);
/* Per-level tables (nlevels entries each; any pointer may be NULL); umax has 16 entries. */
int orb_extractor_tables(const orb_extractor* e, float* scale_factors, float* inv_scale_factors, float* sigma2,
                         float* inv_sigma2, int32_t* features_per_level, int32_t* umax);
/* Upper bound on keypoints one image can produce (>= nfeatures; the quadtree may overshoot). */
int orb_extractor_max_keypoints(const orb_extractor* e);

/* operator(): one 8-bit single-channel image (row stride in bytes).  Writes *n keypoints and
 * n*32 descriptor bytes.  An empty image (w or h == 0) yields *n = 0 like the reference (:1095). */
int orb_extract(orb_extractor* e, const uint8_t* img, int w, int h, int stride, orb_keypoint* kps,
                uint8_t* desc, int cap, int32_t* n);
/* A batch of same-sized images laid out img_stride bytes apart.  Image b writes at kps + b*cap,
 * desc + b*cap*32, counts[b]. */
int orb_extract_batch(orb_extractor* e, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride,
                      orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts);
/* Streaming form of orb_extract_batch (a caller that feeds batch after batch, e.g. a monocular sequence processed offline:
 * Examples/Monocular/mono_tum.cc:60-100 loops over the frames one by one): enqueue() returns as soon as the batch is queued, wait()
 * when its outputs are in the caller's buffers.  Two batches may be in flight on one handle (same geometry and batch size);
 * uploads run a whole call ahead of the kernels.  Tickets are waited for in issue order; the buffers of an enqueued batch stay
 * alive and untouched until its wait() returns; orb_extract_batch must not be called on the handle while a ticket is out. */
int orb_extract_enqueue(orb_extractor* e, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride,
                        orb_keypoint* kps, uint8_t* desc, int cap, int* ticket);
int orb_extract_wait(orb_extractor* e, int ticket, int32_t* counts);
/* Same, but the images are already in device memory and the results stay in the handle's device
 * buffers (see orb_extractor_device_results).  Asynchronous on the handle's stream.  The images are
 * read in aligned 16-byte blocks that hold at least one pixel, i.e. up to 15 bytes before the first and past the last
 * pixel of a row (inside any cudaMalloc'ed buffer, whose size and base are multiples of 256 bytes). */
int orb_extract_device(orb_extractor* e, const uint8_t* d_imgs, int batch, int w, int h, int stride, size_t img_stride);
/* Device views of the last batch's results: keypoints [batch][kp_cap], descriptors [batch][kp_cap][32],
 * counts [batch].  Valid until the next extract call on this handle. */
int orb_extractor_device_results(const orb_extractor* e, const orb_keypoint** d_kps, const uint8_t** d_desc,
                                 const int32_t** d_counts, int* kp_cap);
/* Copies the last batch's results to the host (synchronises the handle's stream). */
int orb_extractor_download(orb_extractor* e, int batch, orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts);
int orb_extractor_sync(orb_extractor* e);
/* CUDA stream (cudaStream_t) the handle launches on; for event timing by the caller. */
void* orb_extractor_stream(const orb_extractor* e);
/* Number of kernel launches issued by the handle since creation. */
long long orb_extractor_launch_count(const orb_extractor* e);
/* Optional CUDA-event timing of the stages of the LAST run on this handle (measurement aid, no
 * reference counterpart).  ms[8]: pyramid, FAST cells, quadtree, blur, orient+describe, stereo
 * rows, stereo match, stereo median (the stereo entries belong to the left handle). */
int orb_extractor_set_timing(orb_extractor* e, int on);
int orb_extractor_stage_ms(orb_extractor* e, float* ms);

/* Pyramid level `level` of image `image` of the last call (mvImagePyramid[level]).  border = 0
 * copies the w x h level, border = 19 reproduces the reference's REFLECT_101 bordered buffer
 * (src/ORBextractor.cc:1181-1203).  blurred != 0 returns the 7x7 Gaussian-blurred level the
 * descriptors were sampled from instead.  *w, *h receive the unbordered level size. */
int orb_get_level(orb_extractor* e, int image, int level, int blurred, int border, uint8_t* dst, int dst_stride,
                  int dst_cap_bytes, int* w, int* h);
/* All levels of image `image` of the last call in one call (the whole mvImagePyramid, unbordered): level l lands in dst[l] with
 * row stride dst_stride[l] >= its width.  One staged device-to-host pass and one synchronisation instead of one per level. */
int orb_get_pyramid(orb_extractor* e, int image, uint8_t* const* dst, const int32_t* dst_stride);
/* Size of pyramid level `level` for the geometry of the last call. */
int orb_level_size(const orb_extractor* e, int level, int32_t* w, int32_t* h);
/* Stage outputs of the last call, for stage-level parity tests: FAST candidates fed to the
 * quadtree (x, y, score triples, window-relative, reference order) and the per-level keypoints
 * chosen by the quadtree, level coordinates (x, y, score). */
int orb_get_level_candidates(orb_extractor* e, int image, int level, int32_t* xys, int cap, int32_t* n);
int orb_get_level_keypoints(orb_extractor* e, int image, int level, int32_t* xys, int cap, int32_t* n);

/* Self-check of the device sincosf used by the rBRIEF kernel (the reference's cos(float)/sin(float),
 * src/ORBextractor.cc:124-125, are one glibc sincosf call).  Evaluates every float whose bit pattern
 * lies in [first_bits, last_bits] on CUDA device `device` and returns the 64-bit checksum that
 * tests/cpp/sincosf_sweep.c computes with the host libm. */
int orb_sincosf_sweep_device(uint32_t first_bits, uint32_t last_bits, int device, uint64_t* checksum);
/* Self-check of the device cv::fastAtan2 used by the orientation stage (IC_Angle, src/ORBextractor.cc:112): out[i] =
 * fastAtan2(y[i], x[i]) in degrees, evaluated by the same device function the kernel calls.  Host pointers. */
int orb_fast_atan2_device(const float* y, const float* x, int n, int device, float* out);

/* ------------------------------------------------------------------ stereo */
/* Frame::ComputeStereoMatches for image `image` of the last batches of the two extractor handles
 * (left and right pyramids stay device-resident).  Keypoints/descriptors are the extractors' own
 * outputs, re-read from device memory.  max_d = mbf / mb (src/Frame.cc:549; passed explicitly
 * because the reference reads Frame::mb before it is assigned), min_d = 0 in the reference.
 * u_right, depth: n_left floats each (-1 = no match). */
int orb_stereo_match(orb_extractor* left, orb_extractor* right, int image, float mbf, float min_d, float max_d,
                     float* u_right, float* depth, int cap, int32_t* n_left);
/* Whole batch: outputs at u_right + b*cap, depth + b*cap.  Asynchronous variant keeps results on
 * the device (d_u_right / d_depth: [batch][kp_cap] device buffers owned by `left`). */
int orb_stereo_match_batch(orb_extractor* left, orb_extractor* right, int batch, float mbf, float min_d, float max_d,
                           float* u_right, float* depth, int cap, int32_t* n_left);
/* The stereo Frame constructor (src/Frame.cc:60-115): ExtractORB on the left and right image
 * (two threads in the reference, :81-84), then ComputeStereoMatches (:95), for a batch of stereo
 * pairs held in HOST memory.  The batch is cut into chunks of image slots whose upload, kernels
 * and download overlap on three streams.  Outputs as orb_extract_batch (per side) and
 * orb_stereo_match_batch; all arrays are [batch][cap] and rows are written whole (the tail beyond
 * n_l[b] / n_r[b] is unspecified).  Pinned host buffers are needed for the copies to overlap.  Calls of at most 8 pairs whose
 * images / result arrays are ordinary pageable memory (a cv::Mat, a std::vector: the per-frame use) are staged through pinned
 * mirrors inside the handles (CPU copy in and out, asynchronous copies across the link) and their kernels are launched as a
 * programmatic dependent chain; results are the same bytes either way. */
int orb_stereo_frames(orb_extractor* left, orb_extractor* right, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch,
                      int w, int h, int stride, size_t img_stride, float mbf, float min_d, float max_d,
                      orb_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l, orb_keypoint* kps_r, uint8_t* desc_r, int32_t* n_r,
                      float* u_right, float* depth, int cap);
/* Asynchronous form: enqueue() returns as soon as the batch is queued, wait() when its outputs are in the caller's buffers.  Two
 * batches may be in flight on one handle pair (same geometry and batch size); chunk c of the second only waits for chunk c of the
 * first, so consecutive batches hide each other's first upload and last download.  Tickets are waited for in issue order; the
 * buffers of an enqueued batch stay alive and untouched until its wait() returns.  orb_stereo_frames == enqueue + wait. */
int orb_stereo_frames_enqueue(orb_extractor* left, orb_extractor* right, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch,
                              int w, int h, int stride, size_t img_stride, float mbf, float min_d, float max_d,
                              orb_keypoint* kps_l, uint8_t* desc_l, orb_keypoint* kps_r, uint8_t* desc_r, float* u_right,
                              float* depth, int cap, int* ticket);
int orb_stereo_frames_wait(orb_extractor* left, orb_extractor* right, int ticket, int32_t* n_l, int32_t* n_r);
/* Diagnostic: the host-buffer entry points of this handle (orb_extract_batch, orb_stereo_frames*; set it on the LEFT handle for
 * the stereo ones) skip the kernels and only move the bytes: the copy ceiling of the pipeline (bench.py --copy-only). */
int orb_extractor_set_copy_only(orb_extractor* e, int on);
int orb_stereo_match_device(orb_extractor* left, orb_extractor* right, int batch, float mbf, float min_d, float max_d);
int orb_stereo_device_results(const orb_extractor* left, const float** d_u_right, const float** d_depth);

/* ----------------------------------------------------------------- matcher */
typedef struct orb_matcher orb_matcher;
/* Workspace for up to max_pairs frame pairs per call, each with at most max_n features per side. */
int orb_matcher_create(int max_pairs, int max_n, int device, orb_matcher** out);
void orb_matcher_destroy(orb_matcher* m);
void* orb_matcher_stream(const orb_matcher* m);
long long orb_matcher_launch_count(const orb_matcher* m);
/* Candidates (keypoints inside the search windows, each one 256-bit Hamming distance) examined by the last projection search. */
long long orb_matcher_last_candidates(const orb_matcher* m);
int orb_matcher_sync(orb_matcher* m);

/* DescriptorDistance for n descriptor pairs (a[i], b[i]); 32 bytes each. */
int orb_descriptor_distance(orb_matcher* m, const uint8_t* a, const uint8_t* b, int n, int32_t* dist);

/* Brute force: for every row of A (n_a x 32) the best and second-best Hamming distance over all
 * rows of B (n_b x 32) with the reference's scan semantics (strict <, first index wins):
 * best_idx/best_dist/second_dist arrays of n_a entries per pair; pairs are laid out
 * max-size apart (a + p*stride_a*32 ...).  */
int orb_match_bruteforce(orb_matcher* m, const uint8_t* a, const uint8_t* b, int pairs, int n_a, int n_b,
                         int32_t* best_idx, int32_t* best_dist, int32_t* second_dist);
int orb_match_bruteforce_device(orb_matcher* m, const uint8_t* d_a, const uint8_t* d_b, int pairs, int n_a, int n_b,
                                int32_t* d_best_idx, int32_t* d_best_dist, int32_t* d_second_dist);

/* Frame grid (include/Frame.h:39-40: 64 x 48 cells).  Bounds and inverse cell sizes as the Frame
 * stores them (mnMinX, mnMinY, mfGridElementWidthInv, mfGridElementHeightInv). */
typedef struct {
    float min_x, min_y, max_x, max_y;
    float inv_w, inv_h;
} orb_grid_params;
/* cell_start: 64*48+1 offsets, cell_items: n indices; cell (ix, iy) is entry ix*48+iy, items in
 * insertion order (ascending keypoint index), keypoints outside the grid are dropped. */
int orb_assign_features_to_grid(orb_matcher* m, const float* kp_x, const float* kp_y, int n, const orb_grid_params* g,
                                int32_t* cell_start, int32_t* cell_items, int32_t* n_items);

/* The same grid, for callers that may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  This entry point has NO counterpart in the reference.
 * It takes no handle and calls no cudaMalloc / cudaFree or any allocating routine: the device copies of the keypoint
 * coordinates and of the results that the kernel works on live in one file-scope __device__ array of 108 KB (csrc/matcher.cu,
 * g_grid_static) that every GPU loading the library holds until the process exits.
 *   - limit: n <= ORB_GRID_STATIC_MAX_N keypoints; a larger n returns ORB_ERR_CAPACITY with a message in orb_last_error();
 *   - kp_x / kp_y / cell_start / cell_items / n_items may each be host or device memory of the CURRENT device
 *     (cudaMemcpyDefault); same layout and capacities as orb_assign_features_to_grid (cell_items: room for n entries, the first
 *     *n_items are written);
 *   - stream: a cudaStream_t of the current device, NULL = the default stream; the call returns after the stream has drained;
 *   - all calls on one device share the array: calls that overlap on the GPU (two host threads, or two streams) race and
 *     return garbage or ORB_ERR_INTERNAL; serialise them.  Nothing in the array has to be zero on entry (the kernel
 *     overwrites all it reads back; the count word is cleared per call). */
#define ORB_GRID_STATIC_MAX_N 8192
int orb_assign_features_to_grid_static(const float* kp_x, const float* kp_y, int n, const orb_grid_params* g,
                                       int32_t* cell_start, int32_t* cell_items, int32_t* n_items, void* stream);

/* One frame's features as SoA (undistorted keypoints mvKeysUn, mvuRight, mDescriptors). */
typedef struct {
    int32_t n;
    const float* x;
    const float* y;
    const int32_t* octave;
    const float* angle;
    const float* u_right;      /* may be NULL: monocular (all -1) */
    const uint8_t* desc;       /* n x 32 */
    orb_grid_params grid;
    const float* scale_factors; /* mvScaleFactors */
    int32_t n_levels;
} orb_frame_view;

/* SearchByProjection(Frame&, vector<MapPoint*>&, th) (src/ORBmatcher.cc:47-133).  Per MapPoint:
 * valid (mbTrackInView && !isBad()), projection (mTrackProjX/Y/XR), predicted level, view cosine,
 * descriptor, blocks (Observations() > 0: a keypoint this point claims is skipped by later points).
 * occupied[i] != 0: keypoint i already holds a MapPoint with observations (:90-93).
 * match[i] receives the index of the MapPoint assigned to keypoint i or -1; returns nmatches. */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* proj_x;
    const float* proj_y;
    const float* proj_xr;
    const int32_t* level;
    const float* view_cos;
    const uint8_t* desc;
    const uint8_t* blocks;
} orb_mappoint_view;
int orb_search_by_projection_mappoints(orb_matcher* m, const orb_frame_view* f, const orb_mappoint_view* mps, float th,
                                       float nn_ratio, const uint8_t* occupied, int32_t* match, int32_t* nmatches);

/* SearchByProjection(Frame& Cur, const Frame& Last, th, bMono) (src/ORBmatcher.cc:1460-1619).
 * Per Last-frame feature: valid (has a MapPoint and is not an outlier), camera-frame position of
 * its MapPoint in the CURRENT frame (x3Dc = Rcw*x3Dw + tcw, computed by the caller with the
 * reference's cv::Mat float algebra), octave and angle in the last frame, descriptor of the MapPoint,
 * blocks (Observations() > 0).  forward/backward as computed at :1486-1487. */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* xc;
    const float* yc;
    const float* zc;
    const int32_t* octave;
    const float* angle;
    const uint8_t* desc;
    const uint8_t* blocks;
} orb_lastframe_view;
typedef struct {
    float fx, fy, cx, cy, mbf;
} orb_camera;
int orb_search_by_projection_lastframe(orb_matcher* m, const orb_frame_view* cur, const orb_lastframe_view* last,
                                       const orb_camera* cam, float th, int forward, int backward,
                                       int check_orientation, const uint8_t* occupied, int32_t* match,
                                       int32_t* nmatches);

/* Batched forms of the two tracking searches (BASELINE.json configs[3]: "batch of 1024 pairs"): `pairs` independent
 * frame pairs per call.  Every array holds `pairs` consecutive blocks of `stride` elements (descriptors: stride * 32
 * bytes), n[p] of them valid; pointers may be host (pinned for full copy speed) or device memory.  The grid constants,
 * scale factors and camera are shared by the batch (one camera).  occupied / match: [pairs][f->stride]; nmatches:
 * [pairs]; forward / backward (src/ORBmatcher.cc:1486-1487): one flag per pair.  The handle must have been created with
 * max_pairs >= pairs.  Results are those of `pairs` single calls. */
typedef struct {
    int32_t pairs, stride;
    const int32_t* n;
    const float* x;
    const float* y;
    const int32_t* octave;
    const float* angle;
    const float* u_right;      /* may be NULL */
    const uint8_t* desc;
    orb_grid_params grid;
    const float* scale_factors;
    int32_t n_levels;
} orb_frame_batch_view;
typedef struct {
    int32_t pairs, stride;
    const int32_t* n;
    const uint8_t* valid;
    const float* proj_x;
    const float* proj_y;
    const float* proj_xr;      /* may be NULL */
    const int32_t* level;
    const float* view_cos;
    const uint8_t* desc;
    const uint8_t* blocks;
} orb_mappoint_batch_view;
typedef struct {
    int32_t pairs, stride;
    const int32_t* n;
    const uint8_t* valid;
    const float* xc;
    const float* yc;
    const float* zc;
    const int32_t* octave;
    const float* angle;
    const uint8_t* desc;
    const uint8_t* blocks;
} orb_lastframe_batch_view;
int orb_search_by_projection_mappoints_batch(orb_matcher* m, const orb_frame_batch_view* f, const orb_mappoint_batch_view* mps,
                                             float th, float nn_ratio, const uint8_t* occupied, int32_t* match,
                                             int32_t* nmatches);
int orb_search_by_projection_lastframe_batch(orb_matcher* m, const orb_frame_batch_view* cur,
                                             const orb_lastframe_batch_view* last, const orb_camera* cam, float th,
                                             const uint8_t* forward, const uint8_t* backward, int check_orientation,
                                             const uint8_t* occupied, int32_t* match, int32_t* nmatches);

/* FeatureVector (DBoW2::FeatureVector = std::map<NodeId, vector<unsigned>>) flattened as CSR in
 * ascending node id: node_ids[n_nodes], start[n_nodes+1], items[start[n_nodes]]. */
typedef struct {
    int32_t n_nodes;
    const uint32_t* node_ids;
    const int32_t* start;
    const uint32_t* items;
} orb_featvec_view;

/* SearchByBoW(KeyFrame*, Frame&, matches) (src/ORBmatcher.cc:164-306).  kf_has_mappoint[i] != 0:
 * KeyFrame feature i has a good MapPoint.  match_f[j] = KeyFrame feature index matched to Frame
 * feature j or -1. */
int orb_search_by_bow(orb_matcher* m, const orb_frame_view* kf, const orb_featvec_view* kf_fv,
                      const uint8_t* kf_has_mappoint, const orb_frame_view* f, const orb_featvec_view* f_fv,
                      float nn_ratio, int check_orientation, int32_t* match_f, int32_t* nmatches);

/* SearchForTriangulation (src/ORBmatcher.cc:718-900).  f12: 3x3 row-major fundamental matrix;
 * (ex, ey): epipole in KF2 (:728-730); *_has_mappoint: feature already has a MapPoint;
 * level_sigma2: pKF2->mvLevelSigma2.  match12[i] = index in KF2 or -1 (pairs = ascending i). */
int orb_search_for_triangulation(orb_matcher* m, const orb_frame_view* kf1, const orb_featvec_view* fv1,
                                 const uint8_t* kf1_has_mappoint, const orb_frame_view* kf2,
                                 const orb_featvec_view* fv2, const uint8_t* kf2_has_mappoint, const float* f12,
                                 float ex, float ey, const float* level_sigma2, int only_stereo,
                                 int check_orientation, int32_t* match12, int32_t* nmatches);

/* SearchByBoW(KeyFrame*, KeyFrame*, matches12) (src/ORBmatcher.cc:581-716).  *_has_mappoint[i] != 0:
 * feature i holds a good MapPoint.  match12[i] = KF2 feature index matched to KF1 feature i or -1. */
int orb_search_by_bow_keyframes(orb_matcher* m, const orb_frame_view* kf1, const orb_featvec_view* fv1,
                                const uint8_t* kf1_has_mappoint, const orb_frame_view* kf2,
                                const orb_featvec_view* fv2, const uint8_t* kf2_has_mappoint, float nn_ratio,
                                int check_orientation, int32_t* match12, int32_t* nmatches);

/* MapPoints projected by the caller into a frame / keyframe (the reference does this with cv::Mat
 * float algebra, MapPoint::PredictScale and the distance / viewing-angle gates before the search:
 * src/ORBmatcher.cc:326-369, :926-968, :1264-1296, :1643-1672).  Per point: valid (passed every
 * gate), projection (u, v[, ur]), search radius (th * mvScaleFactors[predicted level]), accepted
 * keypoint octaves [min_level, max_level] (Frame::GetFeaturesInArea convention: max_level < 0 = open),
 * descriptor, and (rotation check only) the angle of the source keypoint. */
typedef struct {
    int32_t n;
    const uint8_t* valid;
    const float* u;
    const float* v;
    const float* ur;          /* may be NULL (only read by the Fuse chi-square test) */
    const float* radius;
    const int32_t* min_level;
    const int32_t* max_level;
    const uint8_t* desc;      /* n x 32 */
    const float* angle;       /* may be NULL when check_orientation == 0 */
} orb_projected_view;

/* Projection searches that CLAIM keypoints as they go:
 *   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist) (src/ORBmatcher.cc:1622-1759):
 *     occupied[i] = CurrentFrame.mvpMapPoints[i] != NULL, max_dist = ORBdist, rotation check;
 *   SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th) (:311-427):
 *     occupied[i] = vpMatched[i] != NULL, max_dist = TH_LOW, no rotation check.
 * Points are visited in order; the best unclaimed candidate (strict <, first wins) within max_dist
 * is claimed.  match[i] = point index given to keypoint i or -1 (after the rotation cull);
 * point_best_idx[p] (may be NULL) = keypoint claimed by point p before the cull, or -1. */
int orb_search_by_projection_claiming(orb_matcher* m, const orb_frame_view* f, const orb_projected_view* pts,
                                      const uint8_t* occupied, int max_dist, int check_orientation, int32_t* match,
                                      int32_t* point_best_idx, int32_t* nmatches);

/* Independent best candidate per projected point — the search step of Fuse (src/ORBmatcher.cc:903-1054
 * with inv_level_sigma2 = pKF->mvInvLevelSigma2: chi-square gate 5.99 / 7.8 on the reprojection error;
 * :1056-1207 with inv_level_sigma2 = NULL) and of each direction of SearchBySim3 (:1264-1340).
 * best_idx[p] = keypoint index with the smallest distance (first wins) if it is <= max_dist, else -1;
 * best_dist[p] (may be NULL) = that smallest distance (256 when no candidate).  The caller replays
 * the map updates (Replace / AddObservation) in point order, as they touch host-side map state. */
int orb_search_projected_best(orb_matcher* m, const orb_frame_view* f, const orb_projected_view* pts,
                              const float* inv_level_sigma2, int max_dist, int32_t* best_idx, int32_t* best_dist,
                              int32_t* nfound);

/* SearchBySim3 (src/ORBmatcher.cc:1212-1455): pts1_in_kf2[i] = MapPoint of KF1 feature i projected
 * into KF2 (valid = has a good, not already matched MapPoint that passes the gates), pts2_in_kf1
 * likewise; match12[i] = KF2 feature j iff i -> j and j -> i agree, else -1. */
int orb_search_by_sim3(orb_matcher* m, const orb_frame_view* kf1, const orb_projected_view* pts1_in_kf2,
                       const orb_frame_view* kf2, const orb_projected_view* pts2_in_kf1, int32_t* match12,
                       int32_t* nfound);

/* SearchForInitialization (src/ORBmatcher.cc:429-578): level-0 keypoints of F1 searched in a
 * window_size window of F2 around prev_matched; a later, closer match displaces an earlier one.
 * match12[i] = F2 index or -1.  (The caller refreshes vbPrevMatched from match12, :571-574.) */
int orb_search_for_initialization(orb_matcher* m, const orb_frame_view* f1, const orb_frame_view* f2,
                                  const float* prev_matched_x, const float* prev_matched_y, int window_size,
                                  float nn_ratio, int check_orientation, int32_t* match12, int32_t* nmatches);

/* Frame::isInFrustum (src/Frame.cc:303-366) for n MapPoints at once, as Tracking::SearchLocalPoints calls it per
 * point (src/Tracking.cc:1012-1036) before SearchByProjection(Frame, MapPoints).  rcw / tcw / ow = the Frame's
 * mRcw (row-major 3x3), mtcw, mOw; bounds = mnMinX..mnMaxY (the inverse cell sizes are not read).  Per point:
 * world position, normal, mfMinDistance / mfMaxDistance (the 0.8 / 1.2 invariance factors are applied here).
 * Outputs: in_view = mbTrackInView, u / v / ur = mTrackProjX / mTrackProjY / mTrackProjXR, view_cos = mTrackViewCos,
 * dist = |P - Ow| for MapPoint::PredictScale (a logf the caller evaluates: mnTrackScaleLevel).  For points that
 * fail a gate the values computed up to that gate are returned and in_view = 0. */
int orb_is_in_frustum(orb_matcher* m, const float* rcw, const float* tcw, const float* ow, const orb_camera* cam,
                      const orb_grid_params* bounds, float viewing_cos_limit, int n, const float* xw, const float* normal,
                      const float* min_distance, const float* max_distance, uint8_t* in_view, float* u, float* v, float* ur,
                      float* view_cos, float* dist);

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:250-326) for n_points MapPoints: point p owns the
 * descriptors desc[start[p] .. start[p+1]) of its observations, in the order the reference gathers them (:262-268);
 * best_idx[p] = the observation whose median distance to the others is smallest (median = element
 * int(0.5*(N-1)) of the sorted row, first row wins ties), -1 for a point without observations. */
int orb_distinctive_descriptors(orb_matcher* m, int n_points, const int32_t* start, const uint8_t* desc, int32_t* best_idx);

/* ---- Bag of words: Frame::ComputeBoW / KeyFrame::ComputeBoW (src/Frame.cc:439-449, src/KeyFrame.cc:60-71) ----
 * mpORBvocabulary->transform(vCurrentDesc, mBowVec, mFeatVec, 4) is, per descriptor, the tree descent of
 * DBoW2::TemplatedVocabulary<FORB>::transform(feature, word_id, weight, nid, levelsup)
 * (Thirdparty/DBoW2/DBoW2/TemplatedVocabulary.h:1239-1283): at every node go to the child with the smallest
 * Hamming distance (strict <: the first child wins ties) until a leaf; word_id = the leaf's word, node_id = the
 * node passed at level L - levelsup (0 = root when that level is <= 0).  The std::map assembly of BowVector /
 * FeatureVector from (word_id, node_id) and the word weights is host work (include/ORBVocabulary.h).
 *
 * A vocabulary is handed over as loadFromTextFile reads it (:1362-1450): node i (1 <= i < n_nodes, node 0 = root)
 * has parent[i] < i, is_leaf[i] and a 32-byte descriptor desc[i*32 ..]; children keep file order; words are
 * numbered in file order of the leaves.  max_features bounds n of the host entry point. */
typedef struct orb_vocabulary orb_vocabulary;
int orb_vocabulary_create(int levels, int n_nodes, const int32_t* parent, const uint8_t* is_leaf, const uint8_t* desc,
                          int max_features, int device, orb_vocabulary** out);
void orb_vocabulary_destroy(orb_vocabulary* v);
int orb_vocabulary_info(const orb_vocabulary* v, int32_t* n_nodes, int32_t* n_words, int32_t* max_children);
void* orb_vocabulary_stream(const orb_vocabulary* v);
long long orb_vocabulary_launch_count(const orb_vocabulary* v);
int orb_vocabulary_sync(orb_vocabulary* v);
/* host buffers: desc n x 32 -> word_id[n], node_id[n] (synchronous) */
int orb_bow_transform(orb_vocabulary* v, const uint8_t* desc, int n, int levelsup, int32_t* word_id, int32_t* node_id);
/* device buffers (16-byte aligned descriptors, e.g. an extractor's d_desc), asynchronous on the vocabulary's stream */
int orb_bow_transform_device(orb_vocabulary* v, const uint8_t* d_desc, int n, int levelsup, int32_t* d_word_id,
                             int32_t* d_node_id);

/* ---- In front of the extractor: stereo rectification and gray conversion ----
 * cv::remap(imLeft, imLeftRect, M1l, M2l, cv::INTER_LINEAR) with the CV_32F maps of initUndistortRectifyMap
 * (Examples/Stereo/stereo_euroc.cc:97-98,136-137; BORDER_CONSTANT 0) and cv::cvtColor(im, im, CV_RGB2GRAY | CV_BGR2GRAY |
 * CV_RGBA2GRAY | CV_BGRA2GRAY) (src/Tracking.cc:174-199), OpenCV 4.13's 8-bit fixed point bit for bit.  A rectifier
 * holds one camera's maps (dst_w x dst_h floats each, row-major; both NULL = gray conversion only), converted once to
 * the fixed-point form cv::remap derives per call.  max_batch bounds the host entry points. */
typedef struct orb_rectifier orb_rectifier;
int orb_rectifier_create(const float* map_x, const float* map_y, int dst_w, int dst_h, int src_w, int src_h, int max_batch,
                         int device, orb_rectifier** out);
void orb_rectifier_destroy(orb_rectifier* r);
void* orb_rectifier_stream(const orb_rectifier* r);
long long orb_rectifier_launch_count(const orb_rectifier* r);
int orb_rectifier_sync(orb_rectifier* r);
/* host buffers: batch images stacked vertically (image b starts at row b * src_h / b * dst_h), synchronous */
int orb_remap(orb_rectifier* r, const uint8_t* src, int src_stride, int batch, uint8_t* dst, int dst_stride);
int orb_gray(orb_rectifier* r, const uint8_t* src, int w, int h, int channels, int rgb_order, int src_stride, uint8_t* dst,
             int dst_stride);
/* device buffers, asynchronous on the rectifier's stream (dst may be the input of orb_extract_device) */
int orb_remap_device(orb_rectifier* r, const uint8_t* d_src, int src_stride, size_t src_img_stride, int batch, uint8_t* d_dst,
                     int dst_stride, size_t dst_img_stride);
int orb_gray_device(orb_rectifier* r, const uint8_t* d_src, int w, int h, int channels, int rgb_order, int src_stride,
                    size_t src_img_stride, int batch, uint8_t* d_dst, int dst_stride, size_t dst_img_stride);

/* ------------------------------------------------------------------ multi-GPU front end */
/* Frames / stereo pairs of a batch are independent: one handle pair and one host thread per GPU inside the library, GPU g takes the
 * contiguous block orb_multi_slice(batch, n, g) of a HOST batch (pinned buffers for full copy speed) and writes its rows of the
 * caller's [batch][cap] outputs; no collective on the data path (north_star).  n_devices == 0: every visible GPU. */
typedef struct orb_multi orb_multi;
int orb_multi_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int max_w, int max_h,
                     int max_batch_per_gpu, int n_devices, orb_multi** out);
void orb_multi_destroy(orb_multi* m);
int orb_multi_device_count(const orb_multi* m);
int orb_multi_max_keypoints(const orb_multi* m);
void orb_multi_slice(int batch, int n_parts, int part, int* first, int* count);
int orb_multi_stereo_frames(orb_multi* m, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch, int w, int h, int stride,
                            size_t img_stride, float mbf, float min_d, float max_d, orb_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l,
                            orb_keypoint* kps_r, uint8_t* desc_r, int32_t* n_r, float* u_right, float* depth, int cap);
int orb_multi_extract_batch(orb_multi* m, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride,
                            orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts);

/* Guard mode (environment ORB_B200_GUARD=1 before the first allocation): every device buffer of the library carries a 4 KB canary
 * on both sides; this call synchronises the device and reads all canaries back.  *n_buffers = live buffers checked, *n_damaged =
 * buffers whose canary was overwritten (ORB_ERR_CUDA and a description in orb_last_error() if any).  Without the environment
 * variable it reports 0 buffers.  Stand-in for compute-sanitizer memcheck where that tool is not available. */
int orb_debug_guard_check(int32_t* n_buffers, int32_t* n_damaged);

#ifdef __cplusplus
}
#endif
#endif /* ORB_B200_H */
