/* TEST INFRASTRUCTURE — C API of the CPU oracle (restatement of the reference's ORB front-end).
 * Parity pinning: primitives vs cv2 4.13.0 (tests/golden), extractor control flow vs the
 * reference's own src/ORBextractor.cc compiled verbatim (oracle/_ref, see oracle/Makefile).
 * Matchers, grid, stereo, isInFrustum, ComputeDistinctiveDescriptors and the DBoW2 transform are pinned against the
 * reference's own ORBmatcher.cc / Frame.cc / KeyFrame.cc / MapPoint.cc / DBoW2 compiled verbatim
 * (oracle/_ref/ref_match_driver, tests/test_oracle_refmatch.py); remap / cvtColor against cv2 4.13 goldens.
 * Only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline / --impl reference) may use it. */
#ifndef ORB_ORACLE_H
#define ORB_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct {
    float x, y, size, angle, response;
    int32_t octave;
} orc_keypoint; /* 24 bytes; class_id is always -1 in the reference and is not carried */

typedef struct orc_extractor orc_extractor;

/* ---- primitives (for KATs against cv2) ---- */
void orc_resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride, uint8_t* dst, int dw, int dh, int dstride);
void orc_gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride);
/* cv::remap INTER_LINEAR with CV_32FC1 x / y maps, BORDER_CONSTANT 0 (stereo rectification, Examples/Stereo/stereo_euroc.cc:136-137) */
void orc_remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, uint8_t* dst, int dw, int dh, int dstride);
/* cv::cvtColor 8U RGB/BGR/RGBA/BGRA -> GRAY (src/Tracking.cc:174-199) */
void orc_gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, int rgb_order, uint8_t* dst, int dstride);
void orc_border101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b);
/* cv::FAST(view, th, nms=true); returns count, writes up to cap (x,y,score) int triples */
int orc_fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, int32_t* xys, int cap);
float orc_fast_atan2(float y, float x);
void orc_fast_score_map(const uint8_t* img, int w, int h, int stride, int32_t* score); /* w*h, border 3 = 0 */
void orc_sincosf(float x, float* s, float* c); /* the host libm the oracle uses */

/* ---- extractor (reference src/ORBextractor.cc) ---- */
orc_extractor* orc_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th);
void orc_extractor_destroy(orc_extractor*);
/* tables: each out pointer may be NULL; arrays have nlevels entries, umax has 16 */
void orc_extractor_tables(const orc_extractor*, float* sf, float* inv_sf, float* sigma2, float* inv_sigma2,
                          int32_t* n_per_level, int32_t* umax);
/* operator(): returns number of keypoints (<= cap) or -1 if cap too small; 0 for an empty image */
int orc_extract(orc_extractor*, const uint8_t* img, int w, int h, int stride, orc_keypoint* kps, uint8_t* desc, int cap);
/* pyramid level of the LAST orc_extract call (unbordered view): returns pointer, fills w,h,stride */
const uint8_t* orc_level(const orc_extractor*, int level, int* w, int* h, int* stride);
const uint8_t* orc_blurred_level(const orc_extractor*, int level, int* w, int* h, int* stride);
/* stage outputs of the LAST call, for stage KATs.  Candidates: FAST list fed to the quadtree
 * (coords relative to the 16-px border window, reference src/ORBextractor.cc:857-860). */
int orc_level_candidates(const orc_extractor*, int level, int32_t* xys, int cap);
/* keypoints of one level after the quadtree, before scaling (level coords), with angle */
int orc_level_keypoints(const orc_extractor*, int level, orc_keypoint* kps, int cap);
/* quadtree alone (reference DistributeOctTree, canonical tie-break): candidates (x,y,score)
 * relative to the window; returns kept count, writes indices into the candidate list in
 * reference output order. */
int orc_distribute(const int32_t* xys, int n, int min_x, int max_x, int min_y, int max_y, int n_features,
                   int32_t* kept_idx, int cap);
/* same, plus a witness of the sorted phase: stats[0] rounds, [1] adjacent equal counts in the sorted lists, [2] early breaks between equal counts */
int orc_distribute_stats(const int32_t* xys, int n, int minX, int maxX, int minY, int maxY, int N, int32_t* kept_idx, int cap, int32_t* stats);

/* ---- matcher (reference src/ORBmatcher.cc) and stereo (src/Frame.cc:516-704); match_oracle.cpp ----
 * The view structs (orb_frame_view, ...) are the plain C types of include/orb_b200.h. */
int orc_descriptor_distance(const uint8_t* a, const uint8_t* b);
/* Frame::ComputeStereoMatches; the two oracle extractors hold the pyramids of their last call.
 * Returns the number of matches kept after the median cut. best_sad may be NULL. */
int orc_stereo_match(const orc_extractor* L, const orc_extractor* R, const orc_keypoint* kpL, const uint8_t* dL, int nL,
                     const orc_keypoint* kpR, const uint8_t* dR, int nR, const float* sf, const float* inv_sf, float mbf, float minD,
                     float maxD, float* uRight, float* depth, int32_t* best_sad);
void orc_match_bruteforce(const uint8_t* a, const uint8_t* b, int n_a, int n_b, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist);

#ifdef __cplusplus
}
#endif
#endif
