// Internal definitions shared by the extractor, stereo and matcher translation units.
#pragma once
#include <cuda_runtime.h>
#include <cuda.h>   // CUtensorMap only; the encoder is fetched with cudaGetDriverEntryPoint (no libcuda link)
#include <stdint.h>
#include <stddef.h>
#include <string>
#include <utility>
#include <cstdlib>
#include <vector>
#include "../../include/orb_b200.h"

#define ORB_MAX_LEVELS 16
// timed stages: 0 pyramid (7 resize launches) 1 FAST cells 2 quadtree 3 blur 4 orient+describe
//               5 stereo rows 6 stereo match 7 stereo median
#define ORB_N_STAGES 8
#define ORB_EDGE 19        // EDGE_THRESHOLD   (reference src/ORBextractor.cc:74)
#define ORB_HALF_PATCH 15  // HALF_PATCH_SIZE  (:73)
#define ORB_PATCH 31       // PATCH_SIZE       (:72)
#define ORB_WIN 16         // minBorderX = EDGE_THRESHOLD - 3 (:797)

// Candidate / keypoint packing used between FAST, the quadtree and the descriptor stage:
// bits 0..11 x, 12..23 y (window-relative level coordinates), 24..31 FAST score (<= 254).
#define ORB_KEY_X(k) ((int)((k) & 0xfffu))
#define ORB_KEY_Y(k) ((int)(((k) >> 12) & 0xfffu))
#define ORB_KEY_S(k) ((int)((k) >> 24))
#define ORB_MAX_DIM 4096  // largest level width/height representable in a packed key

struct LevelGeom {
    int w, h, pitch;            // level size and row pitch in bytes
    unsigned long long off;     // byte offset of image 0 of this level inside the pyramid buffer
    unsigned long long img_stride;  // bytes between consecutive images of this level
    // FAST cell grid (reference src/ORBextractor.cc:797-814)
    int n_cols, n_rows, w_cell, h_cell;
    int max_bx, max_by;         // maxBorderX, maxBorderY
    int cell_off;               // first cell of this level in the per-image cell arrays
    int cell_cap;               // staging slots per cell
    unsigned ncols_magic;       // c / n_cols == __umulhi(c, ncols_magic) for every cell index of the level
    unsigned long long stage_off;   // first staging slot (u32 units) of this level within one image
    // quadtree (reference :558-782)
    int win_w, win_h;           // maxBorderX-minBorderX, maxBorderY-minBorderY
    int n_ini;
    float h_x;
    int n_feat;                 // mnFeaturesPerLevel[level]
    int node_cap;               // bound on list nodes == bound on keypoints of this level
    int kp_off;                 // first slot of this level in the per-image level-keypoint arrays
    // tiles of the blur kernel
    int tiles_x, tiles_y, tile_off;
    // resize coefficient tables (device pointers; level l is resampled from level l-1)
    const int* xofs;
    const short* alpha;
    const int* yofs;
    const short* beta;
    unsigned rz_qx, rz_qy;      // floor(2^16 * w[l-1] / w[l]), same for h: tile origin of the TMA-staged resize without a division
    float sf;                   // mvScaleFactor[level]
    float patch_size;           // (float)(int)(PATCH_SIZE * sf)
};

struct Plan {
    int nlevels, w, h;
    int total_cells, total_tiles, kp_cap;
    unsigned long long stage_per_image;   // u32 slots
    unsigned long long pyr_bytes_per_batch_unit;  // unused by kernels
    int ini_th, min_th;
    int umax[16];
    LevelGeom lv[ORB_MAX_LEVELS];
};

// TMA descriptors (cp.async.bulk.tensor): one 3-D map (x, y, image) per pyramid level over the pitched level
// buffers; the box is the staged tile of the kernel that uses the map.  Out-of-image box elements are zero-filled by
// the hardware; the kernels turn the zeros into BORDER_REFLECT_101 on edge tiles only.
#define ORB_BLUR_TW 128                     // blur: output tile width / height, staged box = tile + halo
#define ORB_BLUR_TH 64
#define ORB_BLUR_BOXW (ORB_BLUR_TW + 32)    // x from tile_x - 16: the innermost box coordinate must be a multiple of 16 bytes (measured: illegal instruction otherwise)
#define ORB_BLUR_BOXH (ORB_BLUR_TH + 6)
#define ORB_RZ_DW 128                       // resize: destination tile, staged source box (scale <= 1.35 / 1.33, else the untiled kernel)
#define ORB_RZ_DH 64
#define ORB_RZ_BOXW 192
#define ORB_RZ_BOXH 88
#define ORB_OD_BLUR_BOXW 64                 // orient + describe: 37 x 37 blurred patch / 31 x 31 raw patch, x rounded down to 16 bytes
#define ORB_OD_BLUR_BOXH 37
#define ORB_OD_RAW_BOXW 48
#define ORB_OD_RAW_BOXH 31
struct TmaMaps { CUtensorMap lv[ORB_MAX_LEVELS]; };

struct orb_extractor {
    int nfeatures = 0, nlevels = 0, ini_th = 0, min_th = 0;
    double scale_factor = 0;
    int max_w = 0, max_h = 0, max_batch = 0, device = 0;
    std::vector<float> sf, inv_sf, sigma2, inv_sigma2;
    std::vector<int> n_per_level;
    int umax[16];
    cudaStream_t stream = nullptr;
    long long launches = 0;

    // geometry currently planned (w, h) and capacities reserved for (max_w, max_h)
    Plan plan;
    int plan_w = -1, plan_h = -1;
    int last_batch = 0;
    TmaMaps tm_blur, tm_resize;     // rebuilt with the plan / batch layout
    TmaMaps tm_od_raw, tm_od_blur;  // patch boxes of k_orient_describe over d_pyr / d_blur
    TmaMaps tm_fast;                // cell-view boxes of k_fast_cells over d_pyr: 80 x (tallest cell + 6)
    TmaMaps tm_st_l, tm_st_r;       // SAD window boxes of k_stereo_refine over d_pyr: 32 x 11 (a left handle's), 48 x 11 (a right handle's)
    bool rz_tma[ORB_MAX_LEVELS] = {false};   // level l is resampled from l-1 by the TMA-staged kernel

    // device buffers (sized for max geometry x max_batch)
    uint8_t* d_pyr = nullptr;
    uint8_t* d_blur = nullptr;
    size_t pyr_cap = 0;             // bytes
    int* d_cell_cnt = nullptr;      // [batch][total_cells]
    int* d_cell_prefix = nullptr;   // [batch][total_cells]
    size_t cells_cap = 0;           // per image
    uint32_t* d_stage = nullptr;    // [batch][stage_per_image]
    uint32_t* d_keys = nullptr;     // [batch][2][stage_per_image]
    size_t stage_cap = 0;           // per image (u32)
    uint32_t* d_lvl_kp = nullptr;   // [batch][kp_cap]
    int* d_lvl_cnt = nullptr;       // [batch][ORB_MAX_LEVELS]
    int* d_lvl_ncand = nullptr;     // [batch][ORB_MAX_LEVELS]
    int kp_cap_max = 0;
    orb_keypoint* d_kps = nullptr;  // [batch][kp_cap]
    uint8_t* d_desc = nullptr;      // [batch][kp_cap][32]
    int* d_counts = nullptr;        // [batch]
    int* d_status = nullptr;        // device-side error flag
    void* d_tables = nullptr;       // resize coefficient tables
    size_t tables_cap = 0;
    // stereo outputs (owned by the LEFT handle)
    float* d_u_right = nullptr;     // [batch][kp_cap]
    float* d_depth = nullptr;
    int* d_sad = nullptr;           // [batch][kp_cap] best SAD per left keypoint (-1 none)
    int2* d_st_jobs = nullptr;      // [batch][kp_cap] refinement jobs between k_stereo_search and k_stereo_refine
    int* d_row_start = nullptr;     // [batch][ORB_MAX_DIM + 1] vRowIndices as CSR
    int2* d_row_items = nullptr;    // [batch][row_items_cap]  (x bits of the right keypoint, index | octave << 24)
    int row_items_cap = 0;
    cudaEvent_t ev_done = nullptr;
    // host-buffer batches are cut into chunks of image slots: upload (s_h2d), kernels (stream) and
    // download (s_d2h) of consecutive chunks overlap
    cudaStream_t s_h2d = nullptr, s_d2h = nullptr;
    cudaStream_t s_aux = nullptr;   // the blur runs here, beside FAST + quadtree (it only needs the pyramid)
    cudaEvent_t ev_pyr = nullptr, ev_blur = nullptr;
    std::vector<cudaEvent_t> ev_pool;
    uint8_t* d_in = nullptr;        // dense upload staging [max_batch][max_w * max_h]
    // Pinned mirror of the head of d_in.  A small call (one chunk) whose images lie in ordinary pageable memory is copied here by the CPU and
    // crosses the link asynchronously from here: cudaMemcpy2DAsync of 376 pageable rows of 1241 bytes held the calling thread for 120 us.
    uint8_t* h_in = nullptr;
    size_t h_in_cap = 0;
    cudaEvent_t ev_hin = nullptr;   // behind the last copy out of h_in
    bool hin_used = false;
    // The same for the results of a one-chunk call into pageable buffers: every device-to-host copy into such memory returns only when
    // it is done (eight of them per stereo pair, one after the other).  They go into this pinned mirror asynchronously (one half per
    // ticket parity) and the valid rows are handed to the caller by the CPU once the call's event has fired.
    uint8_t* h_out = nullptr;
    size_t h_out_cap = 0;
    struct OutPlan {
        bool on = false;
        int n = 0, cap = 0;
        orb_keypoint* kps = nullptr; uint8_t* desc = nullptr; float* u_right = nullptr; float* depth = nullptr;
    } out_plan[2];
    // optional per-stage event timing of the last pipeline run (bench.py roofline numbers)
    bool timing = false;
    cudaEvent_t ev_stage[ORB_N_STAGES + 4] = {nullptr};
    bool stage_valid[ORB_N_STAGES] = {false};
    // pinned host staging
    uint8_t* h_pin = nullptr;
    size_t h_pin_cap = 0;
    int* h_counts = nullptr;        // pinned [2][max_batch]: one half per in-flight stereo batch (ticket parity)
    uint8_t* h_pyr = nullptr;       // pinned staging of one image's levels (orb_get_pyramid), h_pyr_cap bytes
    size_t h_pyr_cap = 0;
    int* h_status = nullptr;        // pinned [4]: device status words (left, right) per ticket parity
    // asynchronous stereo batches (orb_stereo_frames_enqueue / _wait); state lives in the LEFT handle
    int inflight = 0, next_ticket = 0, pipe_chunks = 0;
    bool pipe_split = false;        // the last enqueued call kept its ingest kernels on the kernel streams (dense images)
    int pipe_batch[2] = {0, 0}, pipe_cap[2] = {0, 0};
    cudaEvent_t ev_call[2] = {nullptr, nullptr};
    bool copy_only = false;         // diagnostic: host-buffer entry points skip the kernels (copy ceiling of the pipeline)
};

void orb_set_error(const std::string& s);
int orb_sm_count();   // multiprocessors of the current device (cached per device; launch heuristics only)
// slot-wise building blocks (extractor.cu) used by the chunk-pipelined host entry points
int orb_ensure_plan(orb_extractor* e, int w, int h, int batch);
int orb_run_slots(orb_extractor* e, int b0, int n, int phase = 0);   // phase: see run_pipeline (extractor.cu)
int orb_pipe_streams(orb_extractor* e);
cudaEvent_t orb_pool_event(orb_extractor* e, int i);
int orb_chunk_size(int batch);
// chunk boundaries of a host batch: bounds[0] = 0 < ... < bounds[n] = batch (ramped sizes, see extractor.cu)
std::vector<int> orb_chunk_bounds(int batch);
int orb_upload_slots(orb_extractor* e, const uint8_t* imgs, int b0, int n, int w, int h, int stride, size_t img_stride, cudaStream_t st);
int orb_upload_can_split(const orb_extractor* e, int w, int h, int stride, size_t img_stride);
// results of slots [0, n) -> mirror half tk, asynchronously on st (with_stereo: also d_u_right / d_depth); false = mirror not usable now
bool orb_out_mirror_ready(orb_extractor* e, int n, bool with_stereo);
int orb_download_mirror(orb_extractor* e, int n, int tk, cudaStream_t st, int* h_counts);
int orb_download_mirror_stereo(orb_extractor* e, int n, int tk, cudaStream_t st);
int orb_download_mirror_stereo_row(orb_extractor* e, int image, cudaStream_t st, const float** h_ur, const float** h_dp);
// after the call's event: valid rows of the mirror half -> the buffers recorded in out_plan[tk]
void orb_deliver_mirror(orb_extractor* e, int tk, const int* h_counts);
bool orb_host_pageable(const void* p);   // ordinary host memory: an "asynchronous" copy out of it returns only when the bytes are staged
int orb_upload_copy_slots(orb_extractor* e, const uint8_t* imgs, int b0, int n, int w, int h, cudaStream_t st);
int orb_ingest_slots(orb_extractor* e, int b0, int n, int w, int h, cudaStream_t st);
int orb_download_slots(orb_extractor* e, int b0, int n, orb_keypoint* kps, uint8_t* desc, int cap, cudaStream_t st, int* h_counts = nullptr);
// NVTX ranges around the stages (host-side enqueue; profilers project them onto the stream).  Header-only NVTX3: no-ops unless a
// tool injects itself.
#include <nvtx3/nvToolsExt.h>
struct OrbNvtxRange {
    explicit OrbNvtxRange(const char* name) { nvtxRangePushA(name); }
    ~OrbNvtxRange() { nvtxRangePop(); }
    OrbNvtxRange(const OrbNvtxRange&) = delete;
    OrbNvtxRange& operator=(const OrbNvtxRange&) = delete;
};
#define ORB_NVTX_CAT2(a, b) a##b
#define ORB_NVTX_CAT(a, b) ORB_NVTX_CAT2(a, b)
#define ORB_NVTX(name) OrbNvtxRange ORB_NVTX_CAT(orb_nvtx_range_, __LINE__)(name)

// device allocations go through guard.cu (ORB_B200_GUARD=1 adds canaries around every buffer; orb_debug_guard_check reads them)
cudaError_t orb_dev_alloc(void** p, size_t bytes);
cudaError_t orb_dev_free(void* p);

#define ORB_CUDA(call)                                                                      \
    do {                                                                                    \
        cudaError_t err__ = (call);                                                         \
        if (err__ != cudaSuccess) {                                                         \
            orb_set_error(std::string(#call) + ": " + cudaGetErrorString(err__));          \
            return ORB_ERR_CUDA;                                                            \
        }                                                                                   \
    } while (0)

// Programmatic dependent launch for calls of a few images (a live camera): the kernels of one image are a chain of launches none of which
// fills the GPU, and the launch-to-launch gap is a large part of each link.  A kernel launched through orb_launch(pdl = true, ...) may be
// scheduled while its predecessor in the stream still runs; ORB_PDL_SYNC() as the FIRST statement of the kernel lets the successor be
// scheduled in turn and then waits until every predecessor has completed and its writes are visible, so nothing is read or written early.
// Without the launch attribute both instructions do nothing.
// Only links whose predecessor in the stream is a kernel of the same chain are launched this way: the first kernel behind a copy or a cross-stream
// event wait (k_ingest, k_orient_describe behind the side-stream blur, k_stereo_rows behind the right handle's stream) is a plain launch.
#define ORB_PDL_SYNC() asm volatile("griddepcontrol.launch_dependents;\n\tgriddepcontrol.wait;" ::: "memory")
inline bool orb_pdl(int batch) {
    static const bool off = getenv("ORB_B200_NO_PDL") != nullptr;
    return batch <= 8 && !off;
}
template <typename... KA, typename... A>
inline cudaError_t orb_launch(bool pdl, void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, A&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = s;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kern, std::forward<A>(args)...);
}

// Popcount Hamming distance of two 256-bit descriptors held as 8 x u32.
__device__ __forceinline__ int orb_hamming256(const uint4 a0, const uint4 a1, const uint4 b0, const uint4 b1) {
    return __popc(a0.x ^ b0.x) + __popc(a0.y ^ b0.y) + __popc(a0.z ^ b0.z) + __popc(a0.w ^ b0.w) +
           __popc(a1.x ^ b1.x) + __popc(a1.y ^ b1.y) + __popc(a1.z ^ b1.z) + __popc(a1.w ^ b1.w);
}
