// Host side of the extractor C ABI (include/orb_b200.h): planning, buffers, launches, copies.
// Mirrors ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:45-114, src/ORBextractor.cc).
#include "orb_internal.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>

void launch_resize(cudaStream_t s, uint8_t* pyr, const CUtensorMap* src_map, bool tma, const LevelGeom& src, const LevelGeom& dst, int b0, int batch);
cudaError_t orb_upload_moment_table(const int* umax);
void launch_ingest(cudaStream_t s, const uint8_t* src, uint8_t* dst, int w, int h, int stride, size_t img_stride, int dst_pitch, size_t dst_img_stride, int batch);
void launch_gauss(cudaStream_t s, const TmaMaps& maps, const uint8_t* pyr, uint8_t* blur, const Plan& P, int b0, int batch);
cudaError_t launch_fast(cudaStream_t s, const TmaMaps& tm, int* cell_cnt, uint32_t* stage, const Plan& P, int b0, int batch);
int fast_box_rows(const Plan& P);
cudaError_t launch_quadtree(cudaStream_t s, const int* cell_cnt, int* cell_prefix, const uint32_t* stage, uint32_t* keys,
                            uint32_t* lvl_kp, int* lvl_cnt, int* lvl_ncand, int* status, const Plan& P, int batch);
size_t quadtree_smem(int cap, int* p2_out);
cudaError_t launch_orient_describe(cudaStream_t s, const TmaMaps& tm_raw, const TmaMaps& tm_blur, const uint32_t* lvl_kp, const int* lvl_cnt,
                                   orb_keypoint* kps, uint8_t* desc, int* counts, const Plan& P, int b0, int batch, bool after_event_wait);

static thread_local std::string g_err;
void orb_set_error(const std::string& s) { g_err = s; }
int orb_sm_count() {
    static int cached[64] = {0};   // 0: not asked yet; plain ints: a racing second query writes the same value
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return 148;
    if (cached[dev] == 0) {
        int n = 0;
        cached[dev] = (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) ? n : 148;
    }
    return cached[dev];
}
extern "C" const char* orb_last_error(void) { return g_err.c_str(); }
extern "C" int orb_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    return n;
}

namespace {

inline int cv_round_f(float v) { return (int)lrintf(v); }
inline int cv_round_d(double v) { return (int)lrint(v); }
inline int cv_floor_d(double v) { int i = (int)v; return i - (i > v); }
inline int cv_ceil_d(double v) { int i = (int)v; return i + (i < v); }

// Scale / feature tables (reference src/ORBextractor.cc:433-488)
void make_tables(orb_extractor& e) {
    const int L = e.nlevels;
    e.sf.assign(L, 1.f); e.sigma2.assign(L, 1.f); e.inv_sf.assign(L, 1.f); e.inv_sigma2.assign(L, 1.f);
    for (int i = 1; i < L; ++i) {
        e.sf[i] = (float)((double)e.sf[i - 1] * e.scale_factor);  // float member * double member
        e.sigma2[i] = e.sf[i] * e.sf[i];
    }
    for (int i = 0; i < L; ++i) { e.inv_sf[i] = 1.0f / e.sf[i]; e.inv_sigma2[i] = 1.0f / e.sigma2[i]; }
    e.n_per_level.assign(L, 0);
    const float factor = (float)(1.0 / e.scale_factor);
    float want = (float)e.nfeatures * (1.f - factor) / (1.f - (float)std::pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l + 1 < L; ++l) {
        e.n_per_level[l] = cv_round_f(want);
        sum += e.n_per_level[l];
        want *= factor;
    }
    e.n_per_level[L - 1] = std::max(e.nfeatures - sum, 0);
    // rows of the radius-15 disc
    int um[ORB_HALF_PATCH + 2] = {0};
    const int vmax = cv_floor_d(ORB_HALF_PATCH * std::sqrt(2.f) / 2 + 1);
    const int vmin = cv_ceil_d(ORB_HALF_PATCH * std::sqrt(2.f) / 2);
    const double hp2 = ORB_HALF_PATCH * ORB_HALF_PATCH;
    for (int v = 0; v <= vmax; ++v) um[v] = cv_round_d(std::sqrt(hp2 - v * v));
    for (int v = ORB_HALF_PATCH, v0 = 0; v >= vmin; --v) {
        while (um[v0] == um[v0 + 1]) ++v0;
        um[v] = v0;
        ++v0;
    }
    for (int v = 0; v < 16; ++v) e.umax[v] = um[v];
}

struct HostTables {
    std::vector<int> xofs, yofs;
    std::vector<short> alpha, beta;
};

// Per-axis INTER_LINEAR source index and 11-bit weights (OpenCV resize, 8u)
void axis_table(int src, int dst, bool clamp_frac, std::vector<int>& ofs, std::vector<short>& w) {
    const double scale = (double)src / dst;
    ofs.resize(dst); w.resize(2 * (size_t)dst);
    for (int d = 0; d < dst; ++d) {
        float f = (float)((d + 0.5) * scale - 0.5);
        int s = cv_floor_d(f);
        f -= s;
        if (clamp_frac) {
            if (s < 0) { f = 0; s = 0; }
            if (s >= src - 1) { f = 0; s = src - 1; }
        }
        ofs[d] = s;
        w[2 * d] = (short)cv_round_f((1.f - f) * 2048);
        w[2 * d + 1] = (short)cv_round_f(f * 2048);
    }
}

// Geometry of every level for a w x h input.  Returns bytes / slots needed per image.
struct Needs { size_t pyr_bytes, cells, stage, tables_bytes; int kp_cap, node_cap_max; };

int build_plan(const orb_extractor& e, int w, int h, Plan& P, Needs& need, std::vector<HostTables>* tabs) {
    std::memset(&P, 0, sizeof(P));
    P.nlevels = e.nlevels; P.w = w; P.h = h; P.ini_th = e.ini_th; P.min_th = e.min_th;
    for (int v = 0; v < 16; ++v) P.umax[v] = e.umax[v];
    size_t off = 0, stage = 0, tbytes = 0;
    int cells = 0, tiles = 0, kp = 0, ncap_max = 0;
    if (tabs) tabs->assign(e.nlevels, HostTables());
    for (int l = 0; l < e.nlevels; ++l) {
        LevelGeom& G = P.lv[l];
        G.w = cv_round_f((float)w * e.inv_sf[l]);
        G.h = cv_round_f((float)h * e.inv_sf[l]);
        if (G.w < 1 || G.h < 1 || G.w > ORB_MAX_DIM || G.h > ORB_MAX_DIM) return ORB_ERR_BAD_ARG;
        G.pitch = (G.w + 63) & ~63;
        G.off = off;
        G.img_stride = (size_t)G.pitch * G.h;
        G.sf = e.sf[l];
        G.patch_size = (float)(int)((float)ORB_PATCH * e.sf[l]);
        G.n_feat = e.n_per_level[l];
        G.max_bx = G.w - ORB_EDGE + 3;
        G.max_by = G.h - ORB_EDGE + 3;
        G.win_w = G.max_bx - ORB_WIN;
        G.win_h = G.max_by - ORB_WIN;
        G.cell_off = cells;
        G.stage_off = stage;
        G.kp_off = kp;
        G.tile_off = tiles;
        const float width = (float)G.win_w, height = (float)G.win_h;
        const int n_cols = (int)(width / 30.f), n_rows = (int)(height / 30.f);
        if (G.win_w > 0 && G.win_h > 0 && n_cols > 0 && n_rows > 0) {
            G.n_cols = n_cols; G.n_rows = n_rows;
            G.w_cell = (int)std::ceil(width / n_cols);
            G.h_cell = (int)std::ceil(height / n_rows);
            G.cell_cap = ((G.w_cell + 1) / 2) * ((G.h_cell + 1) / 2);  // strict local maxima cannot be 8-adjacent
            G.n_ini = (int)std::round((float)G.win_w / (float)G.win_h);
            if (G.n_ini < 1) return ORB_ERR_BAD_ARG;  // reference divides by zero here (aspect < 1:2)
            G.h_x = (float)G.win_w / (float)G.n_ini;
            G.node_cap = std::max(G.n_feat + 3, 4 * G.n_ini) + 1;
        } else {  // level too small for a single cell: the reference's loops do not execute
            G.n_cols = G.n_rows = 0; G.w_cell = G.h_cell = 1; G.cell_cap = 0; G.n_ini = 1; G.h_x = 1.f; G.node_cap = 1;
        }
        G.ncols_magic = (unsigned)((1ull << 32) / (unsigned)std::max(G.n_cols, 1)) + 1u;   // exact for c < 2^32 / n_cols
        G.tiles_x = (G.w + 127) / 128;
        G.tiles_y = (G.h + 63) / 64;
        cells += G.n_cols * G.n_rows;
        stage += (size_t)G.n_cols * G.n_rows * G.cell_cap;
        kp += G.node_cap;
        tiles += G.tiles_x * G.tiles_y;
        off += G.img_stride;  // scaled by batch when laid out (levels are batch-major, see layout())
        ncap_max = std::max(ncap_max, G.node_cap);
        if (l > 0) {
            G.rz_qx = (unsigned)(((unsigned long long)P.lv[l - 1].w << 16) / (unsigned)G.w);
            G.rz_qy = (unsigned)(((unsigned long long)P.lv[l - 1].h << 16) / (unsigned)G.h);
            tbytes += 4 * 64 + (((size_t)G.w + 3) & ~(size_t)3) * 8 + (size_t)G.h * 8 + 64;
            if (tabs) {
                HostTables& T = (*tabs)[l];
                axis_table(P.lv[l - 1].w, G.w, true, T.xofs, T.alpha);
                axis_table(P.lv[l - 1].h, G.h, false, T.yofs, T.beta);
            }
        }
    }
    P.total_cells = cells; P.total_tiles = tiles; P.kp_cap = kp; P.stage_per_image = stage;
    need.pyr_bytes = off; need.cells = (size_t)cells; need.stage = stage; need.kp_cap = kp;
    need.node_cap_max = ncap_max; need.tables_bytes = tbytes + 256;
    return ORB_OK;
}

// Levels are stored batch-major: level l occupies [lvl_base(l) * batch, ...) so that one launch
// walks all images of a level with a fixed stride.
void layout_for_batch(Plan& P, int batch) {
    size_t off = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        P.lv[l].off = off;
        off += P.lv[l].img_stride * (size_t)batch;
    }
}

// TMA descriptors over the level buffers of the current batch layout (orb_internal.h: TmaMaps).  The encoder is a
// driver entry point; it is looked up through the runtime so that the library does not link libcuda.
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn tma_encoder() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (EncodeTiledFn)p;
    }();
    return fn;
}
int build_tma_maps(orb_extractor* e, int batch) {
    EncodeTiledFn enc = tma_encoder();
    if (!enc) { orb_set_error("cuTensorMapEncodeTiled not available from this driver"); return ORB_ERR_CUDA; }
    const Plan& P = e->plan;
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelGeom& G = P.lv[l];
        const cuuint64_t dims[3] = {(cuuint64_t)G.w, (cuuint64_t)G.h, (cuuint64_t)batch};
        const cuuint64_t strides[2] = {(cuuint64_t)G.pitch, (cuuint64_t)G.img_stride};
        const cuuint32_t estr[3] = {1, 1, 1};
        const cuuint32_t box_blur[3] = {ORB_BLUR_BOXW, ORB_BLUR_BOXH, 1}, box_rz[3] = {ORB_RZ_BOXW, ORB_RZ_BOXH, 1};
        const cuuint32_t box_st_l[3] = {32, 11, 1}, box_st_r[3] = {48, 11, 1};
        const cuuint32_t box_fast[3] = {80, (cuuint32_t)fast_box_rows(P), 1};
        const cuuint32_t box_od_raw[3] = {ORB_OD_RAW_BOXW, ORB_OD_RAW_BOXH, 1}, box_od_blur[3] = {ORB_OD_BLUR_BOXW, ORB_OD_BLUR_BOXH, 1};
        void* base = e->d_pyr + G.off;
        CUresult r = enc(&e->tm_blur.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_blur, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_resize.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_rz, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        // the two patches of k_orient_describe: raw level and blurred level (same geometry, other buffer)
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_od_raw.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_od_raw, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_od_blur.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, e->d_blur + G.off, dims, strides, box_od_blur, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        // the cell views of k_fast_cells
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_fast.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_fast, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        // the SAD windows of k_stereo_refine
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_st_l.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_st_l, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r == CUDA_SUCCESS)
            r = enc(&e->tm_st_r.lv[l], CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box_st_r, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { orb_set_error("cuTensorMapEncodeTiled failed (code " + std::to_string((int)r) + ")"); return ORB_ERR_CUDA; }
        // the staged source box of k_resize_tma covers scales up to 1.35 (x) and 1.33 (y); anything else takes the untiled kernel
        e->rz_tma[l] = false;
        if (l > 0) {
            const LevelGeom& S = P.lv[l - 1];
            e->rz_tma[l] = S.w >= G.w && S.h >= G.h && (double)S.w <= 1.35 * G.w && (double)S.h <= 1.33 * G.h;
        }
    }
    return ORB_OK;
}

void free_buffers(orb_extractor* e) {
    orb_dev_free(e->d_pyr); orb_dev_free(e->d_blur); orb_dev_free(e->d_cell_cnt); orb_dev_free(e->d_cell_prefix); orb_dev_free(e->d_stage);
    orb_dev_free(e->d_keys); orb_dev_free(e->d_lvl_kp); orb_dev_free(e->d_lvl_cnt); orb_dev_free(e->d_lvl_ncand); orb_dev_free(e->d_kps);
    orb_dev_free(e->d_desc); orb_dev_free(e->d_counts); orb_dev_free(e->d_status); orb_dev_free(e->d_tables); orb_dev_free(e->d_u_right);
    orb_dev_free(e->d_depth); orb_dev_free(e->d_sad); orb_dev_free(e->d_st_jobs); orb_dev_free(e->d_in); orb_dev_free(e->d_row_start); orb_dev_free(e->d_row_items);
    if (e->ev_done) cudaEventDestroy(e->ev_done);
    if (e->ev_pyr) cudaEventDestroy(e->ev_pyr);
    if (e->ev_blur) cudaEventDestroy(e->ev_blur);
    if (e->s_aux) cudaStreamDestroy(e->s_aux);
    for (int i = 0; i < ORB_N_STAGES + 4; ++i) if (e->ev_stage[i]) cudaEventDestroy(e->ev_stage[i]);
    if (e->h_pin) cudaFreeHost(e->h_pin);
    if (e->h_in) cudaFreeHost(e->h_in);
    if (e->h_out) cudaFreeHost(e->h_out);
    if (e->ev_hin) cudaEventDestroy(e->ev_hin);
    if (e->h_counts) cudaFreeHost(e->h_counts);
    if (e->h_status) cudaFreeHost(e->h_status);
    if (e->h_pyr) cudaFreeHost(e->h_pyr);
    for (int i = 0; i < 2; ++i) if (e->ev_call[i]) cudaEventDestroy(e->ev_call[i]);
    for (cudaEvent_t ev : e->ev_pool) cudaEventDestroy(ev);
    if (e->s_h2d) cudaStreamDestroy(e->s_h2d);
    if (e->s_d2h) cudaStreamDestroy(e->s_d2h);
    if (e->stream) cudaStreamDestroy(e->stream);
}

int ensure_plan(orb_extractor* e, int w, int h, int batch) {
    if (w > e->max_w || h > e->max_h || batch > e->max_batch) {
        orb_set_error("image or batch larger than the capacity given to orb_extractor_create");
        return ORB_ERR_CAPACITY;
    }
    if (e->plan_w == w && e->plan_h == h) {
        if (e->last_batch != batch) {
            layout_for_batch(e->plan, batch);
            const int st = build_tma_maps(e, batch);
            if (st != ORB_OK) { e->plan_w = e->plan_h = -1; return st; }
        }
        e->last_batch = batch;
        return ORB_OK;
    }
    Plan P; Needs need; std::vector<HostTables> tabs;
    const int st = build_plan(*e, w, h, P, need, &tabs);
    if (st != ORB_OK) { orb_set_error("unsupported image geometry"); return st; }
    if (need.pyr_bytes > e->pyr_cap || need.cells > e->cells_cap || need.stage > e->stage_cap || need.kp_cap > e->kp_cap_max ||
        need.tables_bytes > e->tables_cap) {
        orb_set_error("geometry exceeds reserved capacity");
        return ORB_ERR_CAPACITY;
    }
    // upload the resize coefficient tables
    std::vector<uint8_t> blob(need.tables_bytes, 0);
    size_t o = 0;
    for (int l = 1; l < P.nlevels; ++l) {
        LevelGeom& G = P.lv[l];
        const HostTables& T = tabs[l];
        uint8_t* base = (uint8_t*)e->d_tables;
        const size_t w4 = ((size_t)G.w + 3) & ~(size_t)3;   // int4 loads: pad to 4 entries, keep 64-byte alignment
        G.xofs = (const int*)(base + o); std::memcpy(&blob[o], T.xofs.data(), (size_t)G.w * 4); o = (o + w4 * 4 + 63) & ~(size_t)63;
        G.alpha = (const short*)(base + o); std::memcpy(&blob[o], T.alpha.data(), (size_t)G.w * 4); o = (o + w4 * 4 + 63) & ~(size_t)63;
        G.yofs = (const int*)(base + o); std::memcpy(&blob[o], T.yofs.data(), (size_t)G.h * 4); o = (o + (size_t)G.h * 4 + 63) & ~(size_t)63;
        G.beta = (const short*)(base + o); std::memcpy(&blob[o], T.beta.data(), (size_t)G.h * 4); o = (o + (size_t)G.h * 4 + 63) & ~(size_t)63;
    }
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    ORB_CUDA(cudaMemcpyAsync(e->d_tables, blob.data(), o, cudaMemcpyHostToDevice, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    layout_for_batch(P, batch);
    e->plan = P; e->plan_w = w; e->plan_h = h; e->last_batch = batch;
    const int tst = build_tma_maps(e, batch);
    if (tst != ORB_OK) { e->plan_w = e->plan_h = -1; return tst; }
    return ORB_OK;
}

// Runs the extractor stages for image slots [b0, b0 + batch) of the handle's buffers (planned for
// e->last_batch images) on e->stream.  Slot addressing is done on the host: level offsets and the
// per-image arrays are shifted, the kernels always see "their" batch starting at 0.
// phase 0: everything; 1: the pyramid (and the blur when it runs on the side stream) only; 2: the rest.  A caller that feeds two handles
// from one thread queues phase 1 of both before phase 2 of either (orb_stereo_frames_enqueue, one-chunk calls).
int run_pipeline(orb_extractor* e, int b0, int batch, int phase = 0) {
    Plan P = e->plan;
    for (int l = 0; l < P.nlevels; ++l) P.lv[l].off += (unsigned long long)b0 * P.lv[l].img_stride;
    const Plan& Q = e->plan;
    int* cell_cnt = e->d_cell_cnt + (size_t)b0 * Q.total_cells;
    int* cell_prefix = e->d_cell_prefix + (size_t)b0 * Q.total_cells;
    uint32_t* stage = e->d_stage + (size_t)b0 * Q.stage_per_image;
    uint32_t* keys = e->d_keys + (size_t)b0 * 2 * Q.stage_per_image;
    uint32_t* lvl_kp = e->d_lvl_kp + (size_t)b0 * Q.kp_cap;
    int* lvl_cnt = e->d_lvl_cnt + (size_t)b0 * ORB_MAX_LEVELS;
    int* lvl_ncand = e->d_lvl_ncand + (size_t)b0 * ORB_MAX_LEVELS;
    orb_keypoint* kps = e->d_kps + (size_t)b0 * Q.kp_cap;
    uint8_t* desc = e->d_desc + (size_t)b0 * Q.kp_cap * 32;
    int* counts = e->d_counts + b0;
    cudaStream_t s = e->stream;
    const bool T = e->timing;
    // event i is recorded BEFORE stage i; event 5 after the last extractor stage
    const bool first = phase != 2, second = phase != 1;
    if (first) {
    if (T) cudaEventRecord(e->ev_stage[0], s);
    nvtxRangePushA("orb:pyramid");
    for (int l = 1; l < P.nlevels; ++l) { launch_resize(s, e->d_pyr, &e->tm_resize.lv[l - 1], e->rz_tma[l], Q.lv[l - 1], Q.lv[l], b0, batch); e->launches++; }
    nvtxRangePop();
    if (T) cudaEventRecord(e->ev_stage[1], s);
    }
    // The blur depends on the pyramid only: for a small batch (a live camera: the kernels cannot fill the GPU) it runs on a side stream
    // beside FAST and the quadtree, and orient + describe waits for both.  With a large batch the two handles of a stereo pair already
    // keep every SM busy and the side stream was measured to change nothing (74.4 k vs 74.8 k frames/s), so it is not used there.
    const bool side = !T && e->s_aux && batch * P.nlevels <= 2 * orb_sm_count() && !getenv("ORB_B200_NO_SIDE_BLUR");
    if (side && first) {
        ORB_CUDA(cudaEventRecord(e->ev_pyr, s));
        ORB_CUDA(cudaStreamWaitEvent(e->s_aux, e->ev_pyr, 0));
        { ORB_NVTX("orb:gauss7"); launch_gauss(e->s_aux, e->tm_blur, e->d_pyr, e->d_blur, Q, b0, batch); e->launches++; }
        ORB_CUDA(cudaEventRecord(e->ev_blur, e->s_aux));
    }
    if (!second) { ORB_CUDA(cudaGetLastError()); return ORB_OK; }
    if (P.total_cells > 0) {
        ORB_NVTX("orb:fast_cells");
        ORB_CUDA(launch_fast(s, e->tm_fast, cell_cnt, stage, Q, b0, batch)); e->launches++;
    }
    if (T) cudaEventRecord(e->ev_stage[2], s);
    {
        ORB_NVTX("orb:quadtree");
        ORB_CUDA(launch_quadtree(s, cell_cnt, cell_prefix, stage, keys, lvl_kp, lvl_cnt, lvl_ncand, e->d_status, P, batch));
        e->launches++;
    }
    if (T) cudaEventRecord(e->ev_stage[3], s);
    if (side) ORB_CUDA(cudaStreamWaitEvent(s, e->ev_blur, 0));
    else { ORB_NVTX("orb:gauss7"); launch_gauss(s, e->tm_blur, e->d_pyr, e->d_blur, Q, b0, batch); e->launches++; }
    if (T) cudaEventRecord(e->ev_stage[4], s);
    {
        ORB_NVTX("orb:orient_describe");
        ORB_CUDA(launch_orient_describe(s, e->tm_od_raw, e->tm_od_blur, lvl_kp, lvl_cnt, kps, desc, counts, Q, b0, batch, side));
        e->launches++;
    }
    if (T) { cudaEventRecord(e->ev_stage[5], s); for (int i = 0; i < 5; ++i) e->stage_valid[i] = true; }
    ORB_CUDA(cudaGetLastError());
    return ORB_OK;
}
int run_pipeline(orb_extractor* e, int batch) { return run_pipeline(e, 0, batch); }

int check_status(orb_extractor* e) {
    int st = 0;
    ORB_CUDA(cudaMemcpyAsync(&st, e->d_status, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    if (st != 0) {
        cudaMemsetAsync(e->d_status, 0, sizeof(int), e->stream);
        orb_set_error("device-side bound exceeded");
        return st;
    }
    return ORB_OK;
}

}  // namespace

// ---- slot-wise building blocks of the host-buffer entry points (shared with stereo.cu) ----
int orb_ensure_plan(orb_extractor* e, int w, int h, int batch) { return ensure_plan(e, w, h, batch); }
int orb_run_slots(orb_extractor* e, int b0, int n, int phase) { return run_pipeline(e, b0, n, phase); }

int orb_pipe_streams(orb_extractor* e) {
    if (!e->s_h2d) ORB_CUDA(cudaStreamCreateWithFlags(&e->s_h2d, cudaStreamNonBlocking));
    if (!e->s_d2h) ORB_CUDA(cudaStreamCreateWithFlags(&e->s_d2h, cudaStreamNonBlocking));
    if (!e->d_in && !getenv("ORB_B200_NO_STAGING"))
        ORB_CUDA(orb_dev_alloc((void**)&e->d_in, (size_t)e->max_batch * e->max_w * e->max_h + 64));
    if (e->d_in && !e->h_in && !getenv("ORB_B200_NO_PINNED_MIRROR")) {
        const size_t one = (size_t)e->max_w * e->max_h, all = one * e->max_batch, lim = (size_t)16 << 20;
        const size_t cap = all <= lim ? all : (one <= lim ? (lim / one) * one : 0);
        if (cap) {
            ORB_CUDA(cudaHostAlloc((void**)&e->h_in, cap, cudaHostAllocDefault));
            ORB_CUDA(cudaEventCreateWithFlags(&e->ev_hin, cudaEventDisableTiming));
            e->h_in_cap = cap;
        }
    }
    return ORB_OK;
}
bool orb_host_pageable(const void* p) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return true; }
    return a.type == cudaMemoryTypeUnregistered;
}
// ---- pinned mirror of a small call's results (see orb_internal.h) ----
static size_t out_section(size_t bytes) { return (bytes + 255) & ~(size_t)255; }
static void out_offsets(const orb_extractor* e, int n, size_t* desc, size_t* ur, size_t* dp, size_t* total) {
    const size_t kc = (size_t)e->plan.kp_cap;
    *desc = out_section((size_t)n * kc * sizeof(orb_keypoint));
    *ur = *desc + out_section((size_t)n * kc * 32);
    *dp = *ur + out_section((size_t)n * kc * 4);
    *total = *dp + out_section((size_t)n * kc * 4);
}
bool orb_out_mirror_ready(orb_extractor* e, int n, bool with_stereo) {
    (void)with_stereo;
    if (getenv("ORB_B200_NO_PINNED_MIRROR")) return false;
    size_t o_desc, o_ur, o_dp, total;
    out_offsets(e, n, &o_desc, &o_ur, &o_dp, &total);
    if (total > ((size_t)8 << 20)) return false;
    if (e->h_out_cap < 2 * total) {
        if (e->inflight > 0) return false;   // a half may still be in use
        if (e->h_out) { cudaFreeHost(e->h_out); e->h_out = nullptr; e->h_out_cap = 0; }
        if (cudaHostAlloc((void**)&e->h_out, 2 * total, cudaHostAllocDefault) != cudaSuccess) { cudaGetLastError(); e->h_out = nullptr; return false; }
        e->h_out_cap = 2 * total;
    }
    return true;
}
int orb_download_mirror(orb_extractor* e, int n, int tk, cudaStream_t st, int* h_counts) {
    size_t o_desc, o_ur, o_dp, total;
    out_offsets(e, n, &o_desc, &o_ur, &o_dp, &total);
    uint8_t* base = e->h_out + (size_t)tk * (e->h_out_cap / 2);
    const size_t kc = (size_t)e->plan.kp_cap;
    if (!h_counts) h_counts = e->h_counts;
    ORB_CUDA(cudaMemcpyAsync(h_counts, e->d_counts, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
    ORB_CUDA(cudaMemcpyAsync(base, e->d_kps, (size_t)n * kc * sizeof(orb_keypoint), cudaMemcpyDeviceToHost, st));
    ORB_CUDA(cudaMemcpyAsync(base + o_desc, e->d_desc, (size_t)n * kc * 32, cudaMemcpyDeviceToHost, st));
    return ORB_OK;
}
int orb_download_mirror_stereo(orb_extractor* e, int n, int tk, cudaStream_t st) {
    size_t o_desc, o_ur, o_dp, total;
    out_offsets(e, n, &o_desc, &o_ur, &o_dp, &total);
    uint8_t* base = e->h_out + (size_t)tk * (e->h_out_cap / 2);
    const size_t kc = (size_t)e->plan.kp_cap;
    ORB_CUDA(cudaMemcpyAsync(base + o_ur, e->d_u_right, (size_t)n * kc * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA(cudaMemcpyAsync(base + o_dp, e->d_depth, (size_t)n * kc * 4, cudaMemcpyDeviceToHost, st));
    return ORB_OK;
}
// uRight / depth of ONE image of the last batch -> half 0 of the mirror (layout of a one-image call); host addresses in *h_ur, *h_dp
int orb_download_mirror_stereo_row(orb_extractor* e, int image, cudaStream_t st, const float** h_ur, const float** h_dp) {
    size_t o_desc, o_ur, o_dp, total;
    out_offsets(e, 1, &o_desc, &o_ur, &o_dp, &total);
    const size_t kc = (size_t)e->plan.kp_cap;
    ORB_CUDA(cudaMemcpyAsync(e->h_out + o_ur, e->d_u_right + (size_t)image * kc, kc * 4, cudaMemcpyDeviceToHost, st));
    ORB_CUDA(cudaMemcpyAsync(e->h_out + o_dp, e->d_depth + (size_t)image * kc, kc * 4, cudaMemcpyDeviceToHost, st));
    *h_ur = reinterpret_cast<const float*>(e->h_out + o_ur);
    *h_dp = reinterpret_cast<const float*>(e->h_out + o_dp);
    return ORB_OK;
}
void orb_deliver_mirror(orb_extractor* e, int tk, const int* h_counts) {
    orb_extractor::OutPlan& O = e->out_plan[tk];
    if (!O.on) return;
    O.on = false;
    size_t o_desc, o_ur, o_dp, total;
    out_offsets(e, O.n, &o_desc, &o_ur, &o_dp, &total);
    const uint8_t* base = e->h_out + (size_t)tk * (e->h_out_cap / 2);
    const size_t kc = (size_t)e->plan.kp_cap;
    for (int b = 0; b < O.n; ++b) {
        const size_t m = (size_t)std::max(0, std::min(std::min(h_counts[b], O.cap), (int)kc));
        if (O.kps) std::memcpy(O.kps + (size_t)b * O.cap, base + (size_t)b * kc * sizeof(orb_keypoint), m * sizeof(orb_keypoint));
        if (O.desc) std::memcpy(O.desc + (size_t)b * O.cap * 32, base + o_desc + (size_t)b * kc * 32, m * 32);
        if (O.u_right) std::memcpy(O.u_right + (size_t)b * O.cap, base + o_ur + (size_t)b * kc * 4, m * 4);
        if (O.depth) std::memcpy(O.depth + (size_t)b * O.cap, base + o_dp + (size_t)b * kc * 4, m * 4);
    }
}

// Images [0, n) at src (pageable host memory, any row stride) -> dense at d_dst through the pinned mirror: CPU copy in pieces, each piece
// queued on `st` as soon as it is in place, so that the link works while the next piece is copied.  *done = false: not applicable
// (no mirror, too large, or the source is pinned / device memory), nothing was queued.
static int upload_through_mirror(orb_extractor* e, const uint8_t* src, int n, int w, int h, int stride, size_t img_stride, uint8_t* d_dst,
                                 cudaStream_t st, bool* done) {
    *done = false;
    const size_t bytes = (size_t)n * w * h;
    if (!e->h_in || bytes == 0 || bytes > e->h_in_cap || !orb_host_pageable(src)) return ORB_OK;
    if (e->hin_used) ORB_CUDA(cudaEventSynchronize(e->ev_hin));   // the previous copy out of the mirror (normally long done)
    const size_t piece = (size_t)192 << 10;
    const bool dense = stride == w && img_stride == (size_t)w * h;
    size_t sent = 0, filled = 0;
    int b = 0, y = 0;
    while (sent < bytes) {
        const size_t upto = std::min(bytes, sent + piece);
        if (dense) { std::memcpy(e->h_in + filled, src + filled, upto - filled); filled = upto; }
        else {
            while (filled < upto) {   // whole rows; a piece ends at the first row boundary at or beyond `upto`
                std::memcpy(e->h_in + filled, src + (size_t)b * img_stride + (size_t)y * stride, (size_t)w);
                filled += (size_t)w;
                if (++y == h) { y = 0; ++b; }
            }
        }
        ORB_CUDA(cudaMemcpyAsync(d_dst + sent, e->h_in + sent, filled - sent, cudaMemcpyHostToDevice, st));
        sent = filled;
    }
    ORB_CUDA(cudaEventRecord(e->ev_hin, st));
    e->hin_used = true;
    *done = true;
    return ORB_OK;
}
cudaEvent_t orb_pool_event(orb_extractor* e, int i) {
    while ((int)e->ev_pool.size() <= i) {
        cudaEvent_t ev = nullptr;
        if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) != cudaSuccess) return nullptr;
        e->ev_pool.push_back(ev);
    }
    return e->ev_pool[i];
}
// Chunks are large enough to fill the GPU several times over and small enough that a batch has
// ~8 of them in flight between the upload, kernel and download streams.
int orb_chunk_size(int batch) {
    if (const char* v = getenv("ORB_B200_CHUNK")) { const int c = atoi(v); if (c > 0) return c < batch ? c : batch; }
    if (batch <= 8) return batch;
    // 4 chunks: measured on B200 at batch 256 (chunk 16 / 24 / 32 / 48 / 64 / 128 -> 25.0 / 26.4 / 28.6 / 29.7 / 29.8 / 27.6 k frames/s
    // end to end): small chunks leave the per-(level, image) quadtree CTAs too few to fill the GPU, large ones lengthen fill and drain
    const int c = (batch + 3) / 4;
    return c < 8 ? 8 : c;
}
// Chunks of the upload / kernels / download pipeline: uniform.  (A ramp that starts with a quarter-size chunk and grows
// 1.5x per chunk, to shorten the unhidden first upload, measured 29.2-29.4 k frames/s against 29.7-29.8 k for four equal
// chunks at batch 256 on B200: the small chunks cost more in kernel efficiency than the shorter fill gains.)
std::vector<int> orb_chunk_bounds(int batch) {
    std::vector<int> b(1, 0);
    const int c = orb_chunk_size(batch);
    while (b.back() < batch) b.push_back(std::min(batch, b.back() + c));
    return b;
}
// Host images [b0, b0 + n) of a batch -> level 0 of the same slots.
int orb_upload_slots(orb_extractor* e, const uint8_t* imgs, int b0, int n, int w, int h, int stride, size_t img_stride, cudaStream_t st) {
    ORB_NVTX("orb:upload_ingest");
    const LevelGeom& G0 = e->plan.lv[0];
    uint8_t* dst = e->d_pyr + G0.off + (size_t)b0 * G0.img_stride;
    const uint8_t* src = imgs + (size_t)b0 * img_stride;
    if (e->d_in) {
        uint8_t* stg = e->d_in + (size_t)b0 * w * h;
        bool done = false;
        const int st_m = upload_through_mirror(e, src, n, w, h, stride, img_stride, stg, st, &done);
        if (st_m != ORB_OK) return st_m;
        if (done) {
            launch_ingest(st, stg, dst, w, h, w, (size_t)w * h, G0.pitch, G0.img_stride, n);
            e->launches++;
            return ORB_OK;
        }
    }
    if (img_stride == (size_t)stride * h && stride == w && e->d_in) {
        // dense host batch: one linear copy at full PCIe rate into the staging area, then the ingest kernel
        // re-pitches it on the device (strided copies of w-byte rows are DMA-descriptor bound, host or device)
        uint8_t* stg = e->d_in + (size_t)b0 * w * h;
        ORB_CUDA(cudaMemcpyAsync(stg, src, (size_t)n * w * h, cudaMemcpyHostToDevice, st));
        launch_ingest(st, stg, dst, w, h, w, (size_t)w * h, G0.pitch, G0.img_stride, n);
        e->launches++;
    } else if (img_stride == (size_t)stride * h && G0.img_stride == (unsigned long long)G0.pitch * h) {
        ORB_CUDA(cudaMemcpy2DAsync(dst, G0.pitch, src, stride, w, (size_t)h * n, cudaMemcpyHostToDevice, st));
    } else {
        for (int b = 0; b < n; ++b)
            ORB_CUDA(cudaMemcpy2DAsync(dst + (size_t)b * G0.img_stride, G0.pitch, src + (size_t)b * img_stride, stride, w, h, cudaMemcpyHostToDevice, st));
    }
    return ORB_OK;
}
// The same in two steps for pipelines that keep the copy stream free of kernels (orb_stereo_frames_enqueue): the dense host batch
// goes to the staging area on the copy stream; the ingest kernel that re-pitches it into level 0 runs on the handle's kernel
// stream.  The copy then only depends on the PREVIOUS ingest of the same staging slots, not on every kernel that still reads level 0,
// so uploads run a whole call ahead.  Returns 1 when the split form applies (dense images and a staging area), 0 otherwise.
int orb_upload_can_split(const orb_extractor* e, int w, int h, int stride, size_t img_stride) {
    return img_stride == (size_t)stride * h && stride == w && e->d_in != nullptr;
}
int orb_upload_copy_slots(orb_extractor* e, const uint8_t* imgs, int b0, int n, int w, int h, cudaStream_t st) {
    bool done = false;
    const int st_m = upload_through_mirror(e, imgs + (size_t)b0 * w * h, n, w, h, w, (size_t)w * h, e->d_in + (size_t)b0 * w * h, st, &done);
    if (st_m != ORB_OK || done) return st_m;
    ORB_CUDA(cudaMemcpyAsync(e->d_in + (size_t)b0 * w * h, imgs + (size_t)b0 * w * h, (size_t)n * w * h, cudaMemcpyHostToDevice, st));
    return ORB_OK;
}
int orb_ingest_slots(orb_extractor* e, int b0, int n, int w, int h, cudaStream_t st) {
    ORB_NVTX("orb:ingest");
    const LevelGeom& G0 = e->plan.lv[0];
    launch_ingest(st, e->d_in + (size_t)b0 * w * h, e->d_pyr + G0.off + (size_t)b0 * G0.img_stride, w, h, w, (size_t)w * h, G0.pitch, G0.img_stride, n);
    e->launches++;
    return ORB_OK;
}
// Results of slots [b0, b0 + n) -> the caller's [batch][cap] arrays; counts land in e->h_counts.
// Whole slot rows are copied (one strided copy per array): the tail beyond counts[b] is unspecified.
int orb_download_slots(orb_extractor* e, int b0, int n, orb_keypoint* kps, uint8_t* desc, int cap, cudaStream_t st, int* h_counts) {
    const int kc = e->plan.kp_cap;
    const int m = cap < kc ? cap : kc;
    if (!h_counts) h_counts = e->h_counts;
    ORB_CUDA(cudaMemcpyAsync(h_counts + b0, e->d_counts + b0, (size_t)n * sizeof(int), cudaMemcpyDeviceToHost, st));
    if (m > 0) {
        ORB_CUDA(cudaMemcpy2DAsync(kps + (size_t)b0 * cap, (size_t)cap * sizeof(orb_keypoint), e->d_kps + (size_t)b0 * kc, (size_t)kc * sizeof(orb_keypoint),
                                   (size_t)m * sizeof(orb_keypoint), n, cudaMemcpyDeviceToHost, st));
        ORB_CUDA(cudaMemcpy2DAsync(desc + (size_t)b0 * cap * 32, (size_t)cap * 32, e->d_desc + (size_t)b0 * kc * 32, (size_t)kc * 32, (size_t)m * 32, n,
                                   cudaMemcpyDeviceToHost, st));
    }
    return ORB_OK;
}

extern "C" {

int orb_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th, int max_w, int max_h,
                         int max_batch, int device, orb_extractor** out) {
    if (!out) return ORB_ERR_BAD_ARG;
    *out = nullptr;
    if (nfeatures < 1 || nlevels < 1 || nlevels > ORB_MAX_LEVELS || !(scale_factor > 1.0f) || ini_th < 1 || min_th < 1 ||
        ini_th > 254 || min_th > 254 || max_w < 1 || max_h < 1 || max_batch < 1) {
        orb_set_error("bad extractor parameters");
        return ORB_ERR_BAD_ARG;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        orb_set_error("no CUDA device: this library has no CPU fallback");
        return ORB_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { orb_set_error("bad device index"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(device));
    orb_extractor* e = new orb_extractor();
    e->nfeatures = nfeatures; e->scale_factor = (double)scale_factor; e->nlevels = nlevels; e->ini_th = ini_th; e->min_th = min_th;
    e->max_w = max_w; e->max_h = max_h; e->max_batch = max_batch; e->device = device;
    make_tables(*e);
    if (orb_upload_moment_table(e->umax) != cudaSuccess) { delete e; orb_set_error("moment table upload"); return ORB_ERR_CUDA; }
    Plan P; Needs need;
    int st = build_plan(*e, max_w, max_h, P, need, nullptr);
    if (st != ORB_OK) { delete e; orb_set_error("unsupported maximum geometry (level size must be <= 4096, aspect >= 1:2)"); return st; }
    int p2;
    if (quadtree_smem(need.node_cap_max, &p2) > 200 * 1024) {
        delete e; orb_set_error("nfeatures too large for the on-chip quadtree"); return ORB_ERR_BAD_ARG;
    }
    // Smaller images of a different aspect can need slightly more of some resources; keep slack.
    e->pyr_cap = need.pyr_bytes + 4096;
    e->cells_cap = need.cells + need.cells / 8 + 64;
    e->stage_cap = need.stage + need.stage / 8 + 1024;
    e->kp_cap_max = need.kp_cap + 64;
    e->tables_cap = need.tables_bytes + 4096;
    const size_t B = (size_t)max_batch;
#define ALLOC(ptr, bytes) do { cudaError_t er = orb_dev_alloc((void**)&(ptr), (bytes)); if (er != cudaSuccess) { \
        orb_set_error(std::string("cudaMalloc: ") + cudaGetErrorString(er)); free_buffers(e); delete e; return ORB_ERR_CUDA; } } while (0)
    if (cudaStreamCreateWithFlags(&e->stream, cudaStreamNonBlocking) != cudaSuccess) { delete e; orb_set_error("stream"); return ORB_ERR_CUDA; }
    ALLOC(e->d_pyr, e->pyr_cap * B);
    ALLOC(e->d_blur, e->pyr_cap * B);
    ALLOC(e->d_cell_cnt, e->cells_cap * B * sizeof(int));
    ALLOC(e->d_cell_prefix, e->cells_cap * B * sizeof(int));
    ALLOC(e->d_stage, e->stage_cap * B * sizeof(uint32_t));
    ALLOC(e->d_keys, e->stage_cap * B * 2 * sizeof(uint32_t));
    ALLOC(e->d_lvl_kp, (size_t)e->kp_cap_max * B * sizeof(uint32_t));
    ALLOC(e->d_lvl_cnt, B * ORB_MAX_LEVELS * sizeof(int));
    ALLOC(e->d_lvl_ncand, B * ORB_MAX_LEVELS * sizeof(int));
    ALLOC(e->d_kps, (size_t)e->kp_cap_max * B * sizeof(orb_keypoint));
    ALLOC(e->d_desc, (size_t)e->kp_cap_max * B * 32);
    ALLOC(e->d_counts, B * sizeof(int));
    ALLOC(e->d_status, sizeof(int));
    ALLOC(e->d_tables, e->tables_cap);
    ALLOC(e->d_u_right, (size_t)e->kp_cap_max * B * sizeof(float));
    ALLOC(e->d_depth, (size_t)e->kp_cap_max * B * sizeof(float));
    ALLOC(e->d_sad, (size_t)e->kp_cap_max * B * sizeof(int));
    ALLOC(e->d_st_jobs, (size_t)e->kp_cap_max * B * sizeof(int2));
    e->row_items_cap = e->kp_cap_max * (2 * (int)std::ceil(2.0f * e->sf[nlevels - 1]) + 3);
    ALLOC(e->d_row_start, B * (ORB_MAX_DIM + 1) * sizeof(int));
    ALLOC(e->d_row_items, B * (size_t)e->row_items_cap * sizeof(int2));
    if (cudaEventCreateWithFlags(&e->ev_done, cudaEventDisableTiming) != cudaSuccess) { free_buffers(e); delete e; orb_set_error("event"); return ORB_ERR_CUDA; }
    if (cudaStreamCreateWithFlags(&e->s_aux, cudaStreamNonBlocking) != cudaSuccess || cudaEventCreateWithFlags(&e->ev_pyr, cudaEventDisableTiming) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_blur, cudaEventDisableTiming) != cudaSuccess) { free_buffers(e); delete e; orb_set_error("side stream"); return ORB_ERR_CUDA; }
#undef ALLOC
    cudaMemset(e->d_status, 0, sizeof(int));
    cudaMemset(e->d_lvl_cnt, 0, B * ORB_MAX_LEVELS * sizeof(int));
    cudaMemset(e->d_counts, 0, B * sizeof(int));
    cudaMemset(e->d_pyr, 0, e->pyr_cap * B);
    cudaMemset(e->d_blur, 0, e->pyr_cap * B);
    if (cudaMallocHost((void**)&e->h_counts, B * 2 * sizeof(int)) != cudaSuccess || cudaMallocHost((void**)&e->h_status, 4 * sizeof(int)) != cudaSuccess ||
        cudaEventCreateWithFlags(&e->ev_call[0], cudaEventDisableTiming) != cudaSuccess || cudaEventCreateWithFlags(&e->ev_call[1], cudaEventDisableTiming) != cudaSuccess) {
        free_buffers(e); delete e; orb_set_error("pinned alloc"); return ORB_ERR_CUDA;
    }
    *out = e;
    return ORB_OK;
}

void orb_extractor_destroy(orb_extractor* e) {
    if (!e) return;
    cudaSetDevice(e->device);
    cudaStreamSynchronize(e->stream);
    free_buffers(e);
    delete e;
}

int orb_extractor_tables(const orb_extractor* e, float* sf, float* inv_sf, float* s2, float* inv_s2, int32_t* npl, int32_t* umax) {
    if (!e) return ORB_ERR_BAD_ARG;
    for (int i = 0; i < e->nlevels; ++i) {
        if (sf) sf[i] = e->sf[i];
        if (inv_sf) inv_sf[i] = e->inv_sf[i];
        if (s2) s2[i] = e->sigma2[i];
        if (inv_s2) inv_s2[i] = e->inv_sigma2[i];
        if (npl) npl[i] = e->n_per_level[i];
    }
    if (umax) for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
    return ORB_OK;
}

int orb_extractor_set_timing(orb_extractor* e, int on) {
    if (!e) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    if (on && !e->ev_stage[0])
        for (int i = 0; i < ORB_N_STAGES + 4; ++i) ORB_CUDA(cudaEventCreate(&e->ev_stage[i]));
    e->timing = on != 0;
    for (int i = 0; i < ORB_N_STAGES; ++i) e->stage_valid[i] = false;
    return ORB_OK;
}

int orb_extractor_stage_ms(orb_extractor* e, float* ms) {
    if (!e || !ms) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    for (int i = 0; i < ORB_N_STAGES; ++i) {
        ms[i] = 0.f;
        if (!e->stage_valid[i]) continue;
        // extractor stages use events 0..5, stereo stages events 6..9
        const int a = i < 5 ? i : i + 1, b = a + 1;
        ORB_CUDA(cudaEventElapsedTime(&ms[i], e->ev_stage[a], e->ev_stage[b]));
    }
    return ORB_OK;
}

int orb_extractor_max_keypoints(const orb_extractor* e) { return e ? e->kp_cap_max : 0; }
void* orb_extractor_stream(const orb_extractor* e) { return e ? (void*)e->stream : nullptr; }
long long orb_extractor_launch_count(const orb_extractor* e) { return e ? e->launches : 0; }

int orb_extractor_sync(orb_extractor* e) {
    if (!e) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    return ORB_OK;
}

int orb_extract_device(orb_extractor* e, const uint8_t* d_imgs, int batch, int w, int h, int stride, size_t img_stride) {
    if (!e || !d_imgs || batch < 1 || w < 1 || h < 1 || stride < w) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    int st = ensure_plan(e, w, h, batch);
    if (st != ORB_OK) return st;
    const LevelGeom& G0 = e->plan.lv[0];
    launch_ingest(e->stream, d_imgs, e->d_pyr + G0.off, w, h, stride, img_stride, G0.pitch, G0.img_stride, batch);
    e->launches++;
    return run_pipeline(e, batch);
}

int orb_extractor_device_results(const orb_extractor* e, const orb_keypoint** d_kps, const uint8_t** d_desc, const int32_t** d_counts,
                                 int* kp_cap) {
    if (!e) return ORB_ERR_BAD_ARG;
    if (d_kps) *d_kps = e->d_kps;
    if (d_desc) *d_desc = e->d_desc;
    if (d_counts) *d_counts = e->d_counts;
    if (kp_cap) *kp_cap = e->plan.kp_cap;
    return ORB_OK;
}

int orb_extractor_download(orb_extractor* e, int batch, orb_keypoint* kps, uint8_t* desc, int cap, int32_t* counts) {
    if (!e || !kps || !desc || !counts || batch < 1 || batch > e->last_batch) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    ORB_CUDA(cudaMemcpyAsync(e->h_counts, e->d_counts, (size_t)batch * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    int st = check_status(e);
    if (st != ORB_OK) return st;
    const int kc = e->plan.kp_cap;
    int rc = ORB_OK;
    for (int b = 0; b < batch; ++b) {
        const int n = e->h_counts[b];
        counts[b] = n;
        if (n > cap) { rc = ORB_ERR_CAPACITY; continue; }
        if (n == 0) continue;
        ORB_CUDA(cudaMemcpyAsync(kps + (size_t)b * cap, e->d_kps + (size_t)b * kc, (size_t)n * sizeof(orb_keypoint), cudaMemcpyDeviceToHost, e->stream));
        ORB_CUDA(cudaMemcpyAsync(desc + (size_t)b * cap * 32, e->d_desc + (size_t)b * kc * 32, (size_t)n * 32, cudaMemcpyDeviceToHost, e->stream));
    }
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    if (rc != ORB_OK) orb_set_error("keypoint capacity too small");
    return rc;
}

int orb_extract_batch(orb_extractor* e, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride, orb_keypoint* kps,
                      uint8_t* desc, int cap, int32_t* counts) {
    if (!e || !counts || batch < 0) return ORB_ERR_BAD_ARG;
    if (batch == 0) return ORB_OK;
    if (w == 0 || h == 0) {  // _image.empty() -> return (reference :1095)
        for (int b = 0; b < batch; ++b) counts[b] = 0;
        return ORB_OK;
    }
    if (!imgs || !kps || !desc || w < 0 || h < 0 || stride < w) return ORB_ERR_BAD_ARG;
    if (e->inflight > 0) { orb_set_error("an enqueued batch is still in flight on this handle"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(e->device));
    int st = ensure_plan(e, w, h, batch);
    if (st != ORB_OK) return st;
    st = orb_pipe_streams(e);
    if (st != ORB_OK) return st;
    const std::vector<int> cb = orb_chunk_bounds(batch);
    const int nch = (int)cb.size() - 1;
    if (!orb_pool_event(e, 2 * nch)) { orb_set_error("event"); return ORB_ERR_CUDA; }
    // uploads must not overtake work already queued on the kernel stream
    ORB_CUDA(cudaEventRecord(e->ev_pool[2 * nch], e->stream));
    ORB_CUDA(cudaStreamWaitEvent(e->s_h2d, e->ev_pool[2 * nch], 0));
    // on an error after the first enqueue nothing may stay in flight: the caller is free to release its buffers
    auto fail = [&](int code) { cudaStreamSynchronize(e->s_h2d); cudaStreamSynchronize(e->stream); cudaStreamSynchronize(e->s_d2h); e->out_plan[0].on = false; return code; };
    const bool mirror = nch == 1 && orb_host_pageable(kps) && orb_host_pageable(desc) && orb_out_mirror_ready(e, batch, false);
    for (int c = 0; c < nch; ++c) {
        const int b0 = cb[c], n = cb[c + 1] - cb[c];
        st = orb_upload_slots(e, imgs, b0, n, w, h, stride, img_stride, e->s_h2d);
        if (st != ORB_OK) return fail(st);
        if (cudaEventRecord(e->ev_pool[2 * c], e->s_h2d) != cudaSuccess || cudaStreamWaitEvent(e->stream, e->ev_pool[2 * c], 0) != cudaSuccess) return fail(ORB_ERR_CUDA);
        st = e->copy_only ? ORB_OK : run_pipeline(e, b0, n);
        if (st != ORB_OK) return fail(st);
        if (cudaEventRecord(e->ev_pool[2 * c + 1], e->stream) != cudaSuccess || cudaStreamWaitEvent(e->s_d2h, e->ev_pool[2 * c + 1], 0) != cudaSuccess) return fail(ORB_ERR_CUDA);
        if (mirror) {
            st = orb_download_mirror(e, n, 0, e->s_d2h, nullptr);
            orb_extractor::OutPlan& O = e->out_plan[0];
            O.on = st == ORB_OK; O.n = n; O.cap = cap; O.kps = kps; O.desc = desc; O.u_right = nullptr; O.depth = nullptr;
        } else {
            st = orb_download_slots(e, b0, n, kps, desc, cap, e->s_d2h);
        }
        if (st != ORB_OK) return fail(st);
    }
    if (!e->h_status) ORB_CUDA(cudaMallocHost((void**)&e->h_status, 4 * sizeof(int)));
    ORB_CUDA(cudaMemcpyAsync(&e->h_status[0], e->d_status, sizeof(int), cudaMemcpyDeviceToHost, e->s_d2h));
    ORB_CUDA(cudaStreamSynchronize(e->s_d2h));
    const int dst = e->h_status[0];
    if (dst != 0) { e->out_plan[0].on = false; cudaMemsetAsync(e->d_status, 0, sizeof(int), e->stream); orb_set_error("device-side bound exceeded"); return dst; }
    if (mirror) orb_deliver_mirror(e, 0, e->h_counts);
    int rc = ORB_OK;
    for (int b = 0; b < batch; ++b) { counts[b] = e->h_counts[b]; if (counts[b] > cap) rc = ORB_ERR_CAPACITY; }
    if (rc != ORB_OK) orb_set_error("keypoint capacity too small");
    return rc;
}

// Streaming form of orb_extract_batch (mirror of orb_stereo_frames_enqueue / _wait for one handle): enqueue() queues uploads, kernels
// and downloads of the whole batch on the handle's three streams and returns; wait() blocks until that call's outputs are in the
// caller's buffers.  Two calls may be in flight (tickets alternate): chunk c of the next call only waits for chunk c of the previous
// one, and with dense images the copy stream carries copies only (ingest on the kernel stream), so uploads run a whole call ahead.
int orb_extract_enqueue(orb_extractor* e, const uint8_t* imgs, int batch, int w, int h, int stride, size_t img_stride, orb_keypoint* kps,
                        uint8_t* desc, int cap, int* ticket) {
    if (!e || !imgs || !kps || !desc || !ticket || batch < 1 || w < 1 || h < 1 || stride < w) return ORB_ERR_BAD_ARG;
    if (e->inflight >= 2) { orb_set_error("two batches are already in flight on this handle: wait for one"); return ORB_ERR_CAPACITY; }
    ORB_CUDA(cudaSetDevice(e->device));
    if (e->inflight > 0 && (e->plan_w != w || e->plan_h != h || e->last_batch != batch)) {
        orb_set_error("batches in flight together must share image size and batch size");
        return ORB_ERR_BAD_ARG;
    }
    int st = ensure_plan(e, w, h, batch);
    if (st == ORB_OK) st = orb_pipe_streams(e);
    if (st != ORB_OK) return st;
    const std::vector<int> cb = orb_chunk_bounds(batch);
    const int nch = (int)cb.size() - 1;
    // events per chunk: upload done, kernels done, downloads done, ingest done; + 1 for the call
    if (!orb_pool_event(e, 4 * nch + 1)) { orb_set_error("event"); return ORB_ERR_CUDA; }
    const int tk = e->next_ticket & 1;
    if (e->inflight == 0) {   // uploads must not overtake work already queued on the kernel stream by other entry points
        ORB_CUDA(cudaEventRecord(e->ev_pool[4 * nch], e->stream));
        ORB_CUDA(cudaStreamWaitEvent(e->s_h2d, e->ev_pool[4 * nch], 0));
    }
    if (!e->ev_call[0]) for (int i = 0; i < 2; ++i) ORB_CUDA(cudaEventCreateWithFlags(&e->ev_call[i], cudaEventDisableTiming));
    const bool chained = e->inflight > 0 && e->pipe_chunks == nch;
    if (e->inflight > 0 && !chained) {
        ORB_CUDA(cudaStreamWaitEvent(e->s_h2d, e->ev_call[tk ^ 1], 0));
        ORB_CUDA(cudaStreamWaitEvent(e->stream, e->ev_call[tk ^ 1], 0));
    }
    const bool split = orb_upload_can_split(e, w, h, stride, img_stride) != 0;
    const bool was_split = e->pipe_split;
    int* hc = e->h_counts + (size_t)tk * e->max_batch;
    for (int c = 0; c < nch; ++c) {
        const int b0 = cb[c], n = cb[c + 1] - cb[c];
        cudaEvent_t up = e->ev_pool[4 * c], done = e->ev_pool[4 * c + 1], dl = e->ev_pool[4 * c + 2], ing = e->ev_pool[4 * c + 3];
        if (chained) {   // events still hold the previous call's records for this chunk
            if (split && was_split) ORB_CUDA(cudaStreamWaitEvent(e->s_h2d, ing, 0));   // the copy only overwrites the staging slots
            else ORB_CUDA(cudaStreamWaitEvent(e->s_h2d, done, 0));                    // the copy stream writes level 0 itself
            ORB_CUDA(cudaStreamWaitEvent(e->stream, dl, 0));                          // the slot's previous results have left
        }
        if (split) {
            if ((st = orb_upload_copy_slots(e, imgs, b0, n, w, h, e->s_h2d)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(up, e->s_h2d));
            ORB_CUDA(cudaStreamWaitEvent(e->stream, up, 0));
            if ((st = orb_ingest_slots(e, b0, n, w, h, e->stream)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(ing, e->stream));
        } else {
            if ((st = orb_upload_slots(e, imgs, b0, n, w, h, stride, img_stride, e->s_h2d)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(up, e->s_h2d));
            ORB_CUDA(cudaStreamWaitEvent(e->stream, up, 0));
        }
        if (!e->copy_only && (st = run_pipeline(e, b0, n)) != ORB_OK) break;
        ORB_CUDA(cudaEventRecord(done, e->stream));
        ORB_CUDA(cudaStreamWaitEvent(e->s_d2h, done, 0));
        if ((st = orb_download_slots(e, b0, n, kps, desc, cap, e->s_d2h, hc)) != ORB_OK) break;
        ORB_CUDA(cudaEventRecord(dl, e->s_d2h));
    }
    if (st != ORB_OK) {   // nothing of this call may stay in flight once the caller is told it failed (its buffers may go away)
        cudaStreamSynchronize(e->s_h2d); cudaStreamSynchronize(e->stream); cudaStreamSynchronize(e->s_d2h);
        e->inflight = 0; e->pipe_chunks = 0;
        return st;
    }
    if (!e->h_status) ORB_CUDA(cudaMallocHost((void**)&e->h_status, 4 * sizeof(int)));
    ORB_CUDA(cudaMemcpyAsync(&e->h_status[2 * tk], e->d_status, sizeof(int), cudaMemcpyDeviceToHost, e->s_d2h));
    ORB_CUDA(cudaEventRecord(e->ev_call[tk], e->s_d2h));
    e->pipe_chunks = nch; e->pipe_split = split;
    e->pipe_batch[tk] = batch; e->pipe_cap[tk] = cap;
    *ticket = e->next_ticket++;
    e->inflight++;
    return ORB_OK;
}

int orb_extract_wait(orb_extractor* e, int ticket, int32_t* counts) {
    if (!e || !counts) return ORB_ERR_BAD_ARG;
    if (e->inflight < 1 || ticket != e->next_ticket - e->inflight) { orb_set_error("tickets are waited for in the order they were issued"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(e->device));
    const int tk = ticket & 1;
    ORB_CUDA(cudaEventSynchronize(e->ev_call[tk]));
    e->inflight--;
    if (e->h_status[2 * tk] != 0) {
        const int dst = e->h_status[2 * tk];
        cudaMemsetAsync(e->d_status, 0, sizeof(int), e->stream);
        orb_set_error("device-side bound exceeded");
        return dst;
    }
    const int batch = e->pipe_batch[tk], cap = e->pipe_cap[tk];
    const int* hc = e->h_counts + (size_t)tk * e->max_batch;
    int rc = ORB_OK;
    for (int b = 0; b < batch; ++b) { counts[b] = hc[b]; if (counts[b] > cap) rc = ORB_ERR_CAPACITY; }
    if (rc != ORB_OK) orb_set_error("keypoint capacity too small");
    return rc;
}

int orb_extract(orb_extractor* e, const uint8_t* img, int w, int h, int stride, orb_keypoint* kps, uint8_t* desc, int cap, int32_t* n) {
    return orb_extract_batch(e, img, 1, w, h, stride, (size_t)stride * (size_t)(h > 0 ? h : 0), kps, desc, cap, n);
}

int orb_get_level(orb_extractor* e, int image, int level, int blurred, int border, uint8_t* dst, int dst_stride, int dst_cap_bytes,
                  int* w, int* h) {
    if (!e || !dst || e->plan_w < 0 || level < 0 || level >= e->nlevels || image < 0 || image >= e->last_batch || border < 0)
        return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    const LevelGeom& G = e->plan.lv[level];
    if (w) *w = G.w;
    if (h) *h = G.h;
    const int bw = G.w + 2 * border, bh = G.h + 2 * border;
    if (dst_stride < bw || (size_t)dst_stride * (bh - 1) + bw > (size_t)dst_cap_bytes) return ORB_ERR_CAPACITY;
    const uint8_t* src = (blurred ? e->d_blur : e->d_pyr) + G.off + (size_t)image * G.img_stride;
    ORB_CUDA(cudaMemcpy2DAsync(dst + (size_t)border * dst_stride + border, dst_stride, src, G.pitch, G.w, G.h, cudaMemcpyDeviceToHost, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    if (border > 0) {  // copyMakeBorder(BORDER_REFLECT_101) around the level (reference :1197-1203); never read by the hot path
        auto refl = [](int p, int len) { if (len == 1) return 0; while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p; return p; };
        for (int y = 0; y < bh; ++y) {
            const int sy = refl(y - border, G.h) + border;
            uint8_t* D = dst + (size_t)y * dst_stride;
            const uint8_t* S = dst + (size_t)sy * dst_stride;
            if (sy != y) std::memcpy(D + border, S + border, G.w);
            for (int x = 0; x < border; ++x) {
                D[x] = S[border + refl(x - border, G.w)];
                D[border + G.w + x] = S[border + refl(G.w + x, G.w)];
            }
        }
    }
    return ORB_OK;
}

int orb_level_size(const orb_extractor* e, int level, int32_t* w, int32_t* h) {
    if (!e || !w || !h || e->plan_w < 0 || level < 0 || level >= e->nlevels) return ORB_ERR_BAD_ARG;
    *w = e->plan.lv[level].w; *h = e->plan.lv[level].h;
    return ORB_OK;
}

int orb_get_pyramid(orb_extractor* e, int image, uint8_t* const* dst, const int32_t* dst_stride) {
    if (!e || !dst || !dst_stride || e->plan_w < 0 || image < 0 || image >= e->last_batch) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    size_t need = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        if (!dst[l] || dst_stride[l] < e->plan.lv[l].w) return ORB_ERR_BAD_ARG;
        need += (size_t)e->plan.lv[l].w * e->plan.lv[l].h;
    }
    if (need > e->h_pyr_cap) {
        if (e->h_pyr) { cudaFreeHost(e->h_pyr); e->h_pyr = nullptr; e->h_pyr_cap = 0; }
        ORB_CUDA(cudaMallocHost((void**)&e->h_pyr, need));
        e->h_pyr_cap = need;
    }
    size_t off = 0;
    for (int l = 0; l < e->nlevels; ++l) {   // dense rows into pinned memory: the copies queue back to back
        const LevelGeom& G = e->plan.lv[l];
        ORB_CUDA(cudaMemcpy2DAsync(e->h_pyr + off, G.w, e->d_pyr + G.off + (size_t)image * G.img_stride, G.pitch, G.w, G.h, cudaMemcpyDeviceToHost, e->stream));
        off += (size_t)G.w * G.h;
    }
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    off = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        const LevelGeom& G = e->plan.lv[l];
        if (dst_stride[l] == G.w) std::memcpy(dst[l], e->h_pyr + off, (size_t)G.w * G.h);
        else for (int y = 0; y < G.h; ++y) std::memcpy(dst[l] + (size_t)y * dst_stride[l], e->h_pyr + off + (size_t)y * G.w, G.w);
        off += (size_t)G.w * G.h;
    }
    return ORB_OK;
}

int orb_get_level_candidates(orb_extractor* e, int image, int level, int32_t* xys, int cap, int32_t* n) {
    if (!e || !xys || !n || e->plan_w < 0 || level < 0 || level >= e->nlevels || image < 0 || image >= e->last_batch) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    const Plan& P = e->plan;
    const LevelGeom& G = P.lv[level];
    const int nc = G.n_cols * G.n_rows;
    std::vector<int> cnt(std::max(nc, 1));
    ORB_CUDA(cudaMemcpyAsync(cnt.data(), e->d_cell_cnt + (size_t)image * P.total_cells + G.cell_off, (size_t)nc * sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    std::vector<uint32_t> st((size_t)nc * G.cell_cap + 1);
    ORB_CUDA(cudaMemcpyAsync(st.data(), e->d_stage + (size_t)image * P.stage_per_image + G.stage_off, (size_t)nc * G.cell_cap * 4, cudaMemcpyDeviceToHost, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    int m = 0;
    for (int c = 0; c < nc; ++c)
        for (int i = 0; i < cnt[c]; ++i) {
            if (m < cap) {
                const uint32_t k = st[(size_t)c * G.cell_cap + i];
                xys[3 * m] = ORB_KEY_X(k); xys[3 * m + 1] = ORB_KEY_Y(k); xys[3 * m + 2] = ORB_KEY_S(k);
            }
            ++m;
        }
    *n = m;
    return m > cap ? ORB_ERR_CAPACITY : ORB_OK;
}

int orb_get_level_keypoints(orb_extractor* e, int image, int level, int32_t* xys, int cap, int32_t* n) {
    if (!e || !xys || !n || e->plan_w < 0 || level < 0 || level >= e->nlevels || image < 0 || image >= e->last_batch) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(e->device));
    const Plan& P = e->plan;
    const LevelGeom& G = P.lv[level];
    int cnt = 0;
    ORB_CUDA(cudaMemcpyAsync(&cnt, e->d_lvl_cnt + image * ORB_MAX_LEVELS + level, sizeof(int), cudaMemcpyDeviceToHost, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    std::vector<uint32_t> k(std::max(cnt, 1));
    ORB_CUDA(cudaMemcpyAsync(k.data(), e->d_lvl_kp + (size_t)image * P.kp_cap + G.kp_off, (size_t)cnt * 4, cudaMemcpyDeviceToHost, e->stream));
    ORB_CUDA(cudaStreamSynchronize(e->stream));
    *n = cnt;
    if (cnt > cap) return ORB_ERR_CAPACITY;
    for (int i = 0; i < cnt; ++i) { xys[3 * i] = ORB_KEY_X(k[i]) + ORB_WIN; xys[3 * i + 1] = ORB_KEY_Y(k[i]) + ORB_WIN; xys[3 * i + 2] = ORB_KEY_S(k[i]); }
    return ORB_OK;
}

}  // extern "C"
