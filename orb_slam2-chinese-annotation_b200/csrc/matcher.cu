// ORBmatcher hot paths (reference src/ORBmatcher.cc) and the Frame grid (src/Frame.cc:261-276,
// 368-436) on the GPU.  XOR + POPC over 256-bit descriptors; no tensor cores by design.
//
// Structure shared by the claim-based searches (SearchByProjection x2, SearchByBoW): the Hamming
// distances of every (query, candidate) pair are computed in parallel into per-query candidate
// lists that keep the reference's candidate ORDER; a single warp then replays the reference's
// sequential loop over the queries (skip already-claimed keypoints, strict-< best / second best,
// thresholds, ratio test, claim), so match indices are identical to the CPU code including ties.
#include "orb_internal.h"
#include <algorithm>
#include <cmath>
#include <cstring>
#include <mutex>
#include <vector>

#define GRID_COLS 64
#define GRID_ROWS 48
#define GRID_CELLS (GRID_COLS * GRID_ROWS)
#define M_TH_HIGH 100
#define M_TH_LOW 50
#define M_HISTO 30

struct orb_matcher {
    int device = 0, max_pairs = 0, max_n = 0;
    cudaStream_t stream = nullptr;
    long long launches = 0;
    uint8_t* d_ws = nullptr;   // bump-allocated workspace, reset per call
    size_t ws_cap = 0, ws_off = 0;
    uint32_t* d_cands = nullptr;
    size_t cands_cap = 0;
    int* d_scal = nullptr;     // small device scalars
    long long last_cands = 0;  // candidates examined by the last projection search (bench: popc work)
    // Small host arrays of a call (a frame's SoA fields, the points' fields: a dozen of them) are copied into a pinned mirror of the
    // workspace first; runs of adjacent ones then cross the link as ONE cudaMemcpyAsync, issued right before the first kernel that can
    // read them (ws_stream).  A dozen pageable copies of a few KB each cost more than the kernels of a single-frame search.
    uint8_t* h_stage = nullptr;
    size_t h_stage_cap = 0;
    std::vector<std::pair<size_t, size_t>> pend;   // staged, not yet copied: (workspace offset, bytes), ascending
    cudaEvent_t ev_stage = nullptr;
    bool stage_busy = false;
    // Results into pageable memory take the same way back: queued into the mirror (d_scal: into h_scal), handed to the caller by the CPU
    // after ONE wait (ws_download / ws_finish); every such copy made directly holds the calling thread until it is done.
    struct Deliver { void* dst; const uint8_t* src; size_t bytes; };
    std::vector<Deliver> deliver;
    uint8_t* h_scal = nullptr;   // pinned mirror of d_scal (64 bytes)
};

namespace {

struct DFrame {
    int n;
    const float *x, *y, *angle, *u_right;
    const int* octave;
    const uint8_t* desc;
    orb_grid_params grid;
    float sf[ORB_MAX_LEVELS];
    const int* cell_start;
    const int* cell_items;
    // batched calls: frame p of the batch starts p * stride elements further in every array (0: a single frame)
    int stride;
    const int* n_arr;
};

// Query-side layout of a batched call (one frame pair per CTA / grid row): pair p reads its points at p * qstride, its
// candidate offsets at p * (qstride + 1), its candidates at cand_base[p].  qstride == 0: a single pair.
struct Batch {
    int qstride;
    const int* nq_arr;
    const int* cand_base;
};
__device__ __forceinline__ DFrame frame_at(const DFrame& F, int p) {
    DFrame G = F;
    if (F.stride) {
        const size_t o = (size_t)p * F.stride;
        G.n = F.n_arr[p];
        G.x += o; G.y += o; G.octave += o; G.desc += o * 32;
        if (G.angle) G.angle += o;
        if (G.u_right) G.u_right += o;
        G.cell_start += (size_t)p * (GRID_CELLS + 1);
        G.cell_items += o;
    }
    return G;
}

struct Query {
    float x, y, r, xr;
    int min_level, max_level;
    int valid;
    int pad;
};

template <typename T>
int ws_alloc(orb_matcher* m, T** p, size_t count) {
    const size_t bytes = (count * sizeof(T) + 255) & ~(size_t)255;
    if (m->ws_off + bytes > m->ws_cap) { orb_set_error("matcher workspace too small (raise max_n / max_pairs)"); return ORB_ERR_CAPACITY; }
    *p = reinterpret_cast<T*>(m->d_ws + m->ws_off);
    m->ws_off += bytes;
    return ORB_OK;
}

template <typename T>
int ws_upload(orb_matcher* m, const T** dp, const T* host, size_t count) {
    T* p = nullptr;
    const size_t off = m->ws_off;
    int st = ws_alloc(m, &p, std::max<size_t>(count, 1));
    if (st != ORB_OK) return st;
    const size_t bytes = count * sizeof(T);
    if (count && m->h_stage && bytes <= (1u << 20) && m->ws_off <= m->h_stage_cap) {
        std::memcpy(m->h_stage + off, host, bytes);
        const size_t padded = m->ws_off - off;   // the allocation's padding travels along (it belongs to nobody else)
        if (!m->pend.empty() && m->pend.back().first + m->pend.back().second == off) m->pend.back().second += padded;
        else m->pend.push_back(std::make_pair(off, padded));
    } else if (count) {
        ORB_CUDA(cudaMemcpyAsync(p, host, bytes, cudaMemcpyHostToDevice, m->stream));
    }
    *dp = p;
    return ORB_OK;
}
// The stream argument of every kernel launch of this file: staged uploads go out first.
static cudaStream_t ws_stream(orb_matcher* m) {
    if (!m->pend.empty()) {
        for (const auto& sg : m->pend) cudaMemcpyAsync(m->d_ws + sg.first, m->h_stage + sg.first, sg.second, cudaMemcpyHostToDevice, m->stream);
        cudaEventRecord(m->ev_stage, m->stream);
        m->stage_busy = true;
        m->pend.clear();
    }
    return m->stream;
}

// A result of the call: `dev` (inside the workspace or d_scal) -> `dst`.  Pageable dst: queued into the pinned mirror at the same offset, delivered
// by ws_finish; anything else: copied directly.
static cudaError_t ws_download(orb_matcher* m, void* dst, const void* dev, size_t bytes) {
    const uint8_t* d = static_cast<const uint8_t*>(dev);
    const uint8_t* sc = reinterpret_cast<const uint8_t*>(m->d_scal);
    uint8_t* mirror = nullptr;
    if (bytes == 0) return cudaSuccess;
    if (m->h_stage && d >= m->d_ws && d + bytes <= m->d_ws + m->h_stage_cap) mirror = m->h_stage + (d - m->d_ws);
    else if (m->h_scal && d >= sc && d + bytes <= sc + 64) mirror = m->h_scal + (d - sc);
    if (!mirror || bytes > (1u << 20) || !orb_host_pageable(dst)) return cudaMemcpyAsync(dst, dev, bytes, cudaMemcpyDeviceToHost, m->stream);
    const cudaError_t e = cudaMemcpyAsync(mirror, dev, bytes, cudaMemcpyDeviceToHost, m->stream);
    if (e == cudaSuccess) { orb_matcher::Deliver dl = {dst, mirror, bytes}; m->deliver.push_back(dl); }
    return e;
}
// Waits for the call's stream and hands the mirrored results over.
static cudaError_t ws_finish(orb_matcher* m) {
    const cudaError_t e = cudaStreamSynchronize(m->stream);
    if (e == cudaSuccess) for (const auto& dl : m->deliver) std::memcpy(dl.dst, dl.src, dl.bytes);
    m->deliver.clear();
    return e;
}

// ------------------------------------------------------------------------------------------
__global__ void k_distance_pairs(const uint8_t* __restrict__ a, const uint8_t* __restrict__ b, int n, int* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint4* pa = reinterpret_cast<const uint4*>(a + (size_t)i * 32);
    const uint4* pb = reinterpret_cast<const uint4*>(b + (size_t)i * 32);
    out[i] = orb_hamming256(pa[0], pa[1], pb[0], pb[1]);
}

// Brute force: thread = one row of A (descriptor in registers), B streamed through shared memory
// in tiles that every thread reads at the same address (broadcast).  Scan order over B is
// ascending, so "strict <, first wins" is the natural loop.
#define BF_THREADS 128
#define BF_TILE 256
__global__ void __launch_bounds__(BF_THREADS)
k_bruteforce(const uint8_t* __restrict__ A, const uint8_t* __restrict__ Bm, int n_a, int n_b, int* __restrict__ best_idx,
             int* __restrict__ best_dist, int* __restrict__ second_dist) {
    __shared__ uint4 tile[BF_TILE * 2];
    const int pair = blockIdx.y;
    const uint8_t* a = A + (size_t)pair * n_a * 32;
    const uint8_t* b = Bm + (size_t)pair * n_b * 32;
    const int i = blockIdx.x * BF_THREADS + threadIdx.x;
    uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
    if (i < n_a) { const uint4* pa = reinterpret_cast<const uint4*>(a + (size_t)i * 32); a0 = pa[0]; a1 = pa[1]; }
    int b1 = 256, b2 = 256, bi = -1;
    for (int j0 = 0; j0 < n_b; j0 += BF_TILE) {
        const int m = min(BF_TILE, n_b - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < 2 * m; t += BF_THREADS) tile[t] = reinterpret_cast<const uint4*>(b + (size_t)j0 * 32)[t];
        __syncthreads();
#pragma unroll 4
        for (int j = 0; j < m; ++j) {
            const int d = orb_hamming256(a0, a1, tile[2 * j], tile[2 * j + 1]);
            if (d < b1) { b2 = b1; b1 = d; bi = j0 + j; }
            else if (d < b2) b2 = d;
        }
    }
    if (i < n_a) {
        const size_t o = (size_t)pair * n_a + i;
        best_idx[o] = bi; best_dist[o] = b1; second_dist[o] = b2;
    }
}

// ------------------------------------------------------------------------------------------
// Frame::AssignFeaturesToGrid: cell = round((x - minX) * invW), round((y - minY) * invH); lists in
// ascending keypoint index (insertion order).  One CTA.  Counting sort by cell (shared-memory atomics, a block-wide scan of the 3072
// cell counts, unordered scatter), then every cell's short list is put in ascending index order by one thread (insertion sort).
// A cell with more than GRID_SMALL keypoints (degenerate: all features in a few cells) takes the bitonic sort of (cell << 16 | index).
#define GRID_SMALL 48
// The kernel's static shared memory (two 3072-entry tables: 24720 bytes, see build/matcher.ptxas.log) counts against the 48 KB a launch
// gets without opting in: beyond this many bytes of keys the launchers raise cudaFuncAttributeMaxDynamicSharedMemorySize.
#define GRID_BUILD_DYN_DEFAULT (48 * 1024 - 25 * 1024)
__global__ void __launch_bounds__(1024)
k_grid_build(const float* __restrict__ x, const float* __restrict__ y, int n, int p2, orb_grid_params g, int* __restrict__ cell_start,
             int* __restrict__ cell_items, int* __restrict__ n_items, int stride = 0, const int* __restrict__ n_arr = nullptr) {
    extern __shared__ uint32_t keys[];
    __shared__ int cnt[GRID_CELLS + 1];
    ORB_PDL_SYNC();
    __shared__ int cur[GRID_CELLS];
    __shared__ int s_warp[33];
    __shared__ int s_max;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (stride) {   // one CTA per frame of a batch
        const size_t o = (size_t)blockIdx.x * stride;
        x += o; y += o; cell_items += o; cell_start += (size_t)blockIdx.x * (GRID_CELLS + 1); n_items += blockIdx.x;
        n = n_arr[blockIdx.x];
    }
    for (int i = tid; i <= GRID_CELLS; i += 1024) cnt[i] = 0;
    if (tid == 0) s_max = 0;
    __syncthreads();
    for (int i = tid; i < p2; i += 1024) {
        uint32_t k = 0xffffffffu;
        if (i < n) {
            const int px = (int)roundf(__fmul_rn(__fsub_rn(x[i], g.min_x), g.inv_w));
            const int py = (int)roundf(__fmul_rn(__fsub_rn(y[i], g.min_y), g.inv_h));
            if (px >= 0 && px < GRID_COLS && py >= 0 && py < GRID_ROWS) {
                const int c = px * GRID_ROWS + py;
                k = ((uint32_t)c << 16) | (uint32_t)i;
                atomicMax(&s_max, atomicAdd(&cnt[c], 1) + 1);
            }
        }
        keys[i] = k;
    }
    __syncthreads();
    const bool small = s_max <= GRID_SMALL;
    if (!small) {
        for (int k = 2; k <= p2; k <<= 1)
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = tid; i < p2; i += 1024) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const uint32_t a = keys[i], b = keys[ixj];
                        const bool up = (i & k) == 0;
                        if (up ? (a > b) : (a < b)) { keys[i] = b; keys[ixj] = a; }
                    }
                }
                __syncthreads();
            }
    }
    // exclusive scan of the 3072 counts: three consecutive cells per thread, warp scan, scan of the 32 warp totals
    static_assert(GRID_CELLS == 3 * 1024, "three cells per thread");
    {
        const int c0 = cnt[3 * tid], c1 = cnt[3 * tid + 1], c2 = cnt[3 * tid + 2];
        const int mine = c0 + c1 + c2;
        int inc = mine;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += v; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const int w = s_warp[lane];
            int wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += v; }
            s_warp[lane] = wi - w;
            if (lane == 31) s_warp[32] = wi;
        }
        __syncthreads();
        const int ex = s_warp[warp] + inc - mine;
        cnt[3 * tid] = ex; cnt[3 * tid + 1] = ex + c0; cnt[3 * tid + 2] = ex + c0 + c1;
        cur[3 * tid] = ex; cur[3 * tid + 1] = ex + c0; cur[3 * tid + 2] = ex + c0 + c1;
        if (tid == 0) { cnt[GRID_CELLS] = s_warp[32]; *n_items = s_warp[32]; }
    }
    __syncthreads();
    for (int i = tid; i <= GRID_CELLS; i += 1024) cell_start[i] = cnt[i];
    const int total = cnt[GRID_CELLS];
    if (!small) {
        for (int i = tid; i < total; i += 1024) cell_items[i] = (int)(keys[i] & 0xffffu);
        return;
    }
    for (int i = tid; i < n; i += 1024) {
        const uint32_t k = keys[i];
        if (k != 0xffffffffu) cell_items[atomicAdd(&cur[k >> 16], 1)] = i;
    }
    __syncthreads();
    for (int c = tid; c < GRID_CELLS; c += 1024) {   // ascending index inside the cell
        const int a = cnt[c], b = cnt[c + 1];
        for (int i = a + 1; i < b; ++i) {
            const int v = cell_items[i];
            int j = i - 1;
            while (j >= a && cell_items[j] > v) { cell_items[j + 1] = cell_items[j]; --j; }
            cell_items[j + 1] = v;
        }
    }
}

// ------------------------------------------------------------------------------------------
// Query records.
__global__ void k_prep_mappoints(int n, const uint8_t* __restrict__ valid, const float* __restrict__ px, const float* __restrict__ py,
                                 const float* __restrict__ pxr, const int* __restrict__ level, const float* __restrict__ view_cos,
                                 float th, DFrame F, Query* __restrict__ q, Batch B) {
    ORB_PDL_SYNC();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (B.qstride) {
        const size_t o = (size_t)blockIdx.y * B.qstride;
        n = B.nq_arr[blockIdx.y];
        valid += o; px += o; py += o; level += o; view_cos += o; q += o;
        if (pxr) pxr += o;
    }
    if (i >= n) return;
    Query Q;
    Q.valid = valid[i];
    const int lv = level[i];
    float r = ((double)view_cos[i] > 0.998) ? 2.5f : 4.0f;     // RadiusByViewingCos (reference :135-141)
    if (th != 1.0f) r = __fmul_rn(r, th);
    Q.x = px[i]; Q.y = py[i]; Q.xr = pxr ? pxr[i] : 0.f;
    Q.r = __fmul_rn(r, F.sf[min(max(lv, 0), ORB_MAX_LEVELS - 1)]);
    Q.min_level = lv - 1; Q.max_level = lv; Q.pad = 0;
    q[i] = Q;
}

__global__ void k_prep_lastframe(int n, const uint8_t* __restrict__ valid, const float* __restrict__ xc, const float* __restrict__ yc,
                                 const float* __restrict__ zc, const int* __restrict__ octave, orb_camera cam, float th, int forward,
                                 int backward, DFrame F, Query* __restrict__ q, Batch B, const uint8_t* __restrict__ fwd_arr,
                                 const uint8_t* __restrict__ bwd_arr) {
    ORB_PDL_SYNC();
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (B.qstride) {
        const size_t o = (size_t)blockIdx.y * B.qstride;
        n = B.nq_arr[blockIdx.y];
        valid += o; xc += o; yc += o; zc += o; octave += o; q += o;
        forward = fwd_arr[blockIdx.y]; backward = bwd_arr[blockIdx.y];
    }
    if (i >= n) return;
    Query Q;
    Q.valid = valid[i]; Q.pad = 0; Q.x = Q.y = Q.r = Q.xr = 0.f; Q.min_level = Q.max_level = 0;
    if (Q.valid) {
        const float invzc = (float)(1.0 / (double)zc[i]);      // reference :1507 (double division)
        if (invzc < 0) Q.valid = 0;
        else {
            // x*a + y, fused as GCC does for the reference's flags on FMA hardware
            const float u = __fmaf_rn(__fmul_rn(cam.fx, xc[i]), invzc, cam.cx);
            const float v = __fmaf_rn(__fmul_rn(cam.fy, yc[i]), invzc, cam.cy);
            if (u < F.grid.min_x || u > F.grid.max_x || v < F.grid.min_y || v > F.grid.max_y) Q.valid = 0;
            else {
                const int lo = octave[i];
                Q.x = u; Q.y = v;
                Q.r = __fmul_rn(th, F.sf[min(max(lo, 0), ORB_MAX_LEVELS - 1)]);
                Q.xr = __fmaf_rn(-cam.mbf, invzc, u);
                if (forward) { Q.min_level = lo; Q.max_level = -1; }
                else if (backward) { Q.min_level = 0; Q.max_level = lo; }
                else { Q.min_level = lo - 1; Q.max_level = lo + 1; }
            }
        }
    }
    q[i] = Q;
}

// static per-candidate filters of k_area_candidates
#define FILTER_STEREO_WINDOW 0   // |proj_xr - mvuRight| <= r for stereo keypoints (:95-100, :1555-1561)
#define FILTER_NONE 1
#define FILTER_FUSE_CHI2 2       // chi-square reprojection test of Fuse (:977-1003)

// Frame::GetFeaturesInArea + the static candidate filters, one warp per query, in the
// reference's order (ix, then iy, then insertion order).  FILL=false counts, FILL=true writes
// packed candidates: index (16 bits) | distance << 16 (9 bits) | octave << 25.
template <bool FILL>
__global__ void __launch_bounds__(256)
k_area_candidates(int nq, const Query* __restrict__ q, const uint8_t* __restrict__ qdesc, DFrame F0, int filter,
                  const float* __restrict__ inv_sigma2, int* __restrict__ counts, const int* __restrict__ starts,
                  uint32_t* __restrict__ cands, Batch B, const int cap = 0) {
    ORB_PDL_SYNC();
    // cap > 0 (single-pair calls): the candidate buffer's capacity; a fill that would not fit does nothing (the host grows the buffer and repeats)
    if (FILL && cap > 0 && starts[nq] > cap) return;
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (B.qstride) {
        const size_t o = (size_t)blockIdx.y * B.qstride;
        nq = B.nq_arr[blockIdx.y];
        q += o; counts += o;
        if (FILL) { qdesc += o * 32; starts += (size_t)blockIdx.y * (B.qstride + 1); cands += B.cand_base[blockIdx.y]; }
    }
    if (w >= nq) return;
    const DFrame F = frame_at(F0, blockIdx.y);
    const Query Q = q[w];
    int total = 0;
    if (Q.valid) {
        const float x = Q.x, y = Q.y, r = Q.r;
        const int nMinCellX = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(x, F.grid.min_x), r), F.grid.inv_w)));
        const int nMaxCellX = min(GRID_COLS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(x, F.grid.min_x), r), F.grid.inv_w)));
        const int nMinCellY = max(0, (int)floorf(__fmul_rn(__fsub_rn(__fsub_rn(y, F.grid.min_y), r), F.grid.inv_h)));
        const int nMaxCellY = min(GRID_ROWS - 1, (int)ceilf(__fmul_rn(__fadd_rn(__fsub_rn(y, F.grid.min_y), r), F.grid.inv_h)));
        if (!(nMinCellX >= GRID_COLS || nMaxCellX < 0 || nMinCellY >= GRID_ROWS || nMaxCellY < 0)) {
            const bool check_levels = (Q.min_level > 0) || (Q.max_level >= 0);
            uint4 a0 = make_uint4(0, 0, 0, 0), a1 = a0;
            uint32_t* out = nullptr;
            if (FILL) {
                const uint4* pa = reinterpret_cast<const uint4*>(qdesc + (size_t)w * 32);
                a0 = pa[0]; a1 = pa[1];
                out = cands + starts[w];
            }
            for (int ix = nMinCellX; ix <= nMaxCellX; ++ix) {
                // cells (ix, nMinCellY..nMaxCellY) are consecutive in the CSR
                const int c0 = F.cell_start[ix * GRID_ROWS + nMinCellY], c1 = F.cell_start[ix * GRID_ROWS + nMaxCellY + 1];
                for (int p0 = c0; p0 < c1; p0 += 32) {
                    const int p = p0 + lane;
                    bool ok = false;
                    int idx = 0, oct = 0;
                    if (p < c1) {
                        idx = F.cell_items[p];
                        oct = F.octave[idx];
                        ok = true;
                        if (check_levels) {
                            if (oct < Q.min_level) ok = false;
                            if (Q.max_level >= 0 && oct > Q.max_level) ok = false;
                        }
                        if (ok) {
                            const float dx = __fsub_rn(F.x[idx], x), dy = __fsub_rn(F.y[idx], y);
                            ok = fabsf(dx) < r && fabsf(dy) < r;
                        }
                        if (ok && filter == FILTER_STEREO_WINDOW && F.u_right) {
                            const float ur = F.u_right[idx];
                            if (ur > 0 && fabsf(__fsub_rn(Q.xr, ur)) > r) ok = false;
                        }
                        if (ok && filter == FILTER_FUSE_CHI2) {
                            // reprojection test of Fuse(KF, MapPoints) (reference :977-1003), GCC-style contraction
                            const float ex = __fsub_rn(x, F.x[idx]), ey = __fsub_rn(y, F.y[idx]);
                            const float kpr = F.u_right ? F.u_right[idx] : -1.f;
                            if (kpr >= 0) {
                                const float er = __fsub_rn(Q.xr, kpr);
                                const float e2 = __fmaf_rn(er, er, __fmaf_rn(ex, ex, __fmul_rn(ey, ey)));
                                if ((double)__fmul_rn(e2, inv_sigma2[oct]) > 7.8) ok = false;
                            } else {
                                const float e2 = __fmaf_rn(ex, ex, __fmul_rn(ey, ey));
                                if ((double)__fmul_rn(e2, inv_sigma2[oct]) > 5.99) ok = false;
                            }
                        }
                    }
                    const unsigned m = __ballot_sync(0xffffffffu, ok);
                    if (FILL && ok) {
                        const uint4* pb = reinterpret_cast<const uint4*>(F.desc + (size_t)idx * 32);
                        const unsigned d = (unsigned)orb_hamming256(a0, a1, pb[0], pb[1]);
                        out[total + __popc(m & ((1u << lane) - 1))] = (uint32_t)idx | (d << 16) | ((uint32_t)oct << 25);
                    }
                    total += __popc(m);
                }
            }
        }
    }
    if (!FILL && lane == 0) counts[w] = total;
}

// single-CTA exclusive scan: starts[0..n] from counts[0..n)
__global__ void __launch_bounds__(1024) k_scan(const int* __restrict__ counts, int n, int* __restrict__ starts, Batch B, int* __restrict__ totals) {
    __shared__ int ws[33];
    ORB_PDL_SYNC();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (B.qstride) { n = B.nq_arr[blockIdx.x]; counts += (size_t)blockIdx.x * B.qstride; starts += (size_t)blockIdx.x * (B.qstride + 1); }
    int carry = 0;
    for (int i0 = 0; i0 < n; i0 += 1024) {
        const int i = i0 + tid;
        const int v = i < n ? counts[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += t; }
        if (lane == 31) ws[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int wv = ws[lane], wi = wv;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += t; }
            ws[lane] = wi - wv;
            if (lane == 31) ws[32] = wi;
        }
        __syncthreads();
        if (i < n) starts[i] = carry + inc - v + ws[warp];
        carry += ws[32];
        __syncthreads();
    }
    if (tid == 0) { starts[n] = carry; if (totals) totals[blockIdx.x] = carry; }
}
// candidate base of every pair: exclusive scan of the per-pair totals (one thread: a few thousand pairs at most)
__global__ void k_pair_bases(const int* __restrict__ totals, int pairs, int* __restrict__ base) {
    if (threadIdx.x || blockIdx.x) return;
    long long run = 0;
    for (int p = 0; p < pairs; ++p) { base[p] = (int)run; run += totals[p]; }
    base[pairs] = run > 0x7fffffffLL ? -1 : (int)run;
}

// ------------------------------------------------------------------------------------------
// Sequential replay helpers (one warp).  A candidate key is (dist << 16 | position): the
// minimum is the reference's "strict <, first wins" best; the minimum of the rest its second best.
struct Best2 { unsigned b, s; };
__device__ __forceinline__ Best2 merge2(Best2 a, Best2 o) {
    Best2 r;
    r.b = min(a.b, o.b);
    r.s = min(min(a.s, o.s), max(a.b, o.b));
    return r;
}
__device__ __forceinline__ Best2 warp_best2(Best2 v) {
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        Best2 t;
        t.b = __shfl_xor_sync(0xffffffffu, v.b, o);
        t.s = __shfl_xor_sync(0xffffffffu, v.s, o);
        v = merge2(v, t);
    }
    return v;
}

__device__ __forceinline__ int rot_bin(float a1, float a2) {
    float rot = __fsub_rn(a1, a2);
    if (rot < 0.0f) rot = __fadd_rn(rot, 360.0f);
    int bin = (int)roundf(__fmul_rn(rot, 1.0f / M_HISTO));
    if (bin == M_HISTO) bin = 0;
    return bin;
}

// ComputeThreeMaxima (reference :1765-1809) over bin counts
__device__ void three_maxima(const int* cnt, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    ind1 = ind2 = ind3 = -1;
    for (int i = 0; i < M_HISTO; ++i) {
        const int s = cnt[i];
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if ((float)max2 < __fmul_rn(0.1f, (float)max1)) { ind2 = -1; ind3 = -1; }
    else if ((float)max3 < __fmul_rn(0.1f, (float)max1)) ind3 = -1;
}

// rotation-histogram cull shared by the searches: entries (target index, bin); clears match[target]
__device__ void rot_cull_warp(const int* __restrict__ hist_idx, const int* __restrict__ hist_bin, int n_hist, int* __restrict__ match,
                              int* nmatches_io, int* s_cnt /* 32 ints shared */) {
    const int lane = threadIdx.x & 31;
    if (lane < M_HISTO + 2) s_cnt[lane] = 0;
    __syncwarp();
    for (int i = lane; i < n_hist; i += 32) atomicAdd(&s_cnt[hist_bin[i]], 1);
    __syncwarp();
    int i1, i2, i3;
    three_maxima(s_cnt, i1, i2, i3);
    int removed = 0;
    for (int i = lane; i < n_hist; i += 32) {
        const int b = hist_bin[i];
        if (b != i1 && b != i2 && b != i3) { match[hist_idx[i]] = -1; ++removed; }
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) removed += __shfl_xor_sync(0xffffffffu, removed, o);
    if (lane == 0) *nmatches_io -= removed;
}

#define RP_G 8
// offsets and first 32 candidates (one per lane) of points [base, base + RP_G); points past nq come out empty
__device__ __forceinline__ void rp_load_group(int base, int nq, int lane, const int* __restrict__ starts, const uint32_t* __restrict__ cands,
                                              uint32_t (&c)[RP_G], int (&c0)[RP_G], int (&c1)[RP_G]) {
    const int sv = __ldg(starts + min(base + min(lane, RP_G), nq));
#pragma unroll
    for (int k = 0; k < RP_G; ++k) {
        const int a = __shfl_sync(0xffffffffu, sv, k), b = __shfl_sync(0xffffffffu, sv, k + 1);
        c0[k] = a; c1[k] = base + k < nq ? b : a;
        c[k] = (a + lane < c1[k]) ? __ldg(cands + a + lane) : 0xffffffffu;
    }
}
// candidate at position pos of a point's list: the first 32 sit in the lanes' registers (warp-uniform pos)
__device__ __forceinline__ uint32_t rp_cand(const uint32_t* __restrict__ cands, int c0, unsigned pos, uint32_t myc) {
    return pos < 32u ? __shfl_sync(0xffffffffu, myc, (int)pos) : cands[c0 + pos];
}

// MODE 0: SearchByProjection(Frame, MapPoints)  -- best + second best with levels, ratio test
// MODE 1: SearchByProjection(Cur, Last)         -- best only, rotation histogram
// MODE 2: projected points with claiming (relocalisation :1622-1759, loop :311-427) -- best only, accept
//         <= max_dist, every accepted keypoint is closed to later points, optional rotation histogram
__global__ void __launch_bounds__(32)
k_resolve_projection(int mode, int nq, const Query* __restrict__ q, const int* __restrict__ starts, const uint32_t* __restrict__ cands,
                     const uint8_t* __restrict__ blocks, const float* __restrict__ q_angle, DFrame F0, const uint8_t* __restrict__ occupied,
                     uint8_t* occ, float nn_ratio, int check_ori, int max_dist, int* __restrict__ match,
                     int* __restrict__ q_best_idx, int* __restrict__ q_best_dist, int* __restrict__ hist_idx,
                     int* __restrict__ hist_bin, int* __restrict__ nmatches_out, int occ_shared, Batch B) {
    __shared__ int s_cnt[32];
    extern __shared__ uint8_t s_occ[];   // the claimed-keypoint flags live in shared memory when they fit (occ_shared): they are
    const DFrame F = frame_at(F0, blockIdx.x);
    if (B.qstride) {   // one warp (CTA) per frame pair of a batch
        const int p = blockIdx.x;
        const size_t o = (size_t)p * B.qstride, fo = (size_t)p * F0.stride;
        nq = B.nq_arr[p];
        q += o; starts += (size_t)p * (B.qstride + 1); cands += B.cand_base[p]; blocks += o;
        if (q_angle) q_angle += o;
        if (occupied) occupied += fo;
        occ += fo; match += fo; hist_idx += o; hist_bin += o; nmatches_out += p;
        if (q_best_idx) { q_best_idx += o; q_best_dist += o; }
    }
    if (occ_shared) occ = s_occ;         // read and written on the serial chain, one L2 round trip per point otherwise
    const int lane = threadIdx.x;
    for (int i = lane; i < F.n; i += 32) { match[i] = -1; occ[i] = occupied ? occupied[i] : 0; }
    __syncwarp();
    int nmatches = 0, n_hist = 0;
    if (q_best_idx) for (int i = lane; i < nq; i += 32) { q_best_idx[i] = -1; q_best_dist[i] = 256; }
    __syncwarp();
    // The points are resolved one after the other (a keypoint claimed by an earlier point is closed to later ones), but their
    // candidate lists do not depend on that order: the lists of the next RP_G points (offsets, then the first 32 candidates of
    // each, one per lane) are loaded while the current RP_G points are resolved, which takes the two dependent global loads
    // per point out of the serial chain.
    uint32_t pre[RP_G], nxt[RP_G];
    int p0[RP_G], p1[RP_G], n0[RP_G], n1[RP_G];
    rp_load_group(0, nq, lane, starts, cands, pre, p0, p1);
    for (int base = 0; base < nq; base += RP_G) {
        rp_load_group(base + RP_G, nq, lane, starts, cands, nxt, n0, n1);
        auto serial_one = [&](const int w, const int c0, const int c1, const uint32_t myc) {
        if (c0 == c1) return;
        Best2 v; v.b = 0xffffffffu; v.s = 0xffffffffu;
        if (c0 + lane < c1 && !occ[myc & 0xffffu]) { v.b = (((myc >> 16) & 0x1ffu) << 16) | (unsigned)lane; }
        for (int p = c0 + 32 + lane; p < c1; p += 32) {
            const uint32_t c = cands[p];
            if (occ[c & 0xffffu]) continue;
            Best2 t; t.b = (((c >> 16) & 0x1ffu) << 16) | (unsigned)(p - c0); t.s = 0xffffffffu;
            v = merge2(v, t);
        }
        v = warp_best2(v);
        if (v.b == 0xffffffffu) return;
        const int bestDist = (int)(v.b >> 16);
        const uint32_t cb = rp_cand(cands, c0, v.b & 0xffffu, myc);
        const int bestIdx = (int)(cb & 0xffffu);
        if (mode == 0) {
            if (bestDist <= M_TH_HIGH) {
                bool reject = false;
                if (v.s != 0xffffffffu) {
                    const uint32_t cs = rp_cand(cands, c0, v.s & 0xffffu, myc);
                    const int bestDist2 = (int)(v.s >> 16);
                    if ((cb >> 25) == (cs >> 25) && (float)bestDist > __fmul_rn(nn_ratio, (float)bestDist2)) reject = true;
                }
                // no second candidate: bestLevel2 = -1 never equals a real octave
                if (!reject) {
                    if (lane == 0) { match[bestIdx] = w; occ[bestIdx] = blocks[w]; }
                    ++nmatches;
                }
            }
        } else if (mode == 1) {
            if (bestDist <= M_TH_HIGH) {
                if (lane == 0) {
                    match[bestIdx] = w; occ[bestIdx] = blocks[w];
                    if (check_ori) { hist_idx[n_hist] = bestIdx; hist_bin[n_hist] = rot_bin(q_angle[w], F.angle[bestIdx]); }
                }
                ++nmatches; ++n_hist;
            }
        } else {
            if (bestDist <= max_dist) {
                if (lane == 0) {
                    match[bestIdx] = w; occ[bestIdx] = 1;
                    q_best_idx[w] = bestIdx; q_best_dist[w] = bestDist;
                    if (check_ori) { hist_idx[n_hist] = bestIdx; hist_bin[n_hist] = rot_bin(q_angle[w], F.angle[bestIdx]); }
                }
                ++nmatches; ++n_hist;
            }
        }
        __syncwarp();
        };
        // Four points at a time, 8 lanes each, resolved against the claim state at the start of the group; the result stands
        // unless a later point of the group still had, among its free candidates, a keypoint that an earlier point of the group
        // closes (rare: then, or when a list is longer than 8, the four are resolved one after the other as before).
#pragma unroll
        for (int h = 0; h < RP_G / 4; ++h) {
            const int j = lane >> 3, t = lane & 7;
            int cnt[4];
            bool small = true;
#pragma unroll
            for (int jj = 0; jj < 4; ++jj) { cnt[jj] = p1[4 * h + jj] - p0[4 * h + jj]; small = small && cnt[jj] <= 8; }
            bool done = false;
            if (small) {
                uint32_t c = 0xffffffffu; int n = 0;
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) { const uint32_t vv = __shfl_sync(0xffffffffu, pre[4 * h + jj], t); if (j == jj) { c = vv; n = cnt[jj]; } }
                const int w = base + 4 * h + j;
                const bool fr = t < n && !occ[c & 0xffffu];
                Best2 v; v.b = fr ? ((((c >> 16) & 0x1ffu) << 16) | (unsigned)t) : 0xffffffffu; v.s = 0xffffffffu;
#pragma unroll
                for (int o = 4; o; o >>= 1) { Best2 x; x.b = __shfl_xor_sync(0xffffffffu, v.b, o); x.s = __shfl_xor_sync(0xffffffffu, v.s, o); v = merge2(v, x); }
                const bool any = v.b != 0xffffffffu;
                const int bestDist = (int)(v.b >> 16);
                const uint32_t cb = __shfl_sync(0xffffffffu, c, (lane & 24) + (int)(v.b & 7u));
                const uint32_t cs = __shfl_sync(0xffffffffu, c, (lane & 24) + (int)(v.s & 7u));
                const int bestIdx = (int)(cb & 0xffffu);
                bool accept = false;
                if (any) {
                    if (mode == 0) {
                        accept = bestDist <= M_TH_HIGH;
                        if (accept && v.s != 0xffffffffu && (cb >> 25) == (cs >> 25) && (float)bestDist > __fmul_rn(nn_ratio, (float)(int)(v.s >> 16))) accept = false;
                    } else if (mode == 1) accept = bestDist <= M_TH_HIGH;
                    else accept = bestDist <= max_dist;
                }
                const int blk = accept ? (mode == 2 ? 1 : (int)blocks[w]) : 0;
                const int claim = (accept && blk) ? bestIdx : -1;
                bool conflict = false;
#pragma unroll
                for (int i = 0; i < 3; ++i) { const int ci = __shfl_sync(0xffffffffu, claim, 8 * i); if (j > i && fr && ci >= 0 && (int)(c & 0xffffu) == ci) conflict = true; }
                if (!__any_sync(0xffffffffu, conflict)) {
                    done = true;
                    const unsigned acc = __ballot_sync(0xffffffffu, accept && t == 0);
                    const int slot = n_hist + __popc(acc & ((1u << lane) - 1u));
#pragma unroll
                    for (int i = 0; i < 4; ++i) {     // in point order: a later point overwrites an earlier, non-closing one on the same keypoint
                        if (j == i && t == 0 && accept) {
                            match[bestIdx] = w; occ[bestIdx] = (uint8_t)blk;
                            if (mode == 2) { q_best_idx[w] = bestIdx; q_best_dist[w] = bestDist; }
                            if (mode >= 1 && check_ori) { hist_idx[slot] = bestIdx; hist_bin[slot] = rot_bin(q_angle[w], F.angle[bestIdx]); }
                        }
                        __syncwarp();
                    }
                    nmatches += __popc(acc);
                    if (mode >= 1) n_hist += __popc(acc);
                }
            }
            if (!done) {
#pragma unroll
                for (int jj = 0; jj < 4; ++jj) serial_one(base + 4 * h + jj, p0[4 * h + jj], p1[4 * h + jj], pre[4 * h + jj]);
            }
        }
#pragma unroll
        for (int k = 0; k < RP_G; ++k) { pre[k] = nxt[k]; p0[k] = n0[k]; p1[k] = n1[k]; }
    }
    if (lane == 0) *nmatches_out = nmatches;
    __syncwarp();
    if (mode >= 1 && check_ori) rot_cull_warp(hist_idx, hist_bin, n_hist, match, nmatches_out, s_cnt);
}

// The same resolve for ONE frame pair per call (a live Tracking thread), in parallel.  The reference's loop is sequential only through
// the claims: point w sees a keypoint as closed iff the caller closed it or an EARLIER point u < w claimed it.  So the decisions are the
// fixpoint of  decision[w] = f(first_claim[k] < w for the candidates k of w),  first_claim[k] = min { u : decision[u] claims k }:
// every round all points decide side by side against the claims of the previous round (one thread per point: lists hold a few
// candidates), the claims are rebuilt with atomicMin, and the loop ends when no decision changed.  After round r the decisions of the
// points 0 .. r - 1 are final, so it terminates; on real data three to five rounds.  match[k] is the LAST accepted point on k (a later
// point overwrites an earlier, non-closing one), the rotation histogram does not depend on the order of the matches.
#define RPP_THREADS 1024
__global__ void __launch_bounds__(RPP_THREADS)
k_resolve_projection_par(int mode, int nq, const int* __restrict__ starts, const uint32_t* __restrict__ cands, const uint8_t* __restrict__ blocks,
                         const float* __restrict__ q_angle, DFrame F, const uint8_t* __restrict__ occupied, float nn_ratio, int check_ori,
                         int max_dist, int* __restrict__ match, int* __restrict__ q_best_idx, int* __restrict__ q_best_dist,
                         int* __restrict__ hist_idx, int* __restrict__ hist_bin, int* __restrict__ nmatches_out, int* __restrict__ first_claim,
                         int* __restrict__ decision, const int cap) {
    __shared__ int s_cnt[32];
    __shared__ int s_changed, s_nm, s_nh;
    ORB_PDL_SYNC();
    if (cap > 0 && starts[nq] > cap) return;   // the candidate lists were not written (see k_area_candidates)
    const int tid = threadIdx.x;
    for (int i = tid; i < F.n; i += RPP_THREADS) { first_claim[i] = 0x7fffffff; match[i] = -1; }
    for (int w = tid; w < nq; w += RPP_THREADS) decision[w] = -2;   // no decision yet
    if (tid == 0) { s_nm = 0; s_nh = 0; }
    __syncthreads();
    for (;;) {
        if (tid == 0) s_changed = 0;
        __syncthreads();
        for (int w = tid; w < nq; w += RPP_THREADS) {
            const int c0 = starts[w], c1 = starts[w + 1];
            unsigned b = 0xffffffffu, sc = 0xffffffffu;
            for (int p = c0; p < c1; ++p) {
                const uint32_t c = cands[p];
                const int k = (int)(c & 0xffffu);
                if ((occupied && occupied[k]) || first_claim[k] < w) continue;
                const unsigned key = (((c >> 16) & 0x1ffu) << 16) | (unsigned)(p - c0);
                if (key < b) { sc = b; b = key; } else if (key < sc) sc = key;
            }
            int dec = -1;
            if (b != 0xffffffffu) {
                const int bestDist = (int)(b >> 16);
                const uint32_t cb = cands[c0 + (b & 0xffffu)];
                bool accept;
                if (mode == 0) {
                    accept = bestDist <= M_TH_HIGH;
                    if (accept && sc != 0xffffffffu) {
                        const uint32_t cs = cands[c0 + (sc & 0xffffu)];
                        if ((cb >> 25) == (cs >> 25) && (float)bestDist > __fmul_rn(nn_ratio, (float)(int)(sc >> 16))) accept = false;
                    }
                } else if (mode == 1) accept = bestDist <= M_TH_HIGH;
                else accept = bestDist <= max_dist;
                if (accept) {
                    const int blk = mode == 2 ? 1 : (int)blocks[w];
                    dec = (int)(cb & 0xffffu) | (blk ? 0x40000000 : 0) | (bestDist << 16);   // keypoint | distance (9 bits) | closing flag
                }
            }
            if (dec != decision[w]) { decision[w] = dec; s_changed = 1; }
        }
        __syncthreads();
        if (!s_changed) break;
        for (int i = tid; i < F.n; i += RPP_THREADS) first_claim[i] = 0x7fffffff;
        __syncthreads();
        for (int w = tid; w < nq; w += RPP_THREADS) {
            const int dec = decision[w];
            if (dec >= 0 && (dec & 0x40000000)) atomicMin(&first_claim[dec & 0xffff], w);
        }
        __syncthreads();
    }
    // outputs
    for (int w = tid; w < nq; w += RPP_THREADS) {
        const int dec = decision[w];
        if (mode == 2) { q_best_idx[w] = dec >= 0 ? (dec & 0xffff) : -1; q_best_dist[w] = dec >= 0 ? ((dec >> 16) & 0x1ff) : 256; }
        if (dec < 0) continue;
        const int k = dec & 0xffff;
        atomicMax(&match[k], w);
        atomicAdd(&s_nm, 1);
        if (mode >= 1 && check_ori) {
            const int slot = atomicAdd(&s_nh, 1);
            hist_idx[slot] = k; hist_bin[slot] = rot_bin(q_angle[w], F.angle[k]);
        }
    }
    __syncthreads();
    if (tid == 0) *nmatches_out = s_nm;
    __syncthreads();
    if (tid < 32 && mode >= 1 && check_ori) rot_cull_warp(hist_idx, hist_bin, s_nh, match, nmatches_out, s_cnt);
}

// Projected points WITHOUT claiming (SearchBySim3 :1212-1455 per direction, Fuse :903-1207): every point is
// independent; one warp per point takes the first minimum-distance candidate (strict <).
__global__ void __launch_bounds__(256)
k_best_independent(int nq, const int* __restrict__ starts, const uint32_t* __restrict__ cands, const uint8_t* __restrict__ occupied,
                   int max_dist, int* __restrict__ q_best_idx, int* __restrict__ q_best_dist, int* __restrict__ nmatches_out, const int cap = 0) {
    ORB_PDL_SYNC();
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= nq || (cap > 0 && starts[nq] > cap)) return;
    const int c0 = starts[w], c1 = starts[w + 1];
    unsigned best = 0xffffffffu;
    for (int p = c0 + lane; p < c1; p += 32) {
        const uint32_t c = cands[p];
        if (occupied && occupied[c & 0xffffu]) continue;
        best = min(best, (((c >> 16) & 0x1ffu) << 16) | (unsigned)min(p - c0, 0xffff));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0) {
        int bi = -1, bd = 256;
        if (best != 0xffffffffu) {
            bd = (int)(best >> 16);
            if (bd <= max_dist) { bi = (int)(cands[c0 + (best & 0xffffu)] & 0xffffu); atomicAdd(nmatches_out, 1); }
        }
        q_best_idx[w] = bi; q_best_dist[w] = bd;
    }
}

// ------------------------------------------------------------------------------------------
// BoW-guided searches.  Work entries (built on the host by the FeatureVector merge-join, in the
// reference's processing order): query feature, candidate range [f0, f1) inside items2.
struct BowEntry { int q_idx, f0, f1, c_start; };

__global__ void __launch_bounds__(256)
k_bow_distances(int n_entries, const BowEntry* __restrict__ e, const uint8_t* __restrict__ qdesc, const uint8_t* __restrict__ fdesc,
                const uint32_t* __restrict__ items2, uint32_t* __restrict__ cands) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n_entries) return;
    const BowEntry E = e[w];
    const uint4* pa = reinterpret_cast<const uint4*>(qdesc + (size_t)E.q_idx * 32);
    const uint4 a0 = pa[0], a1 = pa[1];
    for (int p = E.f0 + lane; p < E.f1; p += 32) {
        const uint32_t idx = items2[p];
        const uint4* pb = reinterpret_cast<const uint4*>(fdesc + (size_t)idx * 32);
        cands[E.c_start + (p - E.f0)] = idx | ((uint32_t)orb_hamming256(a0, a1, pb[0], pb[1]) << 16);
    }
}

// Resolve of the BoW searches.  A keypoint of the second frame belongs to exactly one vocabulary node, so the claims of the reference's
// loop (vpMapPointMatches[realIdxF] / vbMatched2[idx2]: a feature taken by an earlier query is skipped by later ones) never cross
// node pairs: one warp per node pair walks that pair's queries in the reference's order; node pairs run side by side.  The accepted
// matches land in hist_idx[entry] (-1: none; memset by the caller), k_bow_finish compacts them, counts and applies the rotation
// histogram (whose outcome does not depend on the order in which matches were found).
#define BOW_WARPS 8
__global__ void __launch_bounds__(BOW_WARPS * 32)
k_resolve_bow(int n_nodes, const int* __restrict__ node_start, const BowEntry* __restrict__ e, const uint32_t* __restrict__ cands,
              const float* __restrict__ kf_angle, const float* __restrict__ f_angle, float nn_ratio, int check_ori, int* __restrict__ match_f,
              int* __restrict__ hist_idx, int* __restrict__ hist_bin) {
    const int lane = threadIdx.x & 31, k = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5);
    if (k >= n_nodes) return;
    const int w1 = node_start[k + 1];
    for (int w = node_start[k]; w < w1; ++w) {
        const BowEntry E = e[w];
        const int cnt = E.f1 - E.f0;
        Best2 v; v.b = 0xffffffffu; v.s = 0xffffffffu;
        for (int p = lane; p < cnt; p += 32) {
            const uint32_t c = cands[E.c_start + p];
            if (match_f[c & 0xffffu] >= 0) continue;
            Best2 t; t.b = (((c >> 16) & 0x1ffu) << 16) | (unsigned)p; t.s = 0xffffffffu;
            v = merge2(v, t);
        }
        v = warp_best2(v);
        if (v.b == 0xffffffffu) continue;
        const int d1 = (int)(v.b >> 16), d2 = v.s == 0xffffffffu ? 256 : (int)(v.s >> 16);
        if (d1 <= M_TH_LOW && (float)d1 < __fmul_rn(nn_ratio, (float)d2)) {
            const int idxF = (int)(cands[E.c_start + (v.b & 0xffffu)] & 0xffffu);
            if (lane == 0) {
                match_f[idxF] = E.q_idx;
                hist_idx[w] = idxF;
                if (check_ori) hist_bin[w] = rot_bin(kf_angle[E.q_idx], f_angle[idxF]);
            }
        }
        __syncwarp();
    }
}

// SearchByBoW(KF1, KF2) (reference :581-716): as k_resolve_bow but both features need a MapPoint, KF2
// features are claimed through vbMatched2 (which the rotation cull does NOT release), the threshold is
// strict (< TH_LOW) and the result is indexed by the KF1 feature.
__global__ void __launch_bounds__(BOW_WARPS * 32)
k_resolve_bow_kf(int n_nodes, const int* __restrict__ node_start, const BowEntry* __restrict__ e, const uint32_t* __restrict__ cands,
                 const float* __restrict__ angle1, const float* __restrict__ angle2, uint8_t* __restrict__ closed2, float nn_ratio,
                 int check_ori, int* __restrict__ match12, int* __restrict__ hist_idx, int* __restrict__ hist_bin) {
    const int lane = threadIdx.x & 31, k = blockIdx.x * BOW_WARPS + (threadIdx.x >> 5);
    if (k >= n_nodes) return;
    const int w1 = node_start[k + 1];
    for (int w = node_start[k]; w < w1; ++w) {
        const BowEntry E = e[w];
        const int cnt = E.f1 - E.f0;
        Best2 v; v.b = 0xffffffffu; v.s = 0xffffffffu;
        for (int p = lane; p < cnt; p += 32) {
            const uint32_t c = cands[E.c_start + p];
            if (closed2[c & 0xffffu]) continue;
            Best2 t; t.b = (((c >> 16) & 0x1ffu) << 16) | (unsigned)p; t.s = 0xffffffffu;
            v = merge2(v, t);
        }
        v = warp_best2(v);
        if (v.b == 0xffffffffu) continue;
        const int d1 = (int)(v.b >> 16), d2 = v.s == 0xffffffffu ? 256 : (int)(v.s >> 16);
        if (d1 < M_TH_LOW && (float)d1 < __fmul_rn(nn_ratio, (float)d2)) {
            const int idx2 = (int)(cands[E.c_start + (v.b & 0xffffu)] & 0xffffu);
            if (lane == 0) {
                match12[E.q_idx] = idx2; closed2[idx2] = 1;
                hist_idx[w] = E.q_idx;
                if (check_ori) hist_bin[w] = rot_bin(angle1[E.q_idx], angle2[idx2]);
            }
        }
        __syncwarp();
    }
}
__global__ void k_bow_closed_init(int n2, const uint8_t* __restrict__ has2, uint8_t* __restrict__ closed2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n2) closed2[i] = has2[i] ? 0 : 1;   // !pMP2 (or bad) -> never a candidate
}
// One warp: compacts the per-entry matches (ascending entry == the reference's order), counts them, rotation histogram.
__global__ void __launch_bounds__(32)
k_bow_finish(int n_entries, int* __restrict__ hist_idx, int* __restrict__ hist_bin, int check_ori, int* __restrict__ match,
             int* __restrict__ nmatches_out) {
    __shared__ int s_cnt[32];
    const int lane = threadIdx.x;
    int n = 0;
    for (int w0 = 0; w0 < n_entries; w0 += 32) {
        const int w = w0 + lane;
        const int idx = w < n_entries ? hist_idx[w] : -1;
        const int bin = (check_ori && idx >= 0) ? hist_bin[w] : 0;
        const unsigned m = __ballot_sync(0xffffffffu, idx >= 0);   // also orders this round's reads before the writes below
        if (idx >= 0) { const int o = n + __popc(m & ((1u << lane) - 1)); hist_idx[o] = idx; hist_bin[o] = bin; }
        n += __popc(m);
        __syncwarp();
    }
    if (lane == 0) *nmatches_out = n;
    __syncwarp();
    if (check_ori) rot_cull_warp(hist_idx, hist_bin, n, match, nmatches_out, s_cnt);
}

// SearchForInitialization (reference :429-578): sequential over the level-0 keypoints of F1; a candidate i2 is
// skipped when an earlier match holds it with a distance <= the new one (vMatchedDistance), otherwise the
// earlier match is displaced.  Displaced entries stay in the rotation histogram (they count towards the bin
// sizes) and the cull only clears entries that are still matched (:556-562).
__global__ void __launch_bounds__(32)
k_resolve_initialization(int nq, const int* __restrict__ starts, const uint32_t* __restrict__ cands, const float* __restrict__ angle1,
                         const float* __restrict__ angle2, int n2, float nn_ratio, int check_ori, int* __restrict__ match12,
                         int* __restrict__ match21, int* __restrict__ mdist, int* __restrict__ hist_idx, int* __restrict__ hist_bin,
                         int* __restrict__ nmatches_out) {
    __shared__ int s_cnt[32];
    const int lane = threadIdx.x;
    for (int i = lane; i < nq; i += 32) match12[i] = -1;
    for (int i = lane; i < n2; i += 32) { match21[i] = -1; mdist[i] = 0x7fffffff; }
    __syncwarp();
    int nmatches = 0, n_hist = 0;
    for (int w = 0; w < nq; ++w) {
        const int c0 = starts[w], c1 = starts[w + 1];
        if (c0 == c1) continue;
        Best2 v; v.b = 0xffffffffu; v.s = 0xffffffffu;
        for (int p = c0 + lane; p < c1; p += 32) {
            const uint32_t c = cands[p];
            const int d = (int)((c >> 16) & 0x1ffu);
            if (mdist[c & 0xffffu] <= d) continue;
            Best2 t; t.b = ((unsigned)d << 16) | (unsigned)(p - c0); t.s = 0xffffffffu;
            v = merge2(v, t);
        }
        v = warp_best2(v);
        if (v.b == 0xffffffffu) continue;
        const int bestDist = (int)(v.b >> 16);
        const float second = v.s == 0xffffffffu ? 2147483648.f : (float)(int)(v.s >> 16);   // (float)INT_MAX
        if (bestDist <= M_TH_LOW && (float)bestDist < __fmul_rn(second, nn_ratio)) {
            const int i2 = (int)(cands[c0 + (v.b & 0xffffu)] & 0xffffu);
            const int old = match21[i2];
            if (lane == 0) {
                if (old >= 0) match12[old] = -1;
                match12[w] = i2; match21[i2] = w; mdist[i2] = bestDist;
                if (check_ori) { hist_idx[n_hist] = w; hist_bin[n_hist] = rot_bin(angle1[w], angle2[i2]); }
            }
            if (old >= 0) --nmatches;
            ++nmatches; ++n_hist;
        }
        __syncwarp();
    }
    __syncwarp();
    if (check_ori) {
        if (lane < M_HISTO + 2) s_cnt[lane] = 0;
        __syncwarp();
        for (int i = lane; i < n_hist; i += 32) atomicAdd(&s_cnt[hist_bin[i]], 1);
        __syncwarp();
        int i1, i2, i3;
        three_maxima(s_cnt, i1, i2, i3);
        int removed = 0;
        for (int i = lane; i < n_hist; i += 32) {
            const int b = hist_bin[i];
            if (b != i1 && b != i2 && b != i3 && match12[hist_idx[i]] >= 0) { match12[hist_idx[i]] = -1; ++removed; }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) removed += __shfl_xor_sync(0xffffffffu, removed, o);
        nmatches -= removed;
    }
    if (lane == 0) *nmatches_out = nmatches;
}

// SearchBySim3 agreement (reference :1423-1439): keep i1 -> i2 only when i2 -> i1
__global__ void k_mutual_check(int n1, int n2, const int* __restrict__ m1, const int* __restrict__ m2, int* __restrict__ match12,
                               int* __restrict__ nfound) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n1) return;
    const int i2 = m1[i];
    int r = -1;
    if (i2 >= 0 && i2 < n2 && m2[i2] == i) { r = i2; atomicAdd(nfound, 1); }
    match12[i] = r;
}

// SearchForTriangulation: no claiming in the reference (vbMatched2 is never set), so every query
// is independent.  The sequential rule "skip if dist > TH_LOW or dist > best; accept if the
// epipolar tests pass" ends on the LAST candidate holding the minimum distance among those that
// pass the tests.
struct TriParams {
    float f12[9];
    float ex, ey;
    float sigma2[ORB_MAX_LEVELS];
    int only_stereo;
};
__global__ void __launch_bounds__(256)
k_triangulation(int n_entries, const BowEntry* __restrict__ e, DFrame K1, DFrame K2, const uint32_t* __restrict__ items2,
                const uint8_t* __restrict__ has_mp2, TriParams T, int* __restrict__ match12) {
    const int w = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (w >= n_entries) return;
    const BowEntry E = e[w];
    const int idx1 = E.q_idx;
    const bool stereo1 = K1.u_right && K1.u_right[idx1] >= 0;
    const uint4* pa = reinterpret_cast<const uint4*>(K1.desc + (size_t)idx1 * 32);
    const uint4 a0 = pa[0], a1 = pa[1];
    const float x1 = K1.x[idx1], y1 = K1.y[idx1];
    // epipolar line of kp1 in image 2 (reference :146-148), GCC-style contraction
    const float la = __fadd_rn(__fmaf_rn(x1, T.f12[0], __fmul_rn(y1, T.f12[3])), T.f12[6]);
    const float lb = __fadd_rn(__fmaf_rn(x1, T.f12[1], __fmul_rn(y1, T.f12[4])), T.f12[7]);
    const float lc = __fadd_rn(__fmaf_rn(x1, T.f12[2], __fmul_rn(y1, T.f12[5])), T.f12[8]);
    const float den = __fmaf_rn(la, la, __fmul_rn(lb, lb));
    unsigned best = 0xffffffffu;  // (dist << 16) | (0xffff - position): min -> smallest dist, last position
    for (int p = E.f0 + lane; p < E.f1; p += 32) {
        const int idx2 = (int)items2[p];
        if (has_mp2[idx2]) continue;
        const bool stereo2 = K2.u_right && K2.u_right[idx2] >= 0;
        if (T.only_stereo && !stereo2) continue;
        const uint4* pb = reinterpret_cast<const uint4*>(K2.desc + (size_t)idx2 * 32);
        const int dist = orb_hamming256(a0, a1, pb[0], pb[1]);
        if (dist > M_TH_LOW) continue;
        const float x2 = K2.x[idx2], y2 = K2.y[idx2];
        const int oct2 = K2.octave[idx2];
        if (!stereo1 && !stereo2) {
            const float dex = __fsub_rn(T.ex, x2), dey = __fsub_rn(T.ey, y2);
            if (__fmaf_rn(dex, dex, __fmul_rn(dey, dey)) < __fmul_rn(100.f, K2.sf[oct2])) continue;
        }
        const float num = __fadd_rn(__fmaf_rn(la, x2, __fmul_rn(lb, y2)), lc);
        if (den == 0) continue;
        const float dsqr = __fdiv_rn(__fmul_rn(num, num), den);
        if (!((double)dsqr < 3.84 * (double)T.sigma2[oct2])) continue;
        best = min(best, ((unsigned)dist << 16) | (unsigned)(0xffff - (p - E.f0)));
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    if (lane == 0 && best != 0xffffffffu) match12[idx1] = (int)items2[E.f0 + (0xffff - (int)(best & 0xffffu))];
}

__global__ void __launch_bounds__(32)
k_tri_rot_cull(int n1, const float* __restrict__ a1, const float* __restrict__ a2, int check_ori, int* __restrict__ match12,
               int* __restrict__ hist_idx, int* __restrict__ hist_bin, int* __restrict__ nmatches_out) {
    __shared__ int s_cnt[32];
    const int lane = threadIdx.x;
    int n = 0;
    for (int i0 = 0; i0 < n1; i0 += 32) {  // ordered compaction of matched idx1 (order is irrelevant to the cull)
        const int i = i0 + lane;
        const int m = i < n1 ? match12[i] : -1;
        const unsigned bal = __ballot_sync(0xffffffffu, m >= 0);
        if (m >= 0 && check_ori) {
            const int pos = n + __popc(bal & ((1u << lane) - 1));
            hist_idx[pos] = i; hist_bin[pos] = rot_bin(a1[i], a2[m]);
        }
        n += __popc(bal);
    }
    if (lane == 0) *nmatches_out = n;
    __syncwarp();
    if (check_ori) rot_cull_warp(hist_idx, hist_bin, n, match12, nmatches_out, s_cnt);
}

// ------------------------------------------------------------------------------------------ host helpers
int upload_frame(orb_matcher* m, const orb_frame_view* f, DFrame* D, bool need_grid) {
    if (!f || f->n < 0 || f->n > m->max_n || f->n > 65535 || (f->n > 0 && (!f->x || !f->y || !f->octave || !f->desc))) {
        orb_set_error("bad frame view (n must be <= min(max_n, 65535))");
        return ORB_ERR_BAD_ARG;
    }
    std::memset(D, 0, sizeof(*D));
    D->n = f->n; D->grid = f->grid;
    for (int i = 0; i < ORB_MAX_LEVELS; ++i) D->sf[i] = (f->scale_factors && i < f->n_levels) ? f->scale_factors[i] : 1.f;
    int st;
    if ((st = ws_upload(m, &D->x, f->x, f->n)) != ORB_OK) return st;
    if ((st = ws_upload(m, &D->y, f->y, f->n)) != ORB_OK) return st;
    if ((st = ws_upload(m, &D->octave, f->octave, f->n)) != ORB_OK) return st;
    if (f->angle) { if ((st = ws_upload(m, &D->angle, f->angle, f->n)) != ORB_OK) return st; }
    if (f->u_right) { if ((st = ws_upload(m, &D->u_right, f->u_right, f->n)) != ORB_OK) return st; }
    if ((st = ws_upload(m, &D->desc, f->desc, (size_t)f->n * 32)) != ORB_OK) return st;
    if (need_grid) {
        int *cs, *ci, *ni;
        if ((st = ws_alloc(m, &cs, GRID_CELLS + 1)) != ORB_OK) return st;
        if ((st = ws_alloc(m, &ci, std::max(f->n, 1))) != ORB_OK) return st;
        if ((st = ws_alloc(m, &ni, 1)) != ORB_OK) return st;
        int p2 = 1;
        while (p2 < std::max(f->n, 1)) p2 <<= 1;
        const size_t smem = (size_t)p2 * 4;
        if (smem > GRID_BUILD_DYN_DEFAULT) ORB_CUDA(cudaFuncSetAttribute(k_grid_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        ORB_CUDA(orb_launch(false, k_grid_build, dim3(1), dim3(1024), smem, ws_stream(m), D->x, D->y, f->n, p2, f->grid, cs, ci, ni, 0, (const int*)nullptr));
        m->launches++;
        D->cell_start = cs; D->cell_items = ci;
    }
    return ORB_OK;
}

// FeatureVector merge-join (reference :187-282): matched node pairs in ascending node id
void merge_join(const orb_featvec_view* a, const orb_featvec_view* b, std::vector<std::pair<int, int>>& out) {
    int i = 0, j = 0;
    while (i < a->n_nodes && j < b->n_nodes) {
        if (a->node_ids[i] == b->node_ids[j]) { out.push_back(std::make_pair(i, j)); ++i; ++j; }
        else if (a->node_ids[i] < b->node_ids[j]) i = (int)(std::lower_bound(a->node_ids, a->node_ids + a->n_nodes, b->node_ids[j]) - a->node_ids);
        else j = (int)(std::lower_bound(b->node_ids, b->node_ids + b->n_nodes, a->node_ids[i]) - b->node_ids);
    }
}

// The candidate list buffer grows to the size the counting pass found (the caller has synchronised the stream).
int ensure_cands(orb_matcher* m, size_t total) {
    if (total <= m->cands_cap) return ORB_OK;
    const size_t want = total + total / 4 + 4096;
    uint32_t* p = nullptr;
    if (orb_dev_alloc((void**)&p, want * 4) != cudaSuccess) { cudaGetLastError(); orb_set_error("candidate workspace: out of device memory"); return ORB_ERR_CAPACITY; }
    orb_dev_free(m->d_cands);
    m->d_cands = p; m->cands_cap = want;
    return ORB_OK;
}

int begin_call(orb_matcher* m) {
    if (!m) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(m->device));
    m->ws_off = 0;
    m->pend.clear();
    m->deliver.clear();
    if (m->stage_busy) { ORB_CUDA(cudaEventSynchronize(m->ev_stage)); m->stage_busy = false; }   // the mirror is about to be rewritten
    return ORB_OK;
}

}  // namespace

extern "C" {

int orb_matcher_create(int max_pairs, int max_n, int device, orb_matcher** out) {
    if (!out) return ORB_ERR_BAD_ARG;
    *out = nullptr;
    if (max_pairs < 1 || max_n < 1) { orb_set_error("bad matcher capacities"); return ORB_ERR_BAD_ARG; }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        orb_set_error("no CUDA device: this library has no CPU fallback");
        return ORB_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) { orb_set_error("bad device index"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(device));
    orb_matcher* m = new orb_matcher();
    m->device = device; m->max_pairs = max_pairs; m->max_n = max_n;
    // per pair and feature: frame SoA (52 B) + grid items, point SoA (54 B), query records, offsets, matches, histogram slots;
    // brute force: 2 x 32 B + 3 ints.  Candidate lists start at 64 per feature and grow on demand (ensure_cands).
    m->ws_cap = (size_t)max_pairs * max_n * 256 + (size_t)max_pairs * 16 * 1024 + (size_t)max_n * 512 + (8u << 20);
    m->cands_cap = std::max<size_t>((size_t)max_pairs * max_n * 64, 1 << 16);
    if (cudaStreamCreateWithFlags(&m->stream, cudaStreamNonBlocking) != cudaSuccess || orb_dev_alloc((void**)&m->d_ws, m->ws_cap) != cudaSuccess ||
        orb_dev_alloc((void**)&m->d_cands, m->cands_cap * 4) != cudaSuccess || orb_dev_alloc((void**)&m->d_scal, 64) != cudaSuccess) {
        orb_set_error(std::string("matcher allocation: ") + cudaGetErrorString(cudaGetLastError()));
        orb_matcher_destroy(m);
        return ORB_ERR_CUDA;
    }
    // pinned mirror of the head of the workspace (single-frame calls place their uploads there); optional: without it uploads go direct
    m->h_stage_cap = std::min<size_t>(m->ws_cap, 4u << 20);
    if (cudaMallocHost((void**)&m->h_stage, m->h_stage_cap) != cudaSuccess ||
        cudaEventCreateWithFlags(&m->ev_stage, cudaEventDisableTiming) != cudaSuccess) {
        cudaGetLastError();
        if (m->h_stage) { cudaFreeHost(m->h_stage); m->h_stage = nullptr; }
        m->h_stage_cap = 0;
    }
    if (cudaMallocHost((void**)&m->h_scal, 64) != cudaSuccess) { cudaGetLastError(); m->h_scal = nullptr; }
    *out = m;
    return ORB_OK;
}

void orb_matcher_destroy(orb_matcher* m) {
    if (!m) return;
    cudaSetDevice(m->device);
    if (m->stream) cudaStreamSynchronize(m->stream);
    if (m->h_stage) cudaFreeHost(m->h_stage);
    if (m->h_scal) cudaFreeHost(m->h_scal);
    if (m->ev_stage) cudaEventDestroy(m->ev_stage);
    orb_dev_free(m->d_ws); orb_dev_free(m->d_cands); orb_dev_free(m->d_scal);
    if (m->stream) cudaStreamDestroy(m->stream);
    delete m;
}

void* orb_matcher_stream(const orb_matcher* m) { return m ? (void*)m->stream : nullptr; }
long long orb_matcher_launch_count(const orb_matcher* m) { return m ? m->launches : 0; }
long long orb_matcher_last_candidates(const orb_matcher* m) { return m ? m->last_cands : 0; }
int orb_matcher_sync(orb_matcher* m) {
    if (!m) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(m->device));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    return ORB_OK;
}

int orb_descriptor_distance(orb_matcher* m, const uint8_t* a, const uint8_t* b, int n, int32_t* dist) {
    ORB_NVTX("orb:descriptor_distance");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!a || !b || !dist || n < 0) return ORB_ERR_BAD_ARG;
    if (n == 0) return ORB_OK;
    const uint8_t *da, *db; int* dd;
    if ((st = ws_upload(m, &da, a, (size_t)n * 32)) != ORB_OK) return st;
    if ((st = ws_upload(m, &db, b, (size_t)n * 32)) != ORB_OK) return st;
    if ((st = ws_alloc(m, &dd, n)) != ORB_OK) return st;
    k_distance_pairs<<<(n + 255) / 256, 256, 0, ws_stream(m)>>>(da, db, n, dd);
    m->launches++;
    ORB_CUDA(ws_download(m, dist, dd, (size_t)n * 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_match_bruteforce_device(orb_matcher* m, const uint8_t* d_a, const uint8_t* d_b, int pairs, int n_a, int n_b, int32_t* d_best_idx,
                                int32_t* d_best_dist, int32_t* d_second_dist) {
    ORB_NVTX("orb:match_bruteforce_device");
    if (!m || !d_a || !d_b || !d_best_idx || !d_best_dist || !d_second_dist || pairs < 1 || n_a < 1 || n_b < 0) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(m->device));
    dim3 grid((n_a + BF_THREADS - 1) / BF_THREADS, pairs);
    k_bruteforce<<<grid, BF_THREADS, 0, ws_stream(m)>>>(d_a, d_b, n_a, n_b, d_best_idx, d_best_dist, d_second_dist);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    return ORB_OK;
}

int orb_match_bruteforce(orb_matcher* m, const uint8_t* a, const uint8_t* b, int pairs, int n_a, int n_b, int32_t* best_idx,
                         int32_t* best_dist, int32_t* second_dist) {
    ORB_NVTX("orb:match_bruteforce");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!a || !b || !best_idx || !best_dist || !second_dist || pairs < 1 || n_a < 1 || n_b < 0) return ORB_ERR_BAD_ARG;
    const uint8_t *da, *db; int *bi, *bd, *sd;
    const size_t na = (size_t)pairs * n_a, nb = (size_t)pairs * n_b;
    if ((st = ws_upload(m, &da, a, na * 32)) != ORB_OK) return st;
    if ((st = ws_upload(m, &db, b, nb * 32)) != ORB_OK) return st;
    if ((st = ws_alloc(m, &bi, na)) != ORB_OK || (st = ws_alloc(m, &bd, na)) != ORB_OK || (st = ws_alloc(m, &sd, na)) != ORB_OK) return st;
    st = orb_match_bruteforce_device(m, da, db, pairs, n_a, n_b, bi, bd, sd);
    if (st != ORB_OK) return st;
    ORB_CUDA(ws_download(m, best_idx, bi, na * 4));
    ORB_CUDA(ws_download(m, best_dist, bd, na * 4));
    ORB_CUDA(ws_download(m, second_dist, sd, na * 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_assign_features_to_grid(orb_matcher* m, const float* kp_x, const float* kp_y, int n, const orb_grid_params* g, int32_t* cell_start,
                                int32_t* cell_items, int32_t* n_items) {
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!g || !cell_start || !cell_items || !n_items || n < 0 || n > 65535 || n > m->max_n || (n > 0 && (!kp_x || !kp_y))) return ORB_ERR_BAD_ARG;
    const float *dx, *dy; int *cs, *ci, *ni;
    if ((st = ws_upload(m, &dx, kp_x, n)) != ORB_OK || (st = ws_upload(m, &dy, kp_y, n)) != ORB_OK) return st;
    if ((st = ws_alloc(m, &cs, GRID_CELLS + 1)) != ORB_OK || (st = ws_alloc(m, &ci, std::max(n, 1))) != ORB_OK || (st = ws_alloc(m, &ni, 1)) != ORB_OK) return st;
    int p2 = 1;
    while (p2 < std::max(n, 1)) p2 <<= 1;
    const size_t smem = (size_t)p2 * 4;
    if (smem > GRID_BUILD_DYN_DEFAULT) ORB_CUDA(cudaFuncSetAttribute(k_grid_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_grid_build<<<1, 1024, smem, ws_stream(m)>>>(dx, dy, n, p2, *g, cs, ci, ni);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    ORB_CUDA(cudaMemcpyAsync(cell_start, cs, (GRID_CELLS + 1) * 4, cudaMemcpyDeviceToHost, m->stream));
    ORB_CUDA(cudaMemcpyAsync(n_items, ni, 4, cudaMemcpyDeviceToHost, m->stream));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    if (*n_items > 0) {
        ORB_CUDA(cudaMemcpyAsync(cell_items, ci, (size_t)*n_items * 4, cudaMemcpyDeviceToHost, m->stream));
        ORB_CUDA(cudaStreamSynchronize(m->stream));
    }
    return ORB_OK;
}

// shared tail of the projection searches: count -> scan -> fill -> resolve
struct ProjOpts {
    int mode = 0;                 // k_resolve_projection mode, or -1: independent points (k_best_independent)
    int filter = FILTER_STEREO_WINDOW;
    const float* d_inv_sigma2 = nullptr;
    float nn_ratio = 0.f;
    int check_ori = 0;
    int max_dist = M_TH_HIGH;
    int32_t* q_best_idx = nullptr;   // host outputs per query (modes 2 and -1)
    int32_t* q_best_dist = nullptr;
};

static int projection_common(orb_matcher* m, const ProjOpts& o, int nq, const Query* dq, const uint8_t* d_qdesc, const uint8_t* d_blocks,
                             const float* d_qangle, const DFrame& F, const uint8_t* occupied, int32_t* match, int32_t* nmatches) {
    int st;
    int *counts, *starts, *dmatch, *hidx, *hbin, *qbi = nullptr, *qbd = nullptr;
    uint8_t* occ; const uint8_t* d_occupied = nullptr;
    if ((st = ws_alloc(m, &counts, std::max(nq, 1))) != ORB_OK || (st = ws_alloc(m, &starts, nq + 1)) != ORB_OK) return st;
    if ((st = ws_alloc(m, &dmatch, std::max(F.n, 1))) != ORB_OK || (st = ws_alloc(m, &occ, std::max(F.n, 1))) != ORB_OK) return st;
    if ((st = ws_alloc(m, &hidx, std::max(nq, 1))) != ORB_OK || (st = ws_alloc(m, &hbin, std::max(nq, 1))) != ORB_OK) return st;
    if (o.mode == 2 || o.mode == -1) {
        if ((st = ws_alloc(m, &qbi, std::max(nq, 1))) != ORB_OK || (st = ws_alloc(m, &qbd, std::max(nq, 1))) != ORB_OK) return st;
    }
    if (occupied) { if ((st = ws_upload(m, &d_occupied, occupied, F.n)) != ORB_OK) return st; }
    const int filter = o.filter; const float* d_inv_sigma2 = o.d_inv_sigma2;
    int *first_claim = nullptr, *decision = nullptr;
    if (o.mode != -1) {
        // one frame pair per call: the parallel fixpoint resolve (the one-warp kernel serves the batched entry points, one warp per pair)
        if ((st = ws_alloc(m, &first_claim, std::max(F.n, 1))) != ORB_OK || (st = ws_alloc(m, &decision, std::max(nq, 1))) != ORB_OK) return st;
    }
    (void)occ;
    const int blocks = (nq * 32 + 255) / 256;
    if (nq > 0) {
        ORB_CUDA(orb_launch(true, k_area_candidates<false>, dim3(blocks), dim3(256), 0, ws_stream(m), nq, dq, d_qdesc, F, filter, d_inv_sigma2, counts, (const int*)nullptr,
                            (uint32_t*)nullptr, Batch{0, nullptr, nullptr}, 0));
        ORB_CUDA(orb_launch(true, k_scan, dim3(1), dim3(1024), 0, ws_stream(m), (const int*)counts, nq, starts, Batch{0, nullptr, nullptr}, (int*)nullptr));
        m->launches += 2;
    } else {
        ORB_CUDA(cudaMemsetAsync(starts, 0, 4, m->stream));
    }
    // The size of the candidate lists is only known on the device.  Instead of waiting for it in the middle of the call, the fill and the
    // resolve are queued against the buffer's present capacity (they do nothing when the lists do not fit) and the total comes back with
    // the results; only a call whose lists did not fit grows the buffer and queues the two steps again.
    int32_t total = 0;
    for (int attempt = 0; attempt < 2; ++attempt) {
        const int cap = (int)std::min<size_t>(m->cands_cap, 0x7fffffff);
        if (nq > 0) {
            ORB_CUDA(orb_launch(true, k_area_candidates<true>, dim3(blocks), dim3(256), 0, ws_stream(m), nq, dq, d_qdesc, F, filter, d_inv_sigma2, (int*)nullptr, (const int*)starts,
                                m->d_cands, Batch{0, nullptr, nullptr}, cap));
            m->launches++;
        }
        if (o.mode == -1) {
            ORB_CUDA(cudaMemsetAsync(m->d_scal, 0, 4, m->stream));
            if (nq > 0) {
                ORB_CUDA(orb_launch(true, k_best_independent, dim3(blocks), dim3(256), 0, ws_stream(m), nq, (const int*)starts, (const uint32_t*)m->d_cands, d_occupied, o.max_dist, qbi, qbd,
                                    m->d_scal, cap));
                m->launches++;
            }
        } else {
            ORB_CUDA(orb_launch(true, k_resolve_projection_par, dim3(1), dim3(RPP_THREADS), 0, ws_stream(m), o.mode, nq, (const int*)starts, (const uint32_t*)m->d_cands, d_blocks, d_qangle, F,
                                d_occupied, o.nn_ratio, o.check_ori, o.max_dist, dmatch, qbi, qbd, hidx, hbin, m->d_scal, first_claim, decision, cap));
            m->launches++;
        }
        ORB_CUDA(cudaGetLastError());
        if (match && o.mode != -1 && F.n > 0) ORB_CUDA(ws_download(m, match, dmatch, (size_t)F.n * 4));
        if (o.q_best_idx && nq > 0) ORB_CUDA(ws_download(m, o.q_best_idx, qbi, (size_t)nq * 4));
        if (o.q_best_dist && nq > 0) ORB_CUDA(ws_download(m, o.q_best_dist, qbd, (size_t)nq * 4));
        ORB_CUDA(ws_download(m, nmatches, m->d_scal, 4));
        ORB_CUDA(ws_download(m, &total, starts + nq, 4));
        ORB_CUDA(ws_finish(m));
        if ((size_t)total <= m->cands_cap) break;
        if (attempt == 1) { orb_set_error("candidate lists changed size between two passes"); return ORB_ERR_INTERNAL; }
        if ((st = ensure_cands(m, (size_t)total)) != ORB_OK) return st;
    }
    m->last_cands = total;
    return ORB_OK;
}

int orb_search_by_projection_mappoints(orb_matcher* m, const orb_frame_view* f, const orb_mappoint_view* mps, float th, float nn_ratio,
                                       const uint8_t* occupied, int32_t* match, int32_t* nmatches) {
    ORB_NVTX("orb:search_by_projection_mappoints");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!f || !mps || !match || !nmatches || mps->n < 0 || (mps->n > 0 && (!mps->valid || !mps->proj_x || !mps->proj_y || !mps->level ||
        !mps->view_cos || !mps->desc || !mps->blocks))) return ORB_ERR_BAD_ARG;
    DFrame F;
    if ((st = upload_frame(m, f, &F, true)) != ORB_OK) return st;
    const int nq = mps->n;
    const uint8_t *dv, *dd, *db; const float *dx, *dy, *dxr = nullptr, *dvc; const int* dl; Query* dq;
    if ((st = ws_upload(m, &dv, mps->valid, nq)) != ORB_OK || (st = ws_upload(m, &dx, mps->proj_x, nq)) != ORB_OK ||
        (st = ws_upload(m, &dy, mps->proj_y, nq)) != ORB_OK || (st = ws_upload(m, &dl, mps->level, nq)) != ORB_OK ||
        (st = ws_upload(m, &dvc, mps->view_cos, nq)) != ORB_OK || (st = ws_upload(m, &dd, mps->desc, (size_t)nq * 32)) != ORB_OK ||
        (st = ws_upload(m, &db, mps->blocks, nq)) != ORB_OK || (st = ws_alloc(m, &dq, std::max(nq, 1))) != ORB_OK) return st;
    if (mps->proj_xr) { if ((st = ws_upload(m, &dxr, mps->proj_xr, nq)) != ORB_OK) return st; }
    if (nq > 0) { ORB_CUDA(orb_launch(true, k_prep_mappoints, dim3((nq + 255) / 256), dim3(256), 0, ws_stream(m), nq, dv, dx, dy, dxr, dl, dvc, th, F, dq, Batch{0, nullptr, nullptr})); m->launches++; }
    ProjOpts o; o.mode = 0; o.nn_ratio = nn_ratio;
    return projection_common(m, o, nq, dq, dd, db, nullptr, F, occupied, match, nmatches);
}

int orb_search_by_projection_lastframe(orb_matcher* m, const orb_frame_view* cur, const orb_lastframe_view* last, const orb_camera* cam,
                                       float th, int forward, int backward, int check_orientation, const uint8_t* occupied, int32_t* match,
                                       int32_t* nmatches) {
    ORB_NVTX("orb:search_by_projection_lastframe");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!cur || !last || !cam || !match || !nmatches || last->n < 0 || (last->n > 0 && (!last->valid || !last->xc || !last->yc || !last->zc ||
        !last->octave || !last->angle || !last->desc || !last->blocks)) || (check_orientation && cur->n > 0 && !cur->angle)) return ORB_ERR_BAD_ARG;
    DFrame F;
    if ((st = upload_frame(m, cur, &F, true)) != ORB_OK) return st;
    const int nq = last->n;
    const uint8_t *dv, *dd, *db; const float *dx, *dy, *dz, *da; const int* doct; Query* dq;
    if ((st = ws_upload(m, &dv, last->valid, nq)) != ORB_OK || (st = ws_upload(m, &dx, last->xc, nq)) != ORB_OK ||
        (st = ws_upload(m, &dy, last->yc, nq)) != ORB_OK || (st = ws_upload(m, &dz, last->zc, nq)) != ORB_OK ||
        (st = ws_upload(m, &doct, last->octave, nq)) != ORB_OK || (st = ws_upload(m, &da, last->angle, nq)) != ORB_OK ||
        (st = ws_upload(m, &dd, last->desc, (size_t)nq * 32)) != ORB_OK || (st = ws_upload(m, &db, last->blocks, nq)) != ORB_OK ||
        (st = ws_alloc(m, &dq, std::max(nq, 1))) != ORB_OK) return st;
    if (nq > 0) { ORB_CUDA(orb_launch(true, k_prep_lastframe, dim3((nq + 255) / 256), dim3(256), 0, ws_stream(m), nq, dv, dx, dy, dz, doct, *cam, th, forward, backward, F, dq, Batch{0, nullptr, nullptr}, (const uint8_t*)nullptr, (const uint8_t*)nullptr)); m->launches++; }
    ProjOpts o; o.mode = 1; o.check_ori = check_orientation;
    return projection_common(m, o, nq, dq, dd, db, da, F, occupied, match, nmatches);
}


// ------------------------------------------------------------------------------------------ batched projection searches
// (BASELINE configs[3]: many frame pairs per call).  Every array of the views holds `pairs` consecutive blocks of `stride`
// elements; each field goes up with ONE copy (cudaMemcpyDefault: host or device pointers alike), the kernels get a pair
// dimension (one resolve warp per pair), and the matches come back with one copy.
}  // extern "C"
namespace {
template <typename T>
int ws_upload_any(orb_matcher* m, const T** dp, const T* src, size_t count) {
    T* p = nullptr;
    int st = ws_alloc(m, &p, std::max<size_t>(count, 1));
    if (st != ORB_OK) return st;
    if (count) ORB_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyDefault, m->stream));
    *dp = p;
    return ORB_OK;
}

int upload_frame_batch(orb_matcher* m, const orb_frame_batch_view* f, DFrame* D, std::vector<int>& n_host) {
    if (!f || f->pairs < 1 || f->pairs > m->max_pairs || f->stride < 1 || f->stride > m->max_n || f->stride > 65535 || !f->n || !f->x || !f->y ||
        !f->octave || !f->desc) {
        orb_set_error("bad frame batch view (pairs <= max_pairs, stride <= min(max_n, 65535))");
        return ORB_ERR_BAD_ARG;
    }
    std::memset(D, 0, sizeof(*D));
    D->grid = f->grid; D->stride = f->stride;
    for (int i = 0; i < ORB_MAX_LEVELS; ++i) D->sf[i] = (f->scale_factors && i < f->n_levels) ? f->scale_factors[i] : 1.f;
    const size_t tot = (size_t)f->pairs * f->stride;
    int st;
    if ((st = ws_upload_any(m, &D->n_arr, f->n, f->pairs)) != ORB_OK || (st = ws_upload_any(m, &D->x, f->x, tot)) != ORB_OK ||
        (st = ws_upload_any(m, &D->y, f->y, tot)) != ORB_OK || (st = ws_upload_any(m, &D->octave, f->octave, tot)) != ORB_OK ||
        (st = ws_upload_any(m, &D->desc, f->desc, tot * 32)) != ORB_OK) return st;
    if (f->angle) { if ((st = ws_upload_any(m, &D->angle, f->angle, tot)) != ORB_OK) return st; }
    if (f->u_right) { if ((st = ws_upload_any(m, &D->u_right, f->u_right, tot)) != ORB_OK) return st; }
    // the counts are needed on the host too (bounds), wherever the caller keeps them
    n_host.resize(f->pairs);
    ORB_CUDA(cudaMemcpyAsync(n_host.data(), f->n, (size_t)f->pairs * 4, cudaMemcpyDefault, m->stream));
    int *cs, *ci, *ni;
    if ((st = ws_alloc(m, &cs, (size_t)f->pairs * (GRID_CELLS + 1))) != ORB_OK || (st = ws_alloc(m, &ci, tot)) != ORB_OK ||
        (st = ws_alloc(m, &ni, f->pairs)) != ORB_OK) return st;
    int p2 = 1;
    while (p2 < f->stride) p2 <<= 1;
    const size_t smem = (size_t)p2 * 4;
    if (smem > GRID_BUILD_DYN_DEFAULT) ORB_CUDA(cudaFuncSetAttribute(k_grid_build, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    k_grid_build<<<f->pairs, 1024, smem, ws_stream(m)>>>(D->x, D->y, 0, p2, f->grid, cs, ci, ni, f->stride, D->n_arr);
    m->launches++;
    D->cell_start = cs; D->cell_items = ci;
    return ORB_OK;
}

int projection_batch(orb_matcher* m, int mode, int pairs, int qstride, const int* d_nq, const Query* dq, const uint8_t* d_qdesc, const uint8_t* d_blocks,
                     const float* d_qangle, const DFrame& F, const uint8_t* occupied, float nn_ratio, int check_ori, int32_t* match,
                     int32_t* nmatches) {
    int st;
    const size_t qtot = (size_t)pairs * qstride, ftot = (size_t)pairs * F.stride;
    int *counts, *starts, *totals, *bases, *dmatch, *hidx, *hbin, *dn;
    uint8_t* occ; const uint8_t* d_occupied = nullptr;
    if ((st = ws_alloc(m, &counts, qtot)) != ORB_OK || (st = ws_alloc(m, &starts, (size_t)pairs * (qstride + 1))) != ORB_OK ||
        (st = ws_alloc(m, &totals, pairs)) != ORB_OK || (st = ws_alloc(m, &bases, pairs + 1)) != ORB_OK ||
        (st = ws_alloc(m, &dmatch, ftot)) != ORB_OK || (st = ws_alloc(m, &occ, ftot)) != ORB_OK ||
        (st = ws_alloc(m, &hidx, qtot)) != ORB_OK || (st = ws_alloc(m, &hbin, qtot)) != ORB_OK || (st = ws_alloc(m, &dn, pairs)) != ORB_OK) return st;
    if (occupied) { if ((st = ws_upload_any(m, &d_occupied, occupied, ftot)) != ORB_OK) return st; }
    Batch B{qstride, d_nq, nullptr};
    const dim3 grid((unsigned)((qstride * 32 + 255) / 256), (unsigned)pairs);
    k_area_candidates<false><<<grid, 256, 0, ws_stream(m)>>>(0, dq, d_qdesc, F, FILTER_STEREO_WINDOW, nullptr, counts, nullptr, nullptr, B);
    k_scan<<<pairs, 1024, 0, ws_stream(m)>>>(counts, 0, starts, B, totals);
    k_pair_bases<<<1, 32, 0, ws_stream(m)>>>(totals, pairs, bases);
    int total = 0;
    ORB_CUDA(cudaMemcpyAsync(&total, bases + pairs, 4, cudaMemcpyDeviceToHost, m->stream));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    if (total < 0) { orb_set_error("more than 2^31 candidates in one batch: split the call"); return ORB_ERR_CAPACITY; }
    if ((st = ensure_cands(m, (size_t)total)) != ORB_OK) return st;
    m->last_cands = total;
    B.cand_base = bases;
    k_area_candidates<true><<<grid, 256, 0, ws_stream(m)>>>(0, dq, d_qdesc, F, FILTER_STEREO_WINDOW, nullptr, nullptr, starts, m->d_cands, B);
    const int occ_shared = F.stride <= 40 * 1024;
    k_resolve_projection<<<pairs, 32, occ_shared ? (size_t)((F.stride + 15) & ~15) : 0, ws_stream(m)>>>(mode, 0, dq, starts, m->d_cands, d_blocks, d_qangle, F,
                                                  d_occupied, occ, nn_ratio, check_ori, M_TH_HIGH, dmatch, nullptr, nullptr, hidx, hbin, dn, occ_shared, B);
    m->launches += 5;
    ORB_CUDA(cudaGetLastError());
    ORB_CUDA(cudaMemcpyAsync(match, dmatch, ftot * 4, cudaMemcpyDefault, m->stream));
    ORB_CUDA(cudaMemcpyAsync(nmatches, dn, (size_t)pairs * 4, cudaMemcpyDefault, m->stream));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    return ORB_OK;
}
}  // namespace
extern "C" {

int orb_search_by_projection_mappoints_batch(orb_matcher* m, const orb_frame_batch_view* f, const orb_mappoint_batch_view* mps, float th, float nn_ratio,
                                             const uint8_t* occupied, int32_t* match, int32_t* nmatches) {
    ORB_NVTX("orb:search_by_projection_mappoints_batch");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!f || !mps || !match || !nmatches || mps->pairs != f->pairs || mps->stride < 1 || mps->stride > m->max_n || !mps->n || !mps->valid ||
        !mps->proj_x || !mps->proj_y || !mps->level || !mps->view_cos || !mps->desc || !mps->blocks) return ORB_ERR_BAD_ARG;
    DFrame F; std::vector<int> nf;
    if ((st = upload_frame_batch(m, f, &F, nf)) != ORB_OK) return st;
    const int pairs = f->pairs, qs = mps->stride;
    const size_t qtot = (size_t)pairs * qs;
    const uint8_t *dv, *dd, *db; const float *dx, *dy, *dxr = nullptr, *dvc; const int *dl, *dnq; Query* dq;
    if ((st = ws_upload_any(m, &dnq, mps->n, pairs)) != ORB_OK || (st = ws_upload_any(m, &dv, mps->valid, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &dx, mps->proj_x, qtot)) != ORB_OK || (st = ws_upload_any(m, &dy, mps->proj_y, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &dl, mps->level, qtot)) != ORB_OK || (st = ws_upload_any(m, &dvc, mps->view_cos, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &dd, mps->desc, qtot * 32)) != ORB_OK || (st = ws_upload_any(m, &db, mps->blocks, qtot)) != ORB_OK ||
        (st = ws_alloc(m, &dq, qtot)) != ORB_OK) return st;
    if (mps->proj_xr) { if ((st = ws_upload_any(m, &dxr, mps->proj_xr, qtot)) != ORB_OK) return st; }
    std::vector<int> nq(pairs);
    ORB_CUDA(cudaMemcpyAsync(nq.data(), mps->n, (size_t)pairs * 4, cudaMemcpyDefault, m->stream));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < pairs; ++p)
        if (nf[p] < 0 || nf[p] > f->stride || nq[p] < 0 || nq[p] > qs) { orb_set_error("batch view: n[p] outside [0, stride]"); return ORB_ERR_BAD_ARG; }
    k_prep_mappoints<<<dim3((unsigned)((qs + 255) / 256), (unsigned)pairs), 256, 0, ws_stream(m)>>>(0, dv, dx, dy, dxr, dl, dvc, th, F, dq, Batch{qs, dnq, nullptr});
    m->launches++;
    return projection_batch(m, 0, pairs, qs, dnq, dq, dd, db, nullptr, F, occupied, nn_ratio, 0, match, nmatches);
}

int orb_search_by_projection_lastframe_batch(orb_matcher* m, const orb_frame_batch_view* cur, const orb_lastframe_batch_view* last, const orb_camera* cam,
                                             float th, const uint8_t* forward, const uint8_t* backward, int check_orientation, const uint8_t* occupied,
                                             int32_t* match, int32_t* nmatches) {
    ORB_NVTX("orb:search_by_projection_lastframe_batch");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!cur || !last || !cam || !match || !nmatches || !forward || !backward || last->pairs != cur->pairs || last->stride < 1 || last->stride > m->max_n ||
        !last->n || !last->valid || !last->xc || !last->yc || !last->zc || !last->octave || !last->angle || !last->desc || !last->blocks ||
        (check_orientation && !cur->angle)) return ORB_ERR_BAD_ARG;
    DFrame F; std::vector<int> nf;
    if ((st = upload_frame_batch(m, cur, &F, nf)) != ORB_OK) return st;
    const int pairs = cur->pairs, qs = last->stride;
    const size_t qtot = (size_t)pairs * qs;
    const uint8_t *dv, *dd, *db, *dfw, *dbw; const float *dx, *dy, *dz, *da; const int *doct, *dnq; Query* dq;
    if ((st = ws_upload_any(m, &dnq, last->n, pairs)) != ORB_OK || (st = ws_upload_any(m, &dv, last->valid, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &dx, last->xc, qtot)) != ORB_OK || (st = ws_upload_any(m, &dy, last->yc, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &dz, last->zc, qtot)) != ORB_OK || (st = ws_upload_any(m, &doct, last->octave, qtot)) != ORB_OK ||
        (st = ws_upload_any(m, &da, last->angle, qtot)) != ORB_OK || (st = ws_upload_any(m, &dd, last->desc, qtot * 32)) != ORB_OK ||
        (st = ws_upload_any(m, &db, last->blocks, qtot)) != ORB_OK || (st = ws_upload_any(m, &dfw, forward, pairs)) != ORB_OK ||
        (st = ws_upload_any(m, &dbw, backward, pairs)) != ORB_OK || (st = ws_alloc(m, &dq, qtot)) != ORB_OK) return st;
    std::vector<int> nq(pairs);
    ORB_CUDA(cudaMemcpyAsync(nq.data(), last->n, (size_t)pairs * 4, cudaMemcpyDefault, m->stream));
    ORB_CUDA(cudaStreamSynchronize(m->stream));
    for (int p = 0; p < pairs; ++p)
        if (nf[p] < 0 || nf[p] > cur->stride || nq[p] < 0 || nq[p] > qs) { orb_set_error("batch view: n[p] outside [0, stride]"); return ORB_ERR_BAD_ARG; }
    k_prep_lastframe<<<dim3((unsigned)((qs + 255) / 256), (unsigned)pairs), 256, 0, ws_stream(m)>>>(0, dv, dx, dy, dz, doct, *cam, th, 0, 0, F, dq, Batch{qs, dnq, nullptr},
                                                                                              dfw, dbw);
    m->launches++;
    return projection_batch(m, 1, pairs, qs, dnq, dq, dd, db, da, F, occupied, 0.f, check_orientation, match, nmatches);
}

static int check_fv(const orb_featvec_view* v) {
    return v && v->n_nodes >= 0 && (v->n_nodes == 0 || (v->node_ids && v->start && v->items));
}

int orb_search_by_bow(orb_matcher* m, const orb_frame_view* kf, const orb_featvec_view* kf_fv, const uint8_t* kf_has_mappoint,
                      const orb_frame_view* f, const orb_featvec_view* f_fv, float nn_ratio, int check_orientation, int32_t* match_f,
                      int32_t* nmatches) {
    ORB_NVTX("orb:search_by_bow");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!kf || !f || !check_fv(kf_fv) || !check_fv(f_fv) || !kf_has_mappoint || !match_f || !nmatches ||
        (check_orientation && (!kf->angle || !f->angle))) return ORB_ERR_BAD_ARG;
    DFrame K, F;
    if ((st = upload_frame(m, kf, &K, false)) != ORB_OK || (st = upload_frame(m, f, &F, false)) != ORB_OK) return st;
    std::vector<std::pair<int, int>> nodes;
    merge_join(kf_fv, f_fv, nodes);
    std::vector<BowEntry> entries;
    std::vector<int> node_start;   // first entry of every node pair that has entries, then the total
    size_t total = 0;
    for (auto& ab : nodes) {
        const int f0 = f_fv->start[ab.second], f1 = f_fv->start[ab.second + 1];
        const size_t first = entries.size();
        for (int p = kf_fv->start[ab.first]; p < kf_fv->start[ab.first + 1]; ++p) {
            const int qi = (int)kf_fv->items[p];
            if (qi < 0 || qi >= kf->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
            if (!kf_has_mappoint[qi] || f1 == f0) continue;
            BowEntry e; e.q_idx = qi; e.f0 = f0; e.f1 = f1; e.c_start = (int)total;
            entries.push_back(e);
            total += (size_t)(f1 - f0);
        }
        if (entries.size() > first) node_start.push_back((int)first);
    }
    if ((st = ensure_cands(m, total)) != ORB_OK) return st;
    const int ne = (int)entries.size(), nn = (int)node_start.size();
    node_start.push_back(ne);
    const int n_items2 = f_fv->n_nodes ? f_fv->start[f_fv->n_nodes] : 0;
    for (int i = 0; i < n_items2; ++i) if (f_fv->items[i] >= (uint32_t)f->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
    const BowEntry* de; const uint32_t* ditems; const int* dns; int *dmatch, *hidx, *hbin;
    if ((st = ws_upload(m, &de, entries.data(), ne)) != ORB_OK || (st = ws_upload(m, &ditems, f_fv->items, n_items2)) != ORB_OK ||
        (st = ws_upload(m, &dns, node_start.data(), nn + 1)) != ORB_OK ||
        (st = ws_alloc(m, &dmatch, std::max(f->n, 1))) != ORB_OK || (st = ws_alloc(m, &hidx, std::max(ne, 1))) != ORB_OK ||
        (st = ws_alloc(m, &hbin, std::max(ne, 1))) != ORB_OK) return st;
    ORB_CUDA(cudaMemsetAsync(dmatch, 0xff, (size_t)std::max(f->n, 1) * 4, m->stream));
    ORB_CUDA(cudaMemsetAsync(hidx, 0xff, (size_t)std::max(ne, 1) * 4, m->stream));
    if (ne > 0) {
        k_bow_distances<<<(ne * 32 + 255) / 256, 256, 0, ws_stream(m)>>>(ne, de, K.desc, F.desc, ditems, m->d_cands);
        k_resolve_bow<<<(nn + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, ws_stream(m)>>>(nn, dns, de, m->d_cands, K.angle, F.angle, nn_ratio,
                                                                                          check_orientation, dmatch, hidx, hbin);
        m->launches += 2;
    }
    k_bow_finish<<<1, 32, 0, ws_stream(m)>>>(ne, hidx, hbin, check_orientation, dmatch, m->d_scal);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    if (f->n > 0) ORB_CUDA(ws_download(m, match_f, dmatch, (size_t)f->n * 4));
    ORB_CUDA(ws_download(m, nmatches, m->d_scal, 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_search_for_triangulation(orb_matcher* m, const orb_frame_view* kf1, const orb_featvec_view* fv1, const uint8_t* kf1_has_mappoint,
                                 const orb_frame_view* kf2, const orb_featvec_view* fv2, const uint8_t* kf2_has_mappoint, const float* f12,
                                 float ex, float ey, const float* level_sigma2, int only_stereo, int check_orientation, int32_t* match12,
                                 int32_t* nmatches) {
    ORB_NVTX("orb:search_for_triangulation");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!kf1 || !kf2 || !check_fv(fv1) || !check_fv(fv2) || !kf1_has_mappoint || !kf2_has_mappoint || !f12 || !level_sigma2 || !match12 ||
        !nmatches || (check_orientation && (!kf1->angle || !kf2->angle))) return ORB_ERR_BAD_ARG;
    DFrame K1, K2;
    if ((st = upload_frame(m, kf1, &K1, false)) != ORB_OK || (st = upload_frame(m, kf2, &K2, false)) != ORB_OK) return st;
    std::vector<std::pair<int, int>> nodes;
    merge_join(fv1, fv2, nodes);
    std::vector<BowEntry> entries;
    for (auto& ab : nodes) {
        const int f0 = fv2->start[ab.second], f1 = fv2->start[ab.second + 1];
        if (f1 - f0 > 65535) { orb_set_error("FeatureVector node too large"); return ORB_ERR_CAPACITY; }
        for (int p = fv1->start[ab.first]; p < fv1->start[ab.first + 1]; ++p) {
            const int qi = (int)fv1->items[p];
            if (qi < 0 || qi >= kf1->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
            if (kf1_has_mappoint[qi]) continue;
            if (only_stereo && !(kf1->u_right && kf1->u_right[qi] >= 0)) continue;
            BowEntry e; e.q_idx = qi; e.f0 = f0; e.f1 = f1; e.c_start = 0;
            entries.push_back(e);
        }
    }
    const int ne = (int)entries.size();
    const int n_items2 = fv2->n_nodes ? fv2->start[fv2->n_nodes] : 0;
    for (int i = 0; i < n_items2; ++i) if (fv2->items[i] >= (uint32_t)kf2->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
    const BowEntry* de; const uint32_t* ditems; const uint8_t* dmp2; int *dmatch, *hidx, *hbin;
    if ((st = ws_upload(m, &de, entries.data(), ne)) != ORB_OK || (st = ws_upload(m, &ditems, fv2->items, n_items2)) != ORB_OK ||
        (st = ws_upload(m, &dmp2, kf2_has_mappoint, kf2->n)) != ORB_OK || (st = ws_alloc(m, &dmatch, std::max(kf1->n, 1))) != ORB_OK ||
        (st = ws_alloc(m, &hidx, std::max(kf1->n, 1))) != ORB_OK || (st = ws_alloc(m, &hbin, std::max(kf1->n, 1))) != ORB_OK) return st;
    ORB_CUDA(cudaMemsetAsync(dmatch, 0xff, (size_t)std::max(kf1->n, 1) * 4, m->stream));
    TriParams T;
    for (int i = 0; i < 9; ++i) T.f12[i] = f12[i];
    T.ex = ex; T.ey = ey; T.only_stereo = only_stereo;
    for (int i = 0; i < ORB_MAX_LEVELS; ++i) T.sigma2[i] = i < kf2->n_levels ? level_sigma2[i] : 1.f;
    if (ne > 0) { k_triangulation<<<(ne * 32 + 255) / 256, 256, 0, ws_stream(m)>>>(ne, de, K1, K2, ditems, dmp2, T, dmatch); m->launches++; }
    k_tri_rot_cull<<<1, 32, 0, ws_stream(m)>>>(kf1->n, K1.angle, K2.angle, check_orientation, dmatch, hidx, hbin, m->d_scal);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    if (kf1->n > 0) ORB_CUDA(ws_download(m, match12, dmatch, (size_t)kf1->n * 4));
    ORB_CUDA(ws_download(m, nmatches, m->d_scal, 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}


// ---- searches over caller-projected points (reference :311-427, :903-1207, :1212-1455, :1622-1759) ----
static int check_points(const orb_projected_view* p, int need_angle) {
    return p && p->n >= 0 && (p->n == 0 || (p->valid && p->u && p->v && p->radius && p->min_level && p->max_level && p->desc &&
                                            (!need_angle || p->angle)));
}

static int upload_points(orb_matcher* m, const orb_projected_view* p, const Query** dq, const uint8_t** dd, const float** da) {
    std::vector<Query> q((size_t)std::max(p->n, 1));
    for (int i = 0; i < p->n; ++i) {
        Query& Q = q[i];
        Q.valid = p->valid[i]; Q.x = p->u[i]; Q.y = p->v[i]; Q.r = p->radius[i]; Q.xr = p->ur ? p->ur[i] : 0.f;
        Q.min_level = p->min_level[i]; Q.max_level = p->max_level[i]; Q.pad = 0;
    }
    int st;
    *da = nullptr;
    if ((st = ws_upload(m, dq, q.data(), (size_t)p->n)) != ORB_OK || (st = ws_upload(m, dd, p->desc, (size_t)p->n * 32)) != ORB_OK) return st;
    if (p->angle) { if ((st = ws_upload(m, da, p->angle, (size_t)p->n)) != ORB_OK) return st; }
    return ORB_OK;
}

int orb_search_by_projection_claiming(orb_matcher* m, const orb_frame_view* f, const orb_projected_view* pts, const uint8_t* occupied,
                                      int max_dist, int check_orientation, int32_t* match, int32_t* point_best_idx, int32_t* nmatches) {
    ORB_NVTX("orb:search_by_projection_claiming");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!f || !check_points(pts, check_orientation) || !match || !nmatches || max_dist < 0 || max_dist > 255 ||
        (check_orientation && f->n > 0 && !f->angle)) return ORB_ERR_BAD_ARG;
    DFrame F;
    if ((st = upload_frame(m, f, &F, true)) != ORB_OK) return st;
    const Query* dq; const uint8_t* dd; const float* da;
    if ((st = upload_points(m, pts, &dq, &dd, &da)) != ORB_OK) return st;
    ProjOpts o; o.mode = 2; o.filter = FILTER_NONE; o.check_ori = check_orientation; o.max_dist = max_dist; o.q_best_idx = point_best_idx;
    return projection_common(m, o, pts->n, dq, dd, nullptr, da, F, occupied, match, nmatches);
}

int orb_search_projected_best(orb_matcher* m, const orb_frame_view* f, const orb_projected_view* pts, const float* inv_level_sigma2,
                              int max_dist, int32_t* best_idx, int32_t* best_dist, int32_t* nfound) {
    ORB_NVTX("orb:search_projected_best");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!f || !check_points(pts, 0) || !best_idx || !nfound || max_dist < 0 || max_dist > 255) return ORB_ERR_BAD_ARG;
    DFrame F;
    if ((st = upload_frame(m, f, &F, true)) != ORB_OK) return st;
    const Query* dq; const uint8_t* dd; const float* da;
    if ((st = upload_points(m, pts, &dq, &dd, &da)) != ORB_OK) return st;
    ProjOpts o; o.mode = -1; o.filter = FILTER_NONE; o.max_dist = max_dist; o.q_best_idx = best_idx; o.q_best_dist = best_dist;
    if (inv_level_sigma2) {
        float s[ORB_MAX_LEVELS];
        for (int i = 0; i < ORB_MAX_LEVELS; ++i) s[i] = i < f->n_levels ? inv_level_sigma2[i] : 1.f;
        if ((st = ws_upload(m, &o.d_inv_sigma2, s, (size_t)ORB_MAX_LEVELS)) != ORB_OK) return st;
        o.filter = FILTER_FUSE_CHI2;
    }
    return projection_common(m, o, pts->n, dq, dd, nullptr, nullptr, F, nullptr, nullptr, nfound);
}

int orb_search_by_sim3(orb_matcher* m, const orb_frame_view* kf1, const orb_projected_view* pts1_in_kf2, const orb_frame_view* kf2,
                       const orb_projected_view* pts2_in_kf1, int32_t* match12, int32_t* nfound) {
    ORB_NVTX("orb:search_by_sim3");
    if (!m || !kf1 || !kf2 || !check_points(pts1_in_kf2, 0) || !check_points(pts2_in_kf1, 0) || !match12 || !nfound ||
        pts1_in_kf2->n != kf1->n || pts2_in_kf1->n != kf2->n) return ORB_ERR_BAD_ARG;
    // vnMatch1 / vnMatch2 (reference :1290-1340, :1343-1420): two independent best-candidate searches
    std::vector<int32_t> m1((size_t)std::max(kf1->n, 1)), m2((size_t)std::max(kf2->n, 1));
    int32_t n1 = 0, n2 = 0;
    int st = orb_search_projected_best(m, kf2, pts1_in_kf2, nullptr, M_TH_HIGH, m1.data(), nullptr, &n1);
    if (st != ORB_OK) return st;
    if ((st = orb_search_projected_best(m, kf1, pts2_in_kf1, nullptr, M_TH_HIGH, m2.data(), nullptr, &n2)) != ORB_OK) return st;
    // mutual agreement (:1423-1439)
    if ((st = begin_call(m)) != ORB_OK) return st;
    const int32_t *d1, *d2; int* d12;
    if ((st = ws_upload(m, &d1, m1.data(), (size_t)kf1->n)) != ORB_OK || (st = ws_upload(m, &d2, m2.data(), (size_t)kf2->n)) != ORB_OK ||
        (st = ws_alloc(m, &d12, std::max(kf1->n, 1))) != ORB_OK) return st;
    ORB_CUDA(cudaMemsetAsync(m->d_scal, 0, 4, m->stream));
    if (kf1->n > 0) {
        k_mutual_check<<<(kf1->n + 255) / 256, 256, 0, ws_stream(m)>>>(kf1->n, kf2->n, d1, d2, d12, m->d_scal);
        m->launches++;
        ORB_CUDA(cudaGetLastError());
        ORB_CUDA(ws_download(m, match12, d12, (size_t)kf1->n * 4));
    }
    ORB_CUDA(ws_download(m, nfound, m->d_scal, 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_search_for_initialization(orb_matcher* m, const orb_frame_view* f1, const orb_frame_view* f2, const float* prev_matched_x,
                                  const float* prev_matched_y, int window_size, float nn_ratio, int check_orientation, int32_t* match12,
                                  int32_t* nmatches) {
    ORB_NVTX("orb:search_for_initialization");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!f1 || !f2 || !match12 || !nmatches || (f1->n > 0 && (!prev_matched_x || !prev_matched_y || !f1->octave)) ||
        (check_orientation && ((f1->n > 0 && !f1->angle) || (f2->n > 0 && !f2->angle)))) return ORB_ERR_BAD_ARG;
    DFrame F2;
    if ((st = upload_frame(m, f2, &F2, true)) != ORB_OK) return st;
    const int nq = f1->n;
    std::vector<Query> q((size_t)std::max(nq, 1));
    for (int i = 0; i < nq; ++i) {
        Query& Q = q[i];
        Q.valid = f1->octave[i] > 0 ? 0 : 1;          // level1 > 0 -> continue (:449-451)
        Q.x = prev_matched_x[i]; Q.y = prev_matched_y[i]; Q.r = (float)window_size; Q.xr = 0.f;
        Q.min_level = 0; Q.max_level = 0; Q.pad = 0;  // GetFeaturesInArea(..., level1, level1) with level1 == 0
    }
    const Query* dq; const uint8_t* dd; const float* da1 = nullptr;
    int *counts, *starts, *d12, *d21, *dmd, *hidx, *hbin;
    if ((st = ws_upload(m, &dq, q.data(), (size_t)nq)) != ORB_OK || (st = ws_upload(m, &dd, f1->desc, (size_t)nq * 32)) != ORB_OK ||
        (st = ws_alloc(m, &counts, std::max(nq, 1))) != ORB_OK || (st = ws_alloc(m, &starts, nq + 1)) != ORB_OK ||
        (st = ws_alloc(m, &d12, std::max(nq, 1))) != ORB_OK || (st = ws_alloc(m, &d21, std::max(f2->n, 1))) != ORB_OK ||
        (st = ws_alloc(m, &dmd, std::max(f2->n, 1))) != ORB_OK || (st = ws_alloc(m, &hidx, std::max(nq, 1))) != ORB_OK ||
        (st = ws_alloc(m, &hbin, std::max(nq, 1))) != ORB_OK) return st;
    if (f1->angle) { if ((st = ws_upload(m, &da1, f1->angle, (size_t)nq)) != ORB_OK) return st; }
    if (nq > 0) {
        const int blocks = (nq * 32 + 255) / 256;
        k_area_candidates<false><<<blocks, 256, 0, ws_stream(m)>>>(nq, dq, dd, F2, FILTER_NONE, nullptr, counts, nullptr, nullptr, Batch{0, nullptr, nullptr});
        k_scan<<<1, 1024, 0, ws_stream(m)>>>(counts, nq, starts, Batch{0, nullptr, nullptr}, nullptr);
        int total = 0;
        ORB_CUDA(cudaMemcpyAsync(&total, starts + nq, 4, cudaMemcpyDeviceToHost, m->stream));
        ORB_CUDA(cudaStreamSynchronize(m->stream));
        if ((st = ensure_cands(m, (size_t)total)) != ORB_OK) return st;
        k_area_candidates<true><<<blocks, 256, 0, ws_stream(m)>>>(nq, dq, dd, F2, FILTER_NONE, nullptr, nullptr, starts, m->d_cands, Batch{0, nullptr, nullptr});
        m->launches += 3;
    } else {
        ORB_CUDA(cudaMemsetAsync(starts, 0, 4, m->stream));
    }
    k_resolve_initialization<<<1, 32, 0, ws_stream(m)>>>(nq, starts, m->d_cands, da1, F2.angle, f2->n, nn_ratio, check_orientation, d12, d21, dmd,
                                                      hidx, hbin, m->d_scal);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    if (nq > 0) ORB_CUDA(ws_download(m, match12, d12, (size_t)nq * 4));
    ORB_CUDA(ws_download(m, nmatches, m->d_scal, 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_search_by_bow_keyframes(orb_matcher* m, const orb_frame_view* kf1, const orb_featvec_view* fv1, const uint8_t* kf1_has_mappoint,
                                const orb_frame_view* kf2, const orb_featvec_view* fv2, const uint8_t* kf2_has_mappoint, float nn_ratio,
                                int check_orientation, int32_t* match12, int32_t* nmatches) {
    ORB_NVTX("orb:search_by_bow_keyframes");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!kf1 || !kf2 || !check_fv(fv1) || !check_fv(fv2) || !kf1_has_mappoint || !kf2_has_mappoint || !match12 || !nmatches ||
        (check_orientation && (!kf1->angle || !kf2->angle))) return ORB_ERR_BAD_ARG;
    DFrame K1, K2;
    if ((st = upload_frame(m, kf1, &K1, false)) != ORB_OK || (st = upload_frame(m, kf2, &K2, false)) != ORB_OK) return st;
    std::vector<std::pair<int, int>> nodes;
    merge_join(fv1, fv2, nodes);
    std::vector<BowEntry> entries;
    std::vector<int> node_start;   // first entry of every node pair that has entries, then the total
    size_t total = 0;
    for (auto& ab : nodes) {
        const int f0 = fv2->start[ab.second], f1 = fv2->start[ab.second + 1];
        const size_t first = entries.size();
        for (int p = fv1->start[ab.first]; p < fv1->start[ab.first + 1]; ++p) {
            const int qi = (int)fv1->items[p];
            if (qi < 0 || qi >= kf1->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
            if (!kf1_has_mappoint[qi] || f1 == f0) continue;
            BowEntry e; e.q_idx = qi; e.f0 = f0; e.f1 = f1; e.c_start = (int)total;
            entries.push_back(e);
            total += (size_t)(f1 - f0);
        }
        if (entries.size() > first) node_start.push_back((int)first);
    }
    if ((st = ensure_cands(m, total)) != ORB_OK) return st;
    const int ne = (int)entries.size(), nn = (int)node_start.size();
    node_start.push_back(ne);
    const int n_items2 = fv2->n_nodes ? fv2->start[fv2->n_nodes] : 0;
    for (int i = 0; i < n_items2; ++i) if (fv2->items[i] >= (uint32_t)kf2->n) { orb_set_error("FeatureVector index out of range"); return ORB_ERR_BAD_ARG; }
    const BowEntry* de; const uint32_t* ditems; const uint8_t* dhas2; const int* dns; uint8_t* dclosed; int *dmatch, *hidx, *hbin;
    if ((st = ws_upload(m, &de, entries.data(), ne)) != ORB_OK || (st = ws_upload(m, &ditems, fv2->items, n_items2)) != ORB_OK ||
        (st = ws_upload(m, &dns, node_start.data(), nn + 1)) != ORB_OK ||
        (st = ws_upload(m, &dhas2, kf2_has_mappoint, kf2->n)) != ORB_OK || (st = ws_alloc(m, &dclosed, std::max(kf2->n, 1))) != ORB_OK ||
        (st = ws_alloc(m, &dmatch, std::max(kf1->n, 1))) != ORB_OK || (st = ws_alloc(m, &hidx, std::max(ne, 1))) != ORB_OK ||
        (st = ws_alloc(m, &hbin, std::max(ne, 1))) != ORB_OK) return st;
    ORB_CUDA(cudaMemsetAsync(dmatch, 0xff, (size_t)std::max(kf1->n, 1) * 4, m->stream));
    ORB_CUDA(cudaMemsetAsync(hidx, 0xff, (size_t)std::max(ne, 1) * 4, m->stream));
    if (ne > 0) {
        k_bow_closed_init<<<(kf2->n + 255) / 256, 256, 0, ws_stream(m)>>>(kf2->n, dhas2, dclosed);
        k_bow_distances<<<(ne * 32 + 255) / 256, 256, 0, ws_stream(m)>>>(ne, de, K1.desc, K2.desc, ditems, m->d_cands);
        k_resolve_bow_kf<<<(nn + BOW_WARPS - 1) / BOW_WARPS, BOW_WARPS * 32, 0, ws_stream(m)>>>(nn, dns, de, m->d_cands, K1.angle, K2.angle, dclosed, nn_ratio,
                                                                                             check_orientation, dmatch, hidx, hbin);
        m->launches += 3;
    }
    k_bow_finish<<<1, 32, 0, ws_stream(m)>>>(ne, hidx, hbin, check_orientation, dmatch, m->d_scal);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    if (kf1->n > 0) ORB_CUDA(ws_download(m, match12, dmatch, (size_t)kf1->n * 4));
    ORB_CUDA(ws_download(m, nmatches, m->d_scal, 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}
}  // extern "C"

// ------------------------------------------------------------------------------------------
// Feeding SearchByProjection: Frame::isInFrustum over the local map (reference src/Frame.cc:303-366,
// called per MapPoint from Tracking::SearchLocalPoints, src/Tracking.cc:1012-1036) and
// MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:250-326).
namespace {

// One thread per MapPoint.  Float sequence of the reference (cv::Mat algebra as OpenCV 4.13 evaluates
// it: R*P+t in float left to right, norm and dot in double; scalar expressions as GCC contracts them
// for the reference's -O3 -march=native build: fx*PcX*invz+cx -> fma(fx*PcX, invz, cx), u-mbf*invz ->
// fma(-mbf, invz, u)).  PredictScale (a logf) stays with the caller; dist is returned for it.
__global__ void k_is_in_frustum(int n, const float* __restrict__ xw, const float* __restrict__ nrm, const float* __restrict__ mind,
                                const float* __restrict__ maxd, float r00, float r01, float r02, float r10, float r11, float r12, float r20,
                                float r21, float r22, float t0, float t1, float t2, float o0, float o1, float o2, orb_camera cam,
                                float min_x, float min_y, float max_x, float max_y, float cos_limit, uint8_t* __restrict__ in_view,
                                float* __restrict__ u_out, float* __restrict__ v_out, float* __restrict__ ur_out, float* __restrict__ cos_out,
                                float* __restrict__ dist_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float X = xw[3 * i], Y = xw[3 * i + 1], Z = xw[3 * i + 2];
    uint8_t ok = 0;
    float u = 0.f, v = 0.f, ur = 0.f, vc = 0.f, dist = 0.f;
    const float pcx = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r00, X), __fmul_rn(r01, Y)), __fmul_rn(r02, Z)), t0);
    const float pcy = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r10, X), __fmul_rn(r11, Y)), __fmul_rn(r12, Z)), t1);
    const float pcz = __fadd_rn(__fadd_rn(__fadd_rn(__fmul_rn(r20, X), __fmul_rn(r21, Y)), __fmul_rn(r22, Z)), t2);
    if (!(pcz < 0.0f)) {
        const float invz = __fdiv_rn(1.0f, pcz);
        u = __fmaf_rn(__fmul_rn(cam.fx, pcx), invz, cam.cx);
        v = __fmaf_rn(__fmul_rn(cam.fy, pcy), invz, cam.cy);
        if (!(u < min_x || u > max_x) && !(v < min_y || v > max_y)) {
            const float px = __fsub_rn(X, o0), py = __fsub_rn(Y, o1), pz = __fsub_rn(Z, o2);
            dist = (float)sqrt(__dadd_rn(__dadd_rn(__dmul_rn((double)px, (double)px), __dmul_rn((double)py, (double)py)), __dmul_rn((double)pz, (double)pz)));
            if (!(dist < __fmul_rn(0.8f, mind[i]) || dist > __fmul_rn(1.2f, maxd[i]))) {
                const double dot = __dadd_rn(__dadd_rn(__dmul_rn((double)px, (double)nrm[3 * i]), __dmul_rn((double)py, (double)nrm[3 * i + 1])),
                                             __dmul_rn((double)pz, (double)nrm[3 * i + 2]));
                vc = (float)__ddiv_rn(dot, (double)dist);
                if (!(vc < cos_limit)) {
                    ok = 1;
                    ur = __fmaf_rn(-cam.mbf, invz, u);
                }
            }
        }
    }
    in_view[i] = ok; u_out[i] = u; v_out[i] = v; ur_out[i] = ur; cos_out[i] = vc; dist_out[i] = dist;
}

// One CTA per MapPoint: N x N Hamming distances of its observations' descriptors, the median of every row
// (element int(0.5 * (N - 1)) of the sorted row, found from a 257-bin histogram), the row with the
// smallest median (first wins).  One warp per row at a time.
#define DD_THREADS 128
__global__ void __launch_bounds__(DD_THREADS) k_distinctive(const int* __restrict__ start, const uint8_t* __restrict__ desc, int* __restrict__ best_idx) {
    __shared__ int hist[DD_THREADS / 32][264];
    __shared__ unsigned best[DD_THREADS / 32];
    const int p = blockIdx.x, s = start[p], N = start[p + 1] - s;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    unsigned my_best = 0xffffffffu;                  // (median << 16) | row
    const int kth = (int)(0.5 * (N - 1));
    for (int i = warp; i < N; i += DD_THREADS / 32) {
        for (int b = lane; b < 264; b += 32) hist[warp][b] = 0;
        __syncwarp();
        const uint4* pi = reinterpret_cast<const uint4*>(desc + (size_t)(s + i) * 32);
        const uint4 a0 = __ldg(pi), a1 = __ldg(pi + 1);
        for (int j = lane; j < N; j += 32) {
            const uint4* pj = reinterpret_cast<const uint4*>(desc + (size_t)(s + j) * 32);
            atomicAdd(&hist[warp][orb_hamming256(a0, a1, __ldg(pj), __ldg(pj + 1))], 1);
        }
        __syncwarp();
        // lane l owns bins [8l, 8l + 8) (+ bin 256 on lane 31 via the 9th slot)
        int c[9], tot = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) { c[b] = hist[warp][lane * 8 + b]; tot += c[b]; }
        c[8] = lane == 31 ? hist[warp][256] : 0; tot += c[8];
        int incl = tot;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
        int before = incl - tot, med = -1;
#pragma unroll
        for (int b = 0; b < 9; ++b) { if (med < 0 && kth < before + c[b]) med = (b == 8 ? 256 : lane * 8 + b); before += c[b]; }
        const unsigned found = __ballot_sync(0xffffffffu, med >= 0);
        const int m = __shfl_sync(0xffffffffu, med, __ffs(found) - 1);
        my_best = min(my_best, ((unsigned)m << 16) | (unsigned)i);
        __syncwarp();
    }
    if (lane == 0) best[warp] = my_best;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned b = best[0];
        for (int w = 1; w < DD_THREADS / 32; ++w) b = min(b, best[w]);
        best_idx[p] = N > 0 ? (int)(b & 0xffffu) : -1;
    }
}

}  // namespace

extern "C" {

int orb_is_in_frustum(orb_matcher* m, const float* rcw, const float* tcw, const float* ow, const orb_camera* cam, const orb_grid_params* bounds,
                      float viewing_cos_limit, int n, const float* xw, const float* normal, const float* min_distance, const float* max_distance,
                      uint8_t* in_view, float* u, float* v, float* ur, float* view_cos, float* dist) {
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (!rcw || !tcw || !ow || !cam || !bounds || n < 0 || (n > 0 && (!xw || !normal || !min_distance || !max_distance || !in_view || !u || !v || !ur || !view_cos || !dist)))
        return ORB_ERR_BAD_ARG;
    if (n == 0) return ORB_OK;
    const float *dx, *dn, *dmin, *dmax; float *du, *dv, *dur, *dc, *dd; uint8_t* dok;
    if ((st = ws_upload(m, &dx, xw, (size_t)n * 3)) != ORB_OK || (st = ws_upload(m, &dn, normal, (size_t)n * 3)) != ORB_OK ||
        (st = ws_upload(m, &dmin, min_distance, n)) != ORB_OK || (st = ws_upload(m, &dmax, max_distance, n)) != ORB_OK ||
        (st = ws_alloc(m, &du, n)) != ORB_OK || (st = ws_alloc(m, &dv, n)) != ORB_OK || (st = ws_alloc(m, &dur, n)) != ORB_OK ||
        (st = ws_alloc(m, &dc, n)) != ORB_OK || (st = ws_alloc(m, &dd, n)) != ORB_OK || (st = ws_alloc(m, &dok, n)) != ORB_OK) return st;
    k_is_in_frustum<<<(n + 255) / 256, 256, 0, ws_stream(m)>>>(n, dx, dn, dmin, dmax, rcw[0], rcw[1], rcw[2], rcw[3], rcw[4], rcw[5], rcw[6], rcw[7], rcw[8],
                                                            tcw[0], tcw[1], tcw[2], ow[0], ow[1], ow[2], *cam, bounds->min_x, bounds->min_y, bounds->max_x,
                                                            bounds->max_y, viewing_cos_limit, dok, du, dv, dur, dc, dd);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    ORB_CUDA(ws_download(m, in_view, dok, n));
    ORB_CUDA(ws_download(m, u, du, (size_t)n * 4));
    ORB_CUDA(ws_download(m, v, dv, (size_t)n * 4));
    ORB_CUDA(ws_download(m, ur, dur, (size_t)n * 4));
    ORB_CUDA(ws_download(m, view_cos, dc, (size_t)n * 4));
    ORB_CUDA(ws_download(m, dist, dd, (size_t)n * 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

int orb_distinctive_descriptors(orb_matcher* m, int n_points, const int32_t* start, const uint8_t* desc, int32_t* best_idx) {
    ORB_NVTX("orb:distinctive_descriptors");
    int st = begin_call(m);
    if (st != ORB_OK) return st;
    if (n_points < 0 || (n_points > 0 && (!start || !desc || !best_idx))) return ORB_ERR_BAD_ARG;
    if (n_points == 0) return ORB_OK;
    for (int p = 0; p < n_points; ++p)
        if (start[p + 1] < start[p] || start[p + 1] - start[p] > 65535) { orb_set_error("bad observation ranges"); return ORB_ERR_BAD_ARG; }
    const int total = start[n_points];
    const int* ds; const uint8_t* dd; int* db;
    if ((st = ws_upload(m, &ds, start, (size_t)n_points + 1)) != ORB_OK || (st = ws_upload(m, &dd, desc, (size_t)total * 32)) != ORB_OK ||
        (st = ws_alloc(m, &db, n_points)) != ORB_OK) return st;
    k_distinctive<<<n_points, DD_THREADS, 0, ws_stream(m)>>>(ds, dd, db);
    m->launches++;
    ORB_CUDA(cudaGetLastError());
    ORB_CUDA(ws_download(m, best_idx, db, (size_t)n_points * 4));
    ORB_CUDA(ws_finish(m));
    return ORB_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------
// orb_assign_features_to_grid_static: Frame::AssignFeaturesToGrid for callers that may not allocate (CUDA graph capture around
// them, serving loops, harnesses that forbid cudaMalloc).  No handle and no workspace: the device copies of the keypoint
// coordinates and of the three results that k_grid_build works on (the bump-allocated workspace of the entry point above) live in
// one file-scope __device__ array, 108 KB, held by every device that loads the library until the process exits.
// In host code the array's NAME is only a stand-in: its device address comes from cudaGetSymbolAddress, once per device.
#define GRID_STATIC_MAX_N ORB_GRID_STATIC_MAX_N
#define GRID_STATIC_MAX_DEVICES 64
namespace {
struct GridStaticScratch {   // every member is a multiple of 256 bytes
    float x[GRID_STATIC_MAX_N];
    float y[GRID_STATIC_MAX_N];
    int cell_items[GRID_STATIC_MAX_N];
    int cell_start[GRID_CELLS + 64];   // GRID_CELLS + 1 in use
    int n_items[64];                   // [0] in use
};
}  // namespace
static __device__ __align__(256) unsigned char g_grid_static[sizeof(GridStaticScratch)];

static GridStaticScratch* g_grid_static_ptr[GRID_STATIC_MAX_DEVICES];   // per device, filled at first use
static std::mutex g_grid_static_mu;

// Device address of the scratch array on the CURRENT device (looked up and cached at the first call per device).
static int grid_static_scratch(GridStaticScratch** out) {
    int dev = -1;
    ORB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= GRID_STATIC_MAX_DEVICES) { orb_set_error("orb_assign_features_to_grid_static: device ordinal out of range"); return ORB_ERR_BAD_ARG; }
    std::lock_guard<std::mutex> lk(g_grid_static_mu);
    if (!g_grid_static_ptr[dev]) {
        void* p = nullptr;
        ORB_CUDA(cudaGetSymbolAddress(&p, g_grid_static));
        if (!p) { orb_set_error("cudaGetSymbolAddress(g_grid_static) returned a null address"); return ORB_ERR_CUDA; }
        // 8192 keys (32 KB of dynamic shared memory) beside the kernel's static 25 KB pass the 48 KB a launch gets by default
        ORB_CUDA(cudaFuncSetAttribute(k_grid_build, cudaFuncAttributeMaxDynamicSharedMemorySize, GRID_STATIC_MAX_N * 4));
        g_grid_static_ptr[dev] = static_cast<GridStaticScratch*>(p);
    }
    *out = g_grid_static_ptr[dev];
    return ORB_OK;
}

extern "C" int orb_assign_features_to_grid_static(const float* kp_x, const float* kp_y, int n, const orb_grid_params* g, int32_t* cell_start,
                                                  int32_t* cell_items, int32_t* n_items, void* stream) {
    ORB_NVTX("orb:assign_features_to_grid_static");
    if (!g || !cell_start || !cell_items || !n_items || n < 0 || (n > 0 && (!kp_x || !kp_y))) {
        orb_set_error("orb_assign_features_to_grid_static: null argument or negative n");
        return ORB_ERR_BAD_ARG;
    }
    if (n > GRID_STATIC_MAX_N) {
        orb_set_error("orb_assign_features_to_grid_static: n = " + std::to_string(n) + " keypoints exceeds the static scratch's maximum of " +
                      std::to_string(GRID_STATIC_MAX_N) + " (use orb_assign_features_to_grid)");
        return ORB_ERR_CAPACITY;
    }
    GridStaticScratch* S = nullptr;
    int st = grid_static_scratch(&S);
    if (st != ORB_OK) return st;
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // The array keeps the last call's contents.  k_grid_build overwrites everything it later reads (its counters live in shared
    // memory and are zeroed there; cell_start is written in full, cell_items up to n_items); only the count is cleared here, so
    // that a launch that did not run cannot hand back the previous call's.
    ORB_CUDA(cudaMemsetAsync(S->n_items, 0, sizeof(int), s));
    if (n > 0) {
        ORB_CUDA(cudaMemcpyAsync(S->x, kp_x, (size_t)n * 4, cudaMemcpyDefault, s));
        ORB_CUDA(cudaMemcpyAsync(S->y, kp_y, (size_t)n * 4, cudaMemcpyDefault, s));
    }
    int p2 = 1;
    while (p2 < std::max(n, 1)) p2 <<= 1;
    k_grid_build<<<1, 1024, (size_t)p2 * 4, s>>>(S->x, S->y, n, p2, *g, S->cell_start, S->cell_items, S->n_items);
    ORB_CUDA(cudaGetLastError());
    int total = 0;
    ORB_CUDA(cudaMemcpyAsync(&total, S->n_items, sizeof(int), cudaMemcpyDeviceToHost, s));
    ORB_CUDA(cudaStreamSynchronize(s));
    if (total < 0 || total > n) { orb_set_error("orb_assign_features_to_grid_static: the scratch array was overwritten (overlapping calls?)"); return ORB_ERR_INTERNAL; }
    ORB_CUDA(cudaMemcpyAsync(cell_start, S->cell_start, (size_t)(GRID_CELLS + 1) * 4, cudaMemcpyDefault, s));
    ORB_CUDA(cudaMemcpyAsync(n_items, S->n_items, sizeof(int), cudaMemcpyDefault, s));
    if (total > 0) ORB_CUDA(cudaMemcpyAsync(cell_items, S->cell_items, (size_t)total * 4, cudaMemcpyDefault, s));
    ORB_CUDA(cudaStreamSynchronize(s));
    return ORB_OK;
}
