// Frame::ComputeStereoMatches (reference src/Frame.cc:516-704) on the device-resident outputs of
// two extractor handles.
//
//   k_stereo_rows    vRowIndices: right keypoint -> rows floor(y - 2 sf) .. ceil(y + 2 sf)      :527-544
//   k_stereo_search  per left keypoint (one warp): row-band Hamming search; writes the refinement job
//                    (window centres on the left keypoint's pyramid level)                        :558-633
//   k_stereo_refine  a warp walks 16 left keypoints: both SAD windows of a job arrive by TMA (double-buffered
//                    across jobs), 11x11 SAD over +-5 px on packed 16-bit lanes, parabola sub-pixel fit  :634-689
//   k_stereo_median  threshold 1.5 * 1.4 * median(SAD), invalidate outliers                      :690-703
//
// The per-row lists are built with atomics (unordered): the reference keeps the first candidate
// on equal distance, i.e. the lowest right index, which a (distance, index) minimum reproduces.
#include "orb_internal.h"
#include <cstring>
#include <chrono>
#include <cstdio>
#include <cstdlib>

#define ST_TH_HIGH 100
#define ST_TH_ORB 75  // (TH_HIGH + TH_LOW) / 2

struct StereoGeom {
    int nlevels;
    int h0;              // rows of level 0
    int kp_cap_l, kp_cap_r;
    int items_cap;       // per image
    float sf[ORB_MAX_LEVELS], inv_sf[ORB_MAX_LEVELS];
    LevelGeom lvl_l[ORB_MAX_LEVELS], lvl_r[ORB_MAX_LEVELS];
};

__global__ void __launch_bounds__(512)
k_stereo_rows(const orb_keypoint* __restrict__ kps_r, const int* __restrict__ counts_r, int* __restrict__ row_start,
              int2* __restrict__ row_items, int* __restrict__ status, const StereoGeom S) {
    extern __shared__ int s_rows[];  // h0 + 1 counters, then h0 + 1 cursors
    __shared__ int s_warp[17];
    ORB_PDL_SYNC();
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int H = S.h0;
    int* cnt = s_rows;
    int* cur = s_rows + H + 1;
    for (int i = tid; i <= H; i += 512) cnt[i] = 0;
    __syncthreads();
    const int nR = counts_r[b];
    const orb_keypoint* K = kps_r + (size_t)b * S.kp_cap_r;
    for (int i = tid; i < nR; i += 512) {
        const float y = K[i].y, r = __fmul_rn(2.0f, S.sf[K[i].octave]);
        const int maxr = min((int)ceilf(__fadd_rn(y, r)), H - 1), minr = max((int)floorf(__fsub_rn(y, r)), 0);
        for (int yi = minr; yi <= maxr; ++yi) atomicAdd(&cnt[yi], 1);
    }
    __syncthreads();
    // exclusive scan of cnt[0..H) -> cur, row_start
    int carry = 0;
    for (int i0 = 0; i0 < H; i0 += 512) {
        const int i = i0 + tid;
        const int v = i < H ? cnt[i] : 0;
        int inc = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, inc, o); if (lane >= o) inc += n; }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int w = lane < 16 ? s_warp[lane] : 0, wi = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int n = __shfl_up_sync(0xffffffffu, wi, o); if (lane >= o) wi += n; }
            if (lane < 16) s_warp[lane] = wi - w;
            if (lane == 31) s_warp[16] = wi;
        }
        __syncthreads();
        if (i < H) cur[i] = carry + inc - v + s_warp[warp];
        carry += s_warp[16];
        __syncthreads();
    }
    if (tid == 0) cur[H] = carry;
    __syncthreads();
    int* rs = row_start + (size_t)b * (ORB_MAX_DIM + 1);
    for (int i = tid; i <= H; i += 512) rs[i] = cur[i];
    if (carry > S.items_cap) { if (tid == 0) atomicExch(status, ORB_ERR_INTERNAL); return; }
    __syncthreads();
    // an item carries what the candidate filter of k_stereo_match reads (x, octave), so that the filter costs one load
    int2* items = row_items + (size_t)b * S.items_cap;
    for (int i = tid; i < nR; i += 512) {
        const int oct = K[i].octave;
        const float y = K[i].y, r = __fmul_rn(2.0f, S.sf[oct]);
        const int maxr = min((int)ceilf(__fadd_rn(y, r)), H - 1), minr = max((int)floorf(__fsub_rn(y, r)), 0);
        const int2 it = make_int2(__float_as_int(K[i].x), i | (oct << 24));
        for (int yi = minr; yi <= maxr; ++yi) items[atomicAdd(&cur[yi], 1)] = it;
    }
}

#define SM_WARPS 8
// Refinement job of one left keypoint: x = su | sv << 16 (window centre on level `oct` of the left image), y = sr | oct << 16 | 1 << 31
// (centre column in the right image; bit 31 = job valid).  y == 0: no match / rejected before the SAD.
// Eight lanes per left keypoint, four keypoints per warp: a row band holds a few dozen candidates of which the octave / disparity filter
// keeps about ten, so a whole warp per keypoint idles most lanes in the Hamming part; the chain of dependent loads (keypoint ->
// row range -> items -> descriptors) is the same length, but four of them are in flight per warp.
#define SM_GROUP 8
#define SM_KP_PER_WARP (32 / SM_GROUP)
__global__ void __launch_bounds__(SM_WARPS * 32)
k_stereo_search(const orb_keypoint* __restrict__ kps_l, const uint8_t* __restrict__ desc_l, const int* __restrict__ counts_l,
                const uint8_t* __restrict__ desc_r, const int* __restrict__ row_start, const int2* __restrict__ row_items,
                float* __restrict__ u_right, float* __restrict__ depth, int* __restrict__ sad, int2* __restrict__ jobs, const StereoGeom S,
                const float min_d, const float max_d) {
    ORB_PDL_SYNC();
    const int b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int sub = lane & (SM_GROUP - 1);
    const int iL = (blockIdx.x * SM_WARPS + warp) * SM_KP_PER_WARP + (lane / SM_GROUP);
    const int nL = counts_l[b];
    bool act = iL < nL;
    const size_t oL = (size_t)b * S.kp_cap_l + (act ? iL : 0);
    if (act && sub == 0) { u_right[oL] = -1.0f; depth[oL] = -1.0f; sad[oL] = -1; jobs[oL] = make_int2(0, 0); }
    float klx = 0.f, kly = 0.f;
    int oct = 0;
    if (act) { const orb_keypoint kl = kps_l[oL]; klx = kl.x; kly = kl.y; oct = kl.octave; }
    const float uL = klx, vL = kly;
    const int row = (int)vL;
    act = act && row >= 0 && row < S.h0;
    int c0 = 0, c1 = 0;
    if (act) {
        const int* rs = row_start + (size_t)b * (ORB_MAX_DIM + 1);
        c0 = rs[row]; c1 = rs[row + 1];
    }
    const float minU = __fsub_rn(uL, max_d), maxU = __fsub_rn(uL, min_d);
    act = act && c0 != c1 && !(maxU < 0);

    unsigned best = ((unsigned)ST_TH_HIGH << 20);  // (dist << 20) | iR ; strict < against TH_HIGH
    float best_x = 0.f;                             // x of this lane's best candidate
    if (act) {
        const uint4* dl = reinterpret_cast<const uint4*>(desc_l + oL * 32);
        const uint4 a0 = dl[0], a1 = dl[1];
        const int2* items = row_items + (size_t)b * S.items_cap;
        const uint8_t* DR = desc_r + (size_t)b * S.kp_cap_r * 32;
        for (int c = c0 + sub; c < c1; c += SM_GROUP) {
            const int2 it = __ldg(&items[c]);
            const int iR = it.y & 0xffffff, octR = it.y >> 24;
            const float xR = __int_as_float(it.x);
            if (octR < oct - 1 || octR > oct + 1) continue;
            if (xR >= minU && xR <= maxU) {
                const uint4* dr = reinterpret_cast<const uint4*>(DR + (size_t)iR * 32);
                const unsigned d = (unsigned)orb_hamming256(a0, a1, dr[0], dr[1]);
                const unsigned key = (d << 20) | (unsigned)iR;
                if (key < best) { best = key; best_x = xR; }
            }
        }
    }
    const unsigned mine = best;
#pragma unroll
    for (int o = SM_GROUP / 2; o; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
    const int bestDist = (int)(best >> 20);
    // the winning lane of the group hands over its candidate's x (keys are unique: they carry the candidate index)
    const unsigned win = (__ballot_sync(0xffffffffu, mine == best) >> (lane & ~(SM_GROUP - 1))) & ((1u << SM_GROUP) - 1u);
    const float uR0 = __shfl_sync(0xffffffffu, best_x, (lane & ~(SM_GROUP - 1)) + __ffs(win) - 1);
    if (!act || sub != 0 || bestDist >= ST_TH_ORB) return;

    // window centres on the left keypoint's pyramid level (reference :620-633)
    const float isf = S.inv_sf[oct];
    const int su = (int)roundf(__fmul_rn(klx, isf)), sv = (int)roundf(__fmul_rn(kly, isf)), sr = (int)roundf(__fmul_rn(uR0, isf));
    const LevelGeom& GL = S.lvl_l[oct];
    const LevelGeom& GR = S.lvl_r[oct];
    const int w = 5, L = 5;
    if (sr + L - w < 0 || sr + L + w + 1 >= GR.w) return;            // reference :633
    // the reference would throw from cv::Mat::rowRange/colRange on these; they cannot occur for
    // keypoints produced by the extractor (>= 19 px from every level edge)
    if (sv - w < 0 || sv + w >= GL.h || sv + w >= GR.h || su - w < 0 || su + w >= GL.w || sr - L - w < 0 || sr + L + w >= GR.w) return;
    jobs[oL] = make_int2(su | (sv << 16), sr | (oct << 16) | (int)0x80000000);
}

#define SR_WARPS 4
#define SR_KPW 16          // left keypoints walked by one warp
#define SR_LBOX 384        // left window box: 32 x 11 bytes (11 + up to 15 bytes of TMA alignment), rounded up to 128
#define SR_RBOX 640        // right window box: 48 x 11 bytes (21 + 15)
#define SR_BUF (SR_LBOX + SR_RBOX)
__device__ __forceinline__ unsigned st_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void st_mbar_wait(uint64_t* bar, unsigned phase) {
    unsigned ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(st_smem_u32(bar)), "r"(phase) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void st_tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(st_smem_u32(smem_dst)), "l"(map), "r"(st_smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}

// max of two packed 16-bit lanes as HMNMX2 (128 lanes/clk/SM, VIMNMX.U16x2: 64): valid for lanes below 0x7c00, bit-exact
__device__ __forceinline__ unsigned st_hmax2(unsigned a, unsigned b) { unsigned r; asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
// SAD of the mean-offset windows (reference :640-659): dist(s) = sum |(L - cl) - (R(s) - cr_s)| over 11 x 11, s = 0..10.
// Lane 2 y + par owns window row y and the shifts of parity par: the row's 11 left pixels and 22 right pixels are unpacked once into
// zero-extended byte pairs (PRMT), every |difference| is one IADD3 (biased so that no borrow crosses the 16-bit lanes), one subtract
// and one HMNMX2; row partials meet in shared memory.
__global__ void __launch_bounds__(SR_WARPS * 32)
k_stereo_refine(const __grid_constant__ TmaMaps tm_l, const __grid_constant__ TmaMaps tm_r, const orb_keypoint* __restrict__ kps_l,
                const int* __restrict__ counts_l, const int2* __restrict__ jobs, float* __restrict__ u_right, float* __restrict__ depth,
                int* __restrict__ sad, const StereoGeom S, const float mbf, const float min_d, const float max_d, const int b0) {
    __shared__ __align__(128) uint8_t s_buf[SR_WARPS][2][SR_BUF];
    __shared__ int s_part[SR_WARPS][11 * 12];
    __shared__ uint64_t s_bar[SR_WARPS][2];
    ORB_PDL_SYNC();
    const int b = blockIdx.y, lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(st_smem_u32(&s_bar[warp][0])) : "memory");
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(st_smem_u32(&s_bar[warp][1])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    const int i0 = (blockIdx.x * SR_WARPS + warp) * SR_KPW;
    const int iL = i0 + lane;
    int2 job = make_int2(0, 0);
    float uL = 0.f;
    if (lane < SR_KPW && iL < counts_l[b]) {
        job = __ldg(&jobs[(size_t)b * S.kp_cap_l + iL]);
        uL = kps_l[(size_t)b * S.kp_cap_l + iL].x;
    }
    unsigned todo = __ballot_sync(0xffffffffu, job.y < 0);
    if (todo == 0) return;
    const int img = b0 + b;

    auto prefetch = [&](int j, int slot) {
        const int jx = __shfl_sync(0xffffffffu, job.x, j), jy = __shfl_sync(0xffffffffu, job.y, j);
        if (lane == 0) {
            const int su = jx & 0xffff, sv = jx >> 16, sr = jy & 0xffff, oct = (jy >> 16) & 0xff;
            uint64_t* bar = &s_bar[warp][slot];
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(st_smem_u32(bar)), "r"(32 * 11 + 48 * 11) : "memory");
            st_tma_load_3d(&s_buf[warp][slot][0], &tm_l.lv[oct], bar, (su - 5) & ~15, sv - 5, img);
            st_tma_load_3d(&s_buf[warp][slot][SR_LBOX], &tm_r.lv[oct], bar, (sr - 10) & ~15, sv - 5, img);
        }
    };
    int cur = __ffs(todo) - 1;
    todo &= todo - 1;
    prefetch(cur, 0);
    int* part = s_part[warp];
#pragma unroll 1
    for (int it = 0; cur >= 0; ++it) {
        const int nxt = todo ? __ffs(todo) - 1 : -1;
        todo &= todo - 1;
        if (nxt >= 0) prefetch(nxt, (it + 1) & 1);   // its buffer was last read by job it - 1, which ended with __syncwarp
        const int jx = __shfl_sync(0xffffffffu, job.x, cur), jy = __shfl_sync(0xffffffffu, job.y, cur);
        const float uLc = __shfl_sync(0xffffffffu, uL, cur);
        const int su = jx & 0xffff, sr = jy & 0xffff, oct = (jy >> 16) & 0xff;
        const uint8_t* bl = &s_buf[warp][it & 1][0];
        const uint8_t* br = bl + SR_LBOX;
        const int alL = (su - 5) & 15, alR = (sr - 10) & 15;
        st_mbar_wait(&s_bar[warp][it & 1], (unsigned)((it >> 1) & 1));

        if (lane < 22) {
            const int y = lane >> 1, par = lane & 1;
            unsigned a[6], bp[11];
            {
                const uint32_t* lw = reinterpret_cast<const uint32_t*>(bl + y * 32) + (alL >> 2);
                const unsigned l0 = lw[0], l1 = lw[1], l2 = lw[2], l3 = lw[3], sh = (unsigned)(alL & 3) * 8u;
                const unsigned A0 = __funnelshift_r(l0, l1, sh), A1 = __funnelshift_r(l1, l2, sh), A2 = __funnelshift_r(l2, l3, sh);
                a[0] = __byte_perm(A0, 0u, 0x4140); a[1] = __byte_perm(A0, 0u, 0x4342);
                a[2] = __byte_perm(A1, 0u, 0x4140); a[3] = __byte_perm(A1, 0u, 0x4342);
                a[4] = __byte_perm(A2, 0u, 0x4140); a[5] = __byte_perm(A2, 0u, 0x4442);   // pixel 10 alone
            }
            {
                const int off = alR + par;
                const uint32_t* rw = reinterpret_cast<const uint32_t*>(br + y * 48) + (off >> 2);
                const unsigned sh = (unsigned)(off & 3) * 8u;
                unsigned r[7];
#pragma unroll
                for (int i = 0; i < 7; ++i) r[i] = rw[i];
#pragma unroll
                for (int i = 0; i < 6; ++i) {
                    const unsigned B = __funnelshift_r(r[i], r[i + 1], sh);   // bytes par + 4 i .. par + 4 i + 3 of the row
                    bp[2 * i] = __byte_perm(B, 0u, 0x4140);
                    if (2 * i + 1 < 11) bp[2 * i + 1] = __byte_perm(B, 0u, 0x4342);
                }
            }
            const int cl = bl[5 * 32 + alL + 5];
            const uint8_t* crp = br + 5 * 48 + alR + 5 + par;
#pragma unroll
            for (int sp = 0; sp < 6; ++sp) {   // shift s = 2 sp + par (s = 11 of the odd lanes is not a shift: discarded below)
                const unsigned bias = (unsigned)(1024 - cl + (int)crp[2 * sp]) * 0x10001u;
                unsigned acc = 0;
#pragma unroll
                for (int j = 0; j < 6; ++j) {
                    const unsigned u = a[j] + bias - bp[sp + j];          // 1024 + (L - cl) - (R - cr) on both 16-bit lanes
                    unsigned m = st_hmax2(u, 0x08000800u - u);            // 1024 + |.| (lanes in [514, 1534]: positive halves, same order as integers)
                    if (j == 5) m &= 0xffffu;
                    acc += m;
                }
                if (2 * sp + par < 11) part[(2 * sp + par) * 12 + y] = (int)((acc & 0xffffu) + (acc >> 16)) - 11 * 1024;
            }
        }
        __syncwarp();
        int dist = 0;
        if (lane < 11) {
#pragma unroll
            for (int y = 0; y < 11; ++y) dist += part[lane * 12 + y];
        }
        // first minimum over s = 0..10
        unsigned key = lane < 11 ? (((unsigned)dist << 8) | (unsigned)lane) : 0xffffffffu;
        key = __reduce_min_sync(0xffffffffu, key);   // REDUX.MIN
        const int L = 5;
        const int bs = (int)(key & 0xffu), bestSad = (int)(key >> 8);
        const int bestinc = bs - L;
        const float d1 = (float)__shfl_sync(0xffffffffu, dist, max(bs - 1, 0));
        const float d2 = (float)bestSad;
        const float d3 = (float)__shfl_sync(0xffffffffu, dist, min(bs + 1, 10));
        if (lane == 0 && bestinc != -L && bestinc != L) {
            const float deltaR = __fdiv_rn(__fsub_rn(d1, d3), __fmul_rn(2.0f, __fsub_rn(__fadd_rn(d1, d3), __fmul_rn(2.0f, d2))));
            if (!(deltaR < -1 || deltaR > 1)) {
                float bestuR = __fmul_rn(S.sf[oct], __fadd_rn(__fadd_rn((float)sr, (float)bestinc), deltaR));
                float disparity = __fsub_rn(uLc, bestuR);
                if (disparity >= min_d && disparity < max_d) {
                    if (disparity <= 0) { disparity = 0.01f; bestuR = (float)((double)uLc - 0.01); }
                    const size_t oL = (size_t)b * S.kp_cap_l + i0 + cur;
                    depth[oL] = __fdiv_rn(mbf, disparity);
                    u_right[oL] = bestuR;
                    sad[oL] = bestSad;
                }
            }
        }
        __syncwarp();   // every lane is done with this buffer and with `part`
        cur = nxt;
    }
}

__global__ void __launch_bounds__(512)
k_stereo_median(const int* __restrict__ counts_l, float* __restrict__ u_right, float* __restrict__ depth, const int* __restrict__ sad,
                const int kp_cap) {
    __shared__ int hist[256];
    __shared__ int s_sel[3];
    ORB_PDL_SYNC();
    const int b = blockIdx.x, tid = threadIdx.x;
    const int n = counts_l[b];
    const int* sd = sad + (size_t)b * kp_cap;
    // SAD <= 121 * 510 < 2^16: two-pass radix select of the element at sorted position size/2
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 512) { const int v = sd[i]; if (v >= 0) atomicAdd(&hist[(v >> 8) & 0xff], 1); }
    __syncthreads();
    if (tid == 0) {
        int total = 0;
        for (int i = 0; i < 256; ++i) total += hist[i];
        int k = total / 2, hi = 0;
        if (total > 0) { while (k >= hist[hi]) { k -= hist[hi]; ++hi; } }
        s_sel[0] = total; s_sel[1] = hi; s_sel[2] = k;
    }
    __syncthreads();
    if (s_sel[0] == 0) return;  // reference indexes an empty vector here (undefined); nothing to invalidate
    const int hi = s_sel[1];
    if (tid < 256) hist[tid] = 0;
    __syncthreads();
    for (int i = tid; i < n; i += 512) { const int v = sd[i]; if (v >= 0 && ((v >> 8) & 0xff) == hi) atomicAdd(&hist[v & 0xff], 1); }
    __syncthreads();
    if (tid == 0) {
        int k = s_sel[2], lo = 0;
        while (k >= hist[lo]) { k -= hist[lo]; ++lo; }
        s_sel[1] = (hi << 8) | lo;
    }
    __syncthreads();
    const float median = (float)s_sel[1];
    const float th = __fmul_rn(1.5f * 1.4f, median);
    for (int i = tid; i < n; i += 512) {
        const int v = sd[i];
        if (v >= 0 && !((float)v < th)) { u_right[(size_t)b * kp_cap + i] = -1.0f; depth[(size_t)b * kp_cap + i] = -1.0f; }
    }
}

// Image slots [b0, b0 + batch) of the two handles' last batches.  fence_right: make later work on
// the right stream wait for this stereo pass (needed when the next call reuses the same slots).
static int stereo_launch(orb_extractor* l, orb_extractor* r, int b0, int batch, float mbf, float min_d, float max_d, bool fence_right = true) {
    if (!l || !r || batch < 1 || b0 < 0 || l->plan_w < 0 || r->plan_w < 0 || b0 + batch > l->last_batch || b0 + batch > r->last_batch) return ORB_ERR_BAD_ARG;
    if (l->device != r->device || l->nlevels != r->nlevels || l->plan_w != r->plan_w || l->plan_h != r->plan_h) {
        orb_set_error("left and right extractor must share device, level count and image size");
        return ORB_ERR_BAD_ARG;
    }
    ORB_CUDA(cudaSetDevice(l->device));
    StereoGeom S;
    S.nlevels = l->nlevels; S.h0 = l->plan.lv[0].h; S.kp_cap_l = l->plan.kp_cap; S.kp_cap_r = r->plan.kp_cap;
    S.items_cap = l->row_items_cap;
    for (int i = 0; i < l->nlevels; ++i) {
        S.sf[i] = l->sf[i]; S.inv_sf[i] = l->inv_sf[i]; S.lvl_l[i] = l->plan.lv[i]; S.lvl_r[i] = r->plan.lv[i];
        S.lvl_l[i].off += (unsigned long long)b0 * S.lvl_l[i].img_stride;
        S.lvl_r[i].off += (unsigned long long)b0 * S.lvl_r[i].img_stride;
    }
    const size_t kl = (size_t)b0 * S.kp_cap_l, kr = (size_t)b0 * S.kp_cap_r;
    int* row_start = l->d_row_start + (size_t)b0 * (ORB_MAX_DIM + 1);
    int2* row_items = l->d_row_items + (size_t)b0 * S.items_cap;
    // the right handle's stream must have finished producing its outputs
    ORB_CUDA(cudaEventRecord(r->ev_done, r->stream));
    ORB_CUDA(cudaStreamWaitEvent(l->stream, r->ev_done, 0));
    const size_t smem = (size_t)(S.h0 + 1) * 2 * sizeof(int);
    if (smem > 48 * 1024) ORB_CUDA(cudaFuncSetAttribute(k_stereo_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const bool T = l->timing;
    if (T) cudaEventRecord(l->ev_stage[6], l->stream);
    nvtxRangePushA("orb:stereo_rows");
    const bool pdl = orb_pdl(batch);
    // the first kernel behind the cross-stream event wait above is launched plainly: only links whose predecessor is a kernel of this chain are relaxed
    orb_launch(false, k_stereo_rows, dim3(batch), dim3(512), smem, l->stream, r->d_kps + kr, r->d_counts + b0, row_start, row_items, l->d_status, S);
    nvtxRangePop();
    if (T) cudaEventRecord(l->ev_stage[7], l->stream);
    nvtxRangePushA("orb:stereo_match");
    dim3 grid((l->plan.kp_cap + SM_WARPS * SM_KP_PER_WARP - 1) / (SM_WARPS * SM_KP_PER_WARP), batch);
    int2* jobs = l->d_st_jobs + kl;
    orb_launch(pdl, k_stereo_search, grid, dim3(SM_WARPS * 32), 0, l->stream, l->d_kps + kl, l->d_desc + kl * 32, l->d_counts + b0, r->d_desc + kr * 32,
               row_start, row_items, l->d_u_right + kl, l->d_depth + kl, l->d_sad + kl, jobs, S, min_d, max_d);
    dim3 grid2((l->plan.kp_cap + SR_WARPS * SR_KPW - 1) / (SR_WARPS * SR_KPW), batch);
    orb_launch(pdl, k_stereo_refine, grid2, dim3(SR_WARPS * 32), 0, l->stream, l->tm_st_l, r->tm_st_r, l->d_kps + kl, l->d_counts + b0, jobs, l->d_u_right + kl,
               l->d_depth + kl, l->d_sad + kl, S, mbf, min_d, max_d, b0);
    nvtxRangePop();
    if (T) cudaEventRecord(l->ev_stage[8], l->stream);
    nvtxRangePushA("orb:stereo_median");
    orb_launch(pdl, k_stereo_median, dim3(batch), dim3(512), 0, l->stream, l->d_counts + b0, l->d_u_right + kl, l->d_depth + kl, l->d_sad + kl, l->plan.kp_cap);
    nvtxRangePop();
    if (T) { cudaEventRecord(l->ev_stage[9], l->stream); l->stage_valid[5] = l->stage_valid[6] = l->stage_valid[7] = true; }
    l->launches += 4;
    ORB_CUDA(cudaGetLastError());
    if (fence_right) {
        // later work on the right stream must not overwrite its buffers before we are done
        ORB_CUDA(cudaEventRecord(l->ev_done, l->stream));
        ORB_CUDA(cudaStreamWaitEvent(r->stream, l->ev_done, 0));
    }
    return ORB_OK;
}
static int stereo_launch(orb_extractor* l, orb_extractor* r, int batch, float mbf, float min_d, float max_d) {
    return stereo_launch(l, r, 0, batch, mbf, min_d, max_d, true);
}

extern "C" {

int orb_stereo_match_device(orb_extractor* l, orb_extractor* r, int batch, float mbf, float min_d, float max_d) {
    return stereo_launch(l, r, batch, mbf, min_d, max_d);
}

int orb_stereo_device_results(const orb_extractor* l, const float** d_u_right, const float** d_depth) {
    if (!l) return ORB_ERR_BAD_ARG;
    if (d_u_right) *d_u_right = l->d_u_right;
    if (d_depth) *d_depth = l->d_depth;
    return ORB_OK;
}

int orb_stereo_match_batch(orb_extractor* l, orb_extractor* r, int batch, float mbf, float min_d, float max_d, float* u_right, float* depth,
                           int cap, int32_t* n_left) {
    if (!u_right || !depth || !n_left) return ORB_ERR_BAD_ARG;
    int st = stereo_launch(l, r, batch, mbf, min_d, max_d);
    if (st != ORB_OK) return st;
    ORB_CUDA(cudaMemcpyAsync(l->h_counts, l->d_counts, (size_t)batch * sizeof(int), cudaMemcpyDeviceToHost, l->stream));
    int dst = 0;
    ORB_CUDA(cudaMemcpyAsync(&dst, l->d_status, sizeof(int), cudaMemcpyDeviceToHost, l->stream));
    ORB_CUDA(cudaStreamSynchronize(l->stream));
    if (dst != 0) { cudaMemsetAsync(l->d_status, 0, sizeof(int), l->stream); orb_set_error("device-side bound exceeded"); return dst; }
    int rc = ORB_OK;
    for (int b = 0; b < batch; ++b) {
        const int n = l->h_counts[b];
        n_left[b] = n;
        if (n > cap) { rc = ORB_ERR_CAPACITY; continue; }
        if (n == 0) continue;
        ORB_CUDA(cudaMemcpyAsync(u_right + (size_t)b * cap, l->d_u_right + (size_t)b * l->plan.kp_cap, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, l->stream));
        ORB_CUDA(cudaMemcpyAsync(depth + (size_t)b * cap, l->d_depth + (size_t)b * l->plan.kp_cap, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, l->stream));
    }
    ORB_CUDA(cudaStreamSynchronize(l->stream));
    if (rc != ORB_OK) orb_set_error("stereo output capacity too small");
    return rc;
}

// Asynchronous form of the stereo Frame constructor: enqueue() queues uploads, kernels and downloads of the whole batch on the
// handle pair's three streams and returns; wait() blocks until that call's outputs are in the caller's buffers.  Two calls may be in
// flight on one handle pair (tickets alternate): chunk c of the next call only waits for chunk c of the previous one (its kernels
// before the slot is uploaded again, its downloads before the slot's results are overwritten), so the first upload and the last
// download of consecutive calls overlap with kernels instead of being exposed.  The caller keeps the input and output buffers of
// an enqueued call alive and untouched until its wait() returns.
int orb_stereo_frames_enqueue(orb_extractor* l, orb_extractor* r, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch, int w, int h, int stride,
                              size_t img_stride, float mbf, float min_d, float max_d, orb_keypoint* kps_l, uint8_t* desc_l, orb_keypoint* kps_r,
                              uint8_t* desc_r, float* u_right, float* depth, int cap, int* ticket) {
    if (!l || !r || l == r || batch < 1 || !ticket) return ORB_ERR_BAD_ARG;
    if (!imgs_l || !imgs_r || !kps_l || !desc_l || !kps_r || !desc_r || !u_right || !depth || w < 1 || h < 1 || stride < w) return ORB_ERR_BAD_ARG;
    if (l->device != r->device) { orb_set_error("left and right extractor must share a device"); return ORB_ERR_BAD_ARG; }
    if (l->inflight >= 2) { orb_set_error("two stereo batches are already in flight on this handle pair: wait for one"); return ORB_ERR_CAPACITY; }
    ORB_CUDA(cudaSetDevice(l->device));
    // a different batch size or geometry re-lays the level buffers out: only with nothing in flight
    if (l->inflight > 0 && (l->plan_w != w || l->plan_h != h || l->last_batch != batch)) {
        orb_set_error("geometry / batch size must not change while a stereo batch is in flight");
        return ORB_ERR_BAD_ARG;
    }
    int st = orb_ensure_plan(l, w, h, batch);
    if (st == ORB_OK) st = orb_ensure_plan(r, w, h, batch);
    if (st == ORB_OK) st = orb_pipe_streams(l);
    if (st == ORB_OK) st = orb_pipe_streams(r);   // the right handle's upload staging area
    if (st != ORB_OK) return st;
    const std::vector<int> cb = orb_chunk_bounds(batch);
    const int nch = (int)cb.size() - 1;
    // events per chunk: upload left / right done, kernels done, downloads done; + 2 for the call; then ingest left / right done per chunk;
    // then one for "left extraction done" of a one-chunk call
    if (!orb_pool_event(l, 6 * nch + 5)) { orb_set_error("event"); return ORB_ERR_CUDA; }
    // dense images + staging area: copies on the copy stream, ingest kernels on the kernel streams (uploads run a call ahead)
    const bool split = orb_upload_can_split(l, w, h, stride, img_stride) && orb_upload_can_split(r, w, h, stride, img_stride);
    const bool was_split = l->pipe_split;
    const int tk = l->next_ticket & 1;
    // uploads must not overtake work already queued on the two kernel streams by other entry points
    if (l->inflight == 0) {
        ORB_CUDA(cudaEventRecord(l->ev_pool[4 * nch], l->stream));
        ORB_CUDA(cudaEventRecord(l->ev_pool[4 * nch + 1], r->stream));
        ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, l->ev_pool[4 * nch], 0));
        ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, l->ev_pool[4 * nch + 1], 0));
    }
    const bool chained = l->inflight > 0 && l->pipe_chunks == nch;   // the previous call used the same chunking: chunk-granular hand-over
    if (l->inflight > 0 && !chained) {
        ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, l->ev_call[tk ^ 1], 0));
        ORB_CUDA(cudaStreamWaitEvent(l->stream, l->ev_call[tk ^ 1], 0));
        ORB_CUDA(cudaStreamWaitEvent(r->stream, l->ev_call[tk ^ 1], 0));
    }
    const int kc = l->plan.kp_cap, m = cap < kc ? cap : kc;
    int* hc_l = l->h_counts + (size_t)tk * l->max_batch;
    int* hc_r = r->h_counts + (size_t)tk * r->max_batch;
    // one chunk, results into pageable memory, nothing else in flight: the downloads go through the handles' pinned mirrors (orb_internal.h)
    const bool mirror = nch == 1 && !l->copy_only && l->inflight == 0 && orb_host_pageable(kps_l) && orb_host_pageable(desc_l) && orb_host_pageable(kps_r) &&
                        orb_host_pageable(desc_r) && orb_host_pageable(u_right) && orb_host_pageable(depth) && orb_out_mirror_ready(l, batch, true) &&
                        orb_out_mirror_ready(r, batch, false);
    for (int c = 0; c < nch; ++c) {
        const int b0 = cb[c], n = cb[c + 1] - cb[c];
        cudaEvent_t up_l = l->ev_pool[4 * c], up_r = l->ev_pool[4 * c + 1], done = l->ev_pool[4 * c + 2], dl = l->ev_pool[4 * c + 3];
        cudaEvent_t ing_l = l->ev_pool[4 * nch + 4 + 2 * c], ing_r = l->ev_pool[4 * nch + 5 + 2 * c];
        if (chained) {   // events still hold the previous call's records for this chunk
            if (split && was_split) {   // the copy only overwrites the staging slots: their previous ingest must be done
                ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, ing_l, 0));
                ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, ing_r, 0));
                ORB_CUDA(cudaStreamWaitEvent(l->stream, done, 0));   // the ingest overwrites level 0: every reader of the slot must be done
                ORB_CUDA(cudaStreamWaitEvent(r->stream, done, 0));
            } else {
                ORB_CUDA(cudaStreamWaitEvent(l->s_h2d, done, 0));     // its kernels have read the slot's images
            }
            ORB_CUDA(cudaStreamWaitEvent(l->stream, dl, 0));      // its results have left the slot
            ORB_CUDA(cudaStreamWaitEvent(r->stream, dl, 0));
        }
        // Frame::Frame: ExtractORB(0, left) and ExtractORB(1, right) run side by side (src/Frame.cc:81-84) ...
        const bool skip = l->copy_only;
        if (split) {
            // A call of one chunk (a pair or a few: the per-frame use) queues the left kernels BEFORE the right copy: a copy from pageable
            // memory holds the calling thread until the bytes are staged, and that time is now spent while the left image is being worked on.
            // The copies keep their order on the copy stream either way; pinned buffers and long batches (copies a call ahead) are unchanged.
            const bool left_first = nch == 1 && orb_host_pageable(imgs_r);
            if ((st = orb_upload_copy_slots(l, imgs_l, b0, n, w, h, l->s_h2d)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(up_l, l->s_h2d));
            if (!left_first) {
                if ((st = orb_upload_copy_slots(r, imgs_r, b0, n, w, h, l->s_h2d)) != ORB_OK) break;
                ORB_CUDA(cudaEventRecord(up_r, l->s_h2d));
            }
            ORB_CUDA(cudaStreamWaitEvent(l->stream, up_l, 0));
            if ((st = orb_ingest_slots(l, b0, n, w, h, l->stream)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(ing_l, l->stream));
            // left_first: only the left pyramid goes out before the right copy, then the right pyramid, then the rest of both, so that the
            // right image's kernel chain (the longer wait: it starts last) begins as early as the calling thread allows
            if (!skip && (st = orb_run_slots(l, b0, n, left_first ? 1 : 0)) != ORB_OK) break;
            if (nch == 1 && !left_first) ORB_CUDA(cudaEventRecord(l->ev_pool[6 * nch + 4], l->stream));
            if (left_first) {
                if ((st = orb_upload_copy_slots(r, imgs_r, b0, n, w, h, l->s_h2d)) != ORB_OK) break;
                ORB_CUDA(cudaEventRecord(up_r, l->s_h2d));
            }
            ORB_CUDA(cudaStreamWaitEvent(r->stream, up_r, 0));
            if ((st = orb_ingest_slots(r, b0, n, w, h, r->stream)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(ing_r, r->stream));
            if (left_first) {
                if (!skip && ((st = orb_run_slots(r, b0, n, 1)) != ORB_OK || (st = orb_run_slots(l, b0, n, 2)) != ORB_OK)) break;
                ORB_CUDA(cudaEventRecord(l->ev_pool[6 * nch + 4], l->stream));
                if (!skip && (st = orb_run_slots(r, b0, n, 2)) != ORB_OK) break;
            } else if (!skip && (st = orb_run_slots(r, b0, n)) != ORB_OK) break;
        } else {
            const bool left_first = nch == 1 && orb_host_pageable(imgs_r);   // as above
            if ((st = orb_upload_slots(l, imgs_l, b0, n, w, h, stride, img_stride, l->s_h2d)) != ORB_OK) break;
            ORB_CUDA(cudaEventRecord(up_l, l->s_h2d));
            if (!left_first) {
                if ((st = orb_upload_slots(r, imgs_r, b0, n, w, h, stride, img_stride, l->s_h2d)) != ORB_OK) break;
                ORB_CUDA(cudaEventRecord(up_r, l->s_h2d));
            }
            ORB_CUDA(cudaStreamWaitEvent(l->stream, up_l, 0));
            if (!skip && (st = orb_run_slots(l, b0, n, left_first ? 1 : 0)) != ORB_OK) break;
            if (nch == 1 && !left_first) ORB_CUDA(cudaEventRecord(l->ev_pool[6 * nch + 4], l->stream));
            if (left_first) {
                if ((st = orb_upload_slots(r, imgs_r, b0, n, w, h, stride, img_stride, l->s_h2d)) != ORB_OK) break;
                ORB_CUDA(cudaEventRecord(up_r, l->s_h2d));
            }
            ORB_CUDA(cudaStreamWaitEvent(r->stream, up_r, 0));
            if (left_first) {
                if (!skip && ((st = orb_run_slots(r, b0, n, 1)) != ORB_OK || (st = orb_run_slots(l, b0, n, 2)) != ORB_OK)) break;
                ORB_CUDA(cudaEventRecord(l->ev_pool[6 * nch + 4], l->stream));
                if (!skip && (st = orb_run_slots(r, b0, n, 2)) != ORB_OK) break;
            } else if (!skip && (st = orb_run_slots(r, b0, n)) != ORB_OK) break;
        }
        // ... then ComputeStereoMatches (:95)
        if (!skip && (st = stereo_launch(l, r, b0, n, mbf, min_d, max_d, false)) != ORB_OK) break;
        if (skip) {   // copy-only diagnostic: keep the stream dependencies of the real pipeline
            ORB_CUDA(cudaEventRecord(l->ev_done, r->stream));
            ORB_CUDA(cudaStreamWaitEvent(l->stream, l->ev_done, 0));
        }
        ORB_CUDA(cudaEventRecord(done, l->stream));
        if (nch == 1 && !skip) {
            // one chunk: the keypoints and descriptors of a side leave as soon as that side's extraction is done, beside the stereo kernels
            // (ext_l: recorded below the left extraction; r->ev_done: recorded by stereo_launch behind the right one)
            ORB_CUDA(cudaStreamWaitEvent(l->s_d2h, l->ev_pool[6 * nch + 4], 0));
            if ((st = mirror ? orb_download_mirror(l, n, tk, l->s_d2h, hc_l) : orb_download_slots(l, b0, n, kps_l, desc_l, cap, l->s_d2h, hc_l)) != ORB_OK) break;
            ORB_CUDA(cudaStreamWaitEvent(l->s_d2h, r->ev_done, 0));
            if ((st = mirror ? orb_download_mirror(r, n, tk, l->s_d2h, hc_r) : orb_download_slots(r, b0, n, kps_r, desc_r, cap, l->s_d2h, hc_r)) != ORB_OK) break;
            ORB_CUDA(cudaStreamWaitEvent(l->s_d2h, done, 0));
        } else {
        ORB_CUDA(cudaStreamWaitEvent(l->s_d2h, done, 0));
        if ((st = orb_download_slots(l, b0, n, kps_l, desc_l, cap, l->s_d2h, hc_l)) != ORB_OK) break;
        if ((st = orb_download_slots(r, b0, n, kps_r, desc_r, cap, l->s_d2h, hc_r)) != ORB_OK) break;
        }
        if (mirror) {
            if ((st = orb_download_mirror_stereo(l, n, tk, l->s_d2h)) != ORB_OK) break;
            orb_extractor::OutPlan& OL = l->out_plan[tk];
            OL.on = true; OL.n = n; OL.cap = cap; OL.kps = kps_l; OL.desc = desc_l; OL.u_right = u_right; OL.depth = depth;
            orb_extractor::OutPlan& OR = r->out_plan[tk];
            OR.on = true; OR.n = n; OR.cap = cap; OR.kps = kps_r; OR.desc = desc_r; OR.u_right = nullptr; OR.depth = nullptr;
        } else if (m > 0) {
            ORB_CUDA(cudaMemcpy2DAsync(u_right + (size_t)b0 * cap, (size_t)cap * 4, l->d_u_right + (size_t)b0 * kc, (size_t)kc * 4, (size_t)m * 4, n,
                                       cudaMemcpyDeviceToHost, l->s_d2h));
            ORB_CUDA(cudaMemcpy2DAsync(depth + (size_t)b0 * cap, (size_t)cap * 4, l->d_depth + (size_t)b0 * kc, (size_t)kc * 4, (size_t)m * 4, n,
                                       cudaMemcpyDeviceToHost, l->s_d2h));
        }
        ORB_CUDA(cudaEventRecord(dl, l->s_d2h));
    }
    if (st != ORB_OK) {   // nothing of this call may stay in flight once the caller is told it failed (its buffers may go away)
        cudaStreamSynchronize(l->s_h2d); cudaStreamSynchronize(l->stream); cudaStreamSynchronize(r->stream); cudaStreamSynchronize(l->s_d2h);
        l->inflight = 0; l->pipe_chunks = 0;
        l->out_plan[tk].on = r->out_plan[tk].on = false;
        return st;
    }
    ORB_CUDA(cudaMemcpyAsync(&l->h_status[2 * tk], l->d_status, sizeof(int), cudaMemcpyDeviceToHost, l->s_d2h));
    ORB_CUDA(cudaMemcpyAsync(&l->h_status[2 * tk + 1], r->d_status, sizeof(int), cudaMemcpyDeviceToHost, l->s_d2h));
    ORB_CUDA(cudaEventRecord(l->ev_call[tk], l->s_d2h));
    l->pipe_chunks = nch; l->pipe_split = split;
    l->pipe_batch[tk] = batch; l->pipe_cap[tk] = cap;
    *ticket = l->next_ticket++;
    l->inflight++;
    return ORB_OK;
}

int orb_stereo_frames_wait(orb_extractor* l, orb_extractor* r, int ticket, int32_t* n_l, int32_t* n_r) {
    if (!l || !r || !n_l || !n_r) return ORB_ERR_BAD_ARG;
    if (l->inflight < 1 || ticket != l->next_ticket - l->inflight) { orb_set_error("tickets are waited for in the order they were issued"); return ORB_ERR_BAD_ARG; }
    ORB_CUDA(cudaSetDevice(l->device));
    const int tk = ticket & 1;
    ORB_CUDA(cudaEventSynchronize(l->ev_call[tk]));
    l->inflight--;
    if (l->inflight == 0) {
        // the right stream has nothing queued beyond what the downloads waited for, but later calls on the right handle alone
        // must see it idle
        ORB_CUDA(cudaStreamSynchronize(r->stream));
    }
    const int st_l = l->h_status[2 * tk], st_r = l->h_status[2 * tk + 1];
    if (st_l != 0 || st_r != 0) {
        l->out_plan[tk].on = r->out_plan[tk].on = false;
        cudaMemsetAsync(l->d_status, 0, sizeof(int), l->stream); cudaMemsetAsync(r->d_status, 0, sizeof(int), r->stream);
        orb_set_error("device-side bound exceeded");
        return st_l ? st_l : st_r;
    }
    const int batch = l->pipe_batch[tk], cap = l->pipe_cap[tk];
    const int* hc_l = l->h_counts + (size_t)tk * l->max_batch;
    const int* hc_r = r->h_counts + (size_t)tk * r->max_batch;
    orb_deliver_mirror(l, tk, hc_l);
    orb_deliver_mirror(r, tk, hc_r);
    int rc = ORB_OK;
    for (int b = 0; b < batch; ++b) {
        n_l[b] = hc_l[b]; n_r[b] = hc_r[b];
        if (n_l[b] > cap || n_r[b] > cap) rc = ORB_ERR_CAPACITY;
    }
    if (rc != ORB_OK) orb_set_error("keypoint capacity too small");
    return rc;
}

int orb_stereo_frames(orb_extractor* l, orb_extractor* r, const uint8_t* imgs_l, const uint8_t* imgs_r, int batch, int w, int h, int stride,
                      size_t img_stride, float mbf, float min_d, float max_d, orb_keypoint* kps_l, uint8_t* desc_l, int32_t* n_l,
                      orb_keypoint* kps_r, uint8_t* desc_r, int32_t* n_r, float* u_right, float* depth, int cap) {
    if (!l || !r || l == r || batch < 0 || !n_l || !n_r) return ORB_ERR_BAD_ARG;
    if (batch == 0) return ORB_OK;
    if (w == 0 || h == 0) {  // empty images: no keypoints, nothing to match (reference src/ORBextractor.cc:1095)
        for (int b = 0; b < batch; ++b) n_l[b] = n_r[b] = 0;
        return ORB_OK;
    }
    if (l->inflight > 0) { orb_set_error("an enqueued stereo batch is still in flight on this handle pair"); return ORB_ERR_BAD_ARG; }
    int ticket = 0;
    int st = orb_stereo_frames_enqueue(l, r, imgs_l, imgs_r, batch, w, h, stride, img_stride, mbf, min_d, max_d, kps_l, desc_l, kps_r, desc_r,
                                       u_right, depth, cap, &ticket);
    if (st != ORB_OK) return st;
    return orb_stereo_frames_wait(l, r, ticket, n_l, n_r);
}

int orb_extractor_set_copy_only(orb_extractor* e, int on) {
    if (!e) return ORB_ERR_BAD_ARG;
    e->copy_only = on != 0;
    return ORB_OK;
}

int orb_stereo_match(orb_extractor* l, orb_extractor* r, int image, float mbf, float min_d, float max_d, float* u_right, float* depth, int cap,
                     int32_t* n_left) {
    // single image of the last batch: run the batch prefix up to `image` is wasteful, so the
    // kernels are launched for the whole last batch only when image == 0 and batch == 1;
    // otherwise compute all and copy the requested one.
    if (!l || !u_right || !depth || !n_left || image < 0 || image >= l->last_batch) return ORB_ERR_BAD_ARG;
    int st = stereo_launch(l, r, l->last_batch, mbf, min_d, max_d);
    if (st != ORB_OK) return st;
    if (l->inflight == 0 && l->h_status && l->h_counts && orb_host_pageable(u_right) && orb_host_pageable(depth) && orb_out_mirror_ready(l, 1, true)) {
        // results into pageable memory: count, status and both rows are queued behind the kernels into pinned memory (four blocking
        // copies otherwise), one wait, then the CPU hands the n valid floats of each row over
        const float *h_ur = nullptr, *h_dp = nullptr;
        ORB_CUDA(cudaMemcpyAsync(l->h_counts + image, l->d_counts + image, sizeof(int), cudaMemcpyDeviceToHost, l->stream));
        ORB_CUDA(cudaMemcpyAsync(&l->h_status[0], l->d_status, sizeof(int), cudaMemcpyDeviceToHost, l->stream));
        if ((st = orb_download_mirror_stereo_row(l, image, l->stream, &h_ur, &h_dp)) != ORB_OK) return st;
        ORB_CUDA(cudaStreamSynchronize(l->stream));
        const int n = l->h_counts[image], dst = l->h_status[0];
        if (dst != 0) { cudaMemsetAsync(l->d_status, 0, sizeof(int), l->stream); orb_set_error("device-side bound exceeded"); return dst; }
        *n_left = n;
        if (n > cap) { orb_set_error("stereo output capacity too small"); return ORB_ERR_CAPACITY; }
        if (n > 0) { std::memcpy(u_right, h_ur, (size_t)n * sizeof(float)); std::memcpy(depth, h_dp, (size_t)n * sizeof(float)); }
        return ORB_OK;
    }
    int n = 0, dst = 0;
    ORB_CUDA(cudaMemcpyAsync(&n, l->d_counts + image, sizeof(int), cudaMemcpyDeviceToHost, l->stream));
    ORB_CUDA(cudaMemcpyAsync(&dst, l->d_status, sizeof(int), cudaMemcpyDeviceToHost, l->stream));
    ORB_CUDA(cudaStreamSynchronize(l->stream));
    if (dst != 0) { cudaMemsetAsync(l->d_status, 0, sizeof(int), l->stream); orb_set_error("device-side bound exceeded"); return dst; }
    *n_left = n;
    if (n > cap) { orb_set_error("stereo output capacity too small"); return ORB_ERR_CAPACITY; }
    if (n > 0) {
        ORB_CUDA(cudaMemcpyAsync(u_right, l->d_u_right + (size_t)image * l->plan.kp_cap, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, l->stream));
        ORB_CUDA(cudaMemcpyAsync(depth, l->d_depth + (size_t)image * l->plan.kp_cap, (size_t)n * sizeof(float), cudaMemcpyDeviceToHost, l->stream));
        ORB_CUDA(cudaStreamSynchronize(l->stream));
    }
    return ORB_OK;
}

}  // extern "C"
