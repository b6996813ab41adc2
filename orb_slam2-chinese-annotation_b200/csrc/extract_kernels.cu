// sm_100a kernels of the ORB extractor (reference src/ORBextractor.cc).  Integer/byte work,
// HBM- and ALU-bound; no tensor cores by design.
//
//   k_resize_level     cv::resize(INTER_LINEAR) 8u, untiled fallback for scale factors beyond the TMA box (pyramid_kernels.cu)
//   k_fast_cells       per-cell FAST-9/16 + NMS + iniTh->minTh retry                  :785-865
//   (k_quadtree: DistributeOctTree / DivideNode, quadtree_kernel.cu)
//   k_orient_describe  IC_Angle + fastAtan2 + rBRIEF + final keypoint assembly        :77-164, 1141-1167
#include "orb_internal.h"
#include "sincosf_glibc.h"

// LDGSTS: global -> shared copies that bypass the register file (4- or 16-byte, both sides aligned)
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

__constant__ __align__(16) signed char c_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};

// ------------------------------------------------------------------------------------------
// Ingest: caller images (any row stride, any alignment) -> the pitched, 16-byte-aligned level-0 slots.
// One thread assembles one aligned 16-byte destination chunk from the two aligned 16-byte source blocks that cover it
// (word select + funnel shift): two LDG.128 and one STG.128 per thread, (row, chunk) flattened over the image so that no
// thread idles on a row tail.  Source reads stay inside aligned 16-byte blocks that hold at least one pixel of the row
// (always inside a cudaMalloc'ed allocation); destination bytes past the row's last pixel are written as 0.
__global__ void __launch_bounds__(256) k_ingest(const uint8_t* __restrict__ src, uint8_t* __restrict__ dst, const int w, const int h,
                                                const int stride, const unsigned long long img_stride, const int dst_pitch,
                                                const unsigned long long dst_img_stride, const int nch, const unsigned magic) {
    ORB_PDL_SYNC();
    const unsigned idx = blockIdx.x * 256u + threadIdx.x;
    const int y = nch == 1 ? (int)idx : (int)__umulhi(idx, magic);   // idx / nch (idx < 2^20, nch <= 256)
    if (y >= h) return;
    const int j = (int)idx - y * nch, b = blockIdx.y;
    const int rem = w - 16 * j;                       // pixels of this chunk (>= 1)
    const unsigned long long a = (unsigned long long)src + (unsigned long long)b * img_stride + (unsigned long long)y * stride + 16ull * j;
    const uint4* q = reinterpret_cast<const uint4*>(a & ~15ull);
    const unsigned mis = (unsigned)(a & 15ull);
    const uint4 A = __ldg(q);
    uint4 B = make_uint4(0u, 0u, 0u, 0u);
    if (mis + (unsigned)min(rem, 16) > 16u) B = __ldg(q + 1);
    // words mis/4 .. mis/4 + 4 of (A, B), then a byte funnel by mis % 4
    unsigned w0 = A.x, w1 = A.y, w2 = A.z, w3 = A.w, w4 = B.x, w5 = B.y, w6 = B.z, w7 = B.w;
    if (mis & 8u) { w0 = w2; w1 = w3; w2 = w4; w3 = w5; w4 = w6; w5 = w7; }
    if (mis & 4u) { w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; }
    const unsigned sh = (mis & 3u) * 8u;
    uint4 v = make_uint4(__funnelshift_r(w0, w1, sh), __funnelshift_r(w1, w2, sh), __funnelshift_r(w2, w3, sh), __funnelshift_r(w3, w4, sh));
    if (rem < 16) {
        const unsigned keep = 0xffffffffu >> (8 * (4 - (rem & 3)));   // rem & 3 == 0 is handled by the word tests
        if (rem < 4) { v.x &= keep; v.y = v.z = v.w = 0u; }
        else if (rem < 8) { if (rem & 3) v.y &= keep; else v.y = 0u; v.z = v.w = 0u; }
        else if (rem < 12) { if (rem & 3) v.z &= keep; else v.z = 0u; v.w = 0u; }
        else { if (rem & 3) v.w &= keep; else v.w = 0u; }
    }
    *reinterpret_cast<uint4*>(dst + (size_t)b * dst_img_stride + (size_t)y * dst_pitch + 16 * j) = v;
}
void launch_ingest(cudaStream_t s, const uint8_t* src, uint8_t* dst, int w, int h, int stride, size_t img_stride, int dst_pitch, size_t dst_img_stride, int batch) {
    const int nch = (w + 15) / 16;
    const unsigned magic = (unsigned)((1ull << 32) / (unsigned)nch) + 1u;
    dim3 grid((unsigned)(((size_t)nch * h + 255) / 256), batch);
    // plain launch: what precedes it in the stream is a copy or a cross-stream event wait, not a kernel of the chain
    orb_launch(false, k_ingest, grid, dim3(256), 0, s, src, dst, w, h, stride, (unsigned long long)img_stride, dst_pitch, (unsigned long long)dst_img_stride, nch, magic);
}

// ------------------------------------------------------------------------------------------
// Pyramid, untiled fallback (scale factors whose source footprint exceeds the TMA box of k_resize_tma):
// level l <- level l-1, OpenCV 8u INTER_LINEAR (11-bit coefficients).  One thread
// produces 4 horizontally adjacent pixels of one row and stores them as one 32-bit word.  The
// two source rows are read as aligned words; a (pixel, right neighbour) byte pair is pulled out
// with one PRMT and weighted with one DP2A (the two 16-bit horizontal coefficients of a pixel are
// stored as one 32-bit word).
__device__ __forceinline__ unsigned pick_pair(unsigned w0, unsigned w1, unsigned w2, int idx) {
    // bytes idx, idx+1 of the 12-byte window (w0,w1,w2), idx in [0, 7]
    const unsigned lo = idx < 4 ? w0 : w1, hi = idx < 4 ? w1 : w2;
    const unsigned j = idx & 3;
    return __byte_perm(lo, hi, j | ((j + 1) << 4));
}

#define RESIZE_ROWS 4   // destination rows per thread: the horizontal tables and byte selectors are per column
__global__ void __launch_bounds__(256) k_resize_level(uint8_t* __restrict__ pyr, const LevelGeom src, const LevelGeom dst) {
    const int x4 = (blockIdx.x * 64 + threadIdx.x) * 4;
    const int y0 = (blockIdx.y * 4 + threadIdx.y) * RESIZE_ROWS;
    if (x4 >= dst.w || y0 >= dst.h) return;
    const uint8_t* S = pyr + src.off + (size_t)blockIdx.z * src.img_stride;
    uint8_t* D = pyr + dst.off + (size_t)blockIdx.z * dst.img_stride;
    // tables are padded to a multiple of 4 entries and 16-byte aligned
    const int4 xo = reinterpret_cast<const int4*>(dst.xofs)[x4 >> 2];
    const uint4 al = reinterpret_cast<const uint4*>(dst.alpha)[x4 >> 2];
    const int base = xo.x & ~3;
    const int sxs[4] = {xo.x, xo.y, xo.z, xo.w};
    const unsigned als[4] = {al.x, al.y, al.z, al.w};
    const bool window = xo.w - base <= 7;
    unsigned sel[4];
    bool hi4[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int idx = sxs[i] - base;
        const unsigned j = idx & 3;
        sel[i] = j | ((j + 1) << 4);
        hi4[i] = idx >= 4;
    }
#pragma unroll
    for (int r = 0; r < RESIZE_ROWS; ++r) {
        const int y = y0 + r;
        if (y >= dst.h) break;
        int sy0 = dst.yofs[y], sy1 = sy0 + 1;
        sy0 = min(max(sy0, 0), src.h - 1);
        sy1 = min(max(sy1, 0), src.h - 1);
        const unsigned bw = reinterpret_cast<const unsigned*>(dst.beta)[y];
        const int b0 = (int)(bw & 0xffffu), b1 = (int)(bw >> 16);
        const uint8_t* r0 = S + (size_t)sy0 * src.pitch;
        const uint8_t* r1 = S + (size_t)sy1 * src.pitch;
        uint32_t out = 0;
        if (window) {
            const unsigned* p0 = reinterpret_cast<const unsigned*>(r0 + base);
            const unsigned* p1 = reinterpret_cast<const unsigned*>(r1 + base);
            const unsigned u0 = p0[0], u1 = p0[1], u2 = p0[2], v0 = p1[0], v1 = p1[1], v2 = p1[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const unsigned pu = __byte_perm(hi4[i] ? u1 : u0, hi4[i] ? u2 : u1, sel[i]);
                const unsigned pv = __byte_perm(hi4[i] ? v1 : v0, hi4[i] ? v2 : v1, sel[i]);
                const int h0 = (int)__dp2a_lo(als[i], pu, 0u);
                const int h1 = (int)__dp2a_lo(als[i], pv, 0u);
                // weights sum to 2048 * 2048: the result is a convex combination of bytes, no clamp needed
                const int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                out |= (uint32_t)v << (8 * i);
            }
        } else {  // generic path (scale factors whose 4-pixel footprint exceeds the 12-byte window)
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int sx = sxs[i], sx1 = min(sx + 1, src.w - 1);
                const int a0 = (int)(als[i] & 0xffffu), a1 = (int)(als[i] >> 16);
                const int h0 = r0[sx] * a0 + r0[sx1] * a1;
                const int h1 = r1[sx] * a0 + r1[sx1] * a1;
                int v = (((b0 * (h0 >> 4)) >> 16) + ((b1 * (h1 >> 4)) >> 16) + 2) >> 2;
                v = min(max(v, 0), 255);
                out |= (uint32_t)v << (8 * i);
            }
        }
        // rows are pitch-padded to a multiple of 64 bytes, so the full word is always in bounds
        *reinterpret_cast<uint32_t*>(D + (size_t)y * dst.pitch + x4) = out;
    }
}

// ------------------------------------------------------------------------------------------
// FAST-9/16.  cv::FAST(cell, th, nms=true) on a cell view is equivalent to: threshold-free score
// s = max(bright, dark) - 1 per interior pixel; corner at th <=> s >= th; a corner survives NMS
// iff s is strictly greater than the raw scores of its 8 neighbours inside the cell interior
// (outside counts as 0).  The set of strict local maxima does not depend on th, so the
// iniTh -> minTh retry (reference :843-850) only re-filters the same list: a cell keeps its maxima
// with s >= iniTh, or, when it has none, those with s >= minTh.
//
// One WARP per cell, FC_CELLS consecutive cells per warp, no CTA-wide barrier anywhere: every warp owns a private
// shared-memory region (tile | score map with a zero frame | list) and every list stays in row-major order, so the
// final write needs no sort, no atomics and no prefix over warps, and the per-cell threshold choice is a warp vote:
//   (1) the cell view (w_cell + 6) x (h_cell + 6) arrives by TMA (one box per cell, requested one cell ahead);
//   (2) byte-SIMD necessary test at minTh, 4 pixels per word, every lane on its own contiguous run of words: test + count,
//       one warp scan, ordered write (no ballot, no POPC per word);
//   (3) exact score of the survivors, two pixels per lane on packed 16-bit lanes (VIMNMX3.S16x2), into the score
//       map; corners compacted in place;
//   (4) strict local maxima (VIMNMX3 over the 8 neighbours), compacted in place;
//   (5) ordered write of the maxima.
// (2) - (4) run at iniTh first and are repeated at minTh only for a cell that came back empty, like the reference's retry.
// The tile pitch TP and score pitch SP are compile-time so that every ring / neighbour offset is an immediate.
#define FC_WARPS 4     // warps per CTA (independent of each other)
#define FC_CELLS 4     // consecutive cells walked by one warp when the batch fills the GPU (1 for small batches: latency)
#define FC_TP 80       // tile pitch: view width (<= 59 + 6) + alignment shift (<= 15); 20 words: consecutive rows land in different banks
                       // (a 64-byte pitch was measured for cells <= 43 wide: 28 instead of 24 resident warps, but rows alias every
                       // second row and the ring gathers conflict: 1.93 -> 2.91 ms)

__device__ __forceinline__ unsigned gt_bytes(unsigned ad, unsigned k) {  // bit 7 of byte j set <=> ad.j > 127 - k.j  (k.j <= 127)
    return ad | ((ad & 0x7f7f7f7fu) + k);
}

// FAST score of two pixels at once.  e[k] = ring[k] (raw bytes) on packed 16-bit lanes, c = the two centres.
// min / max commute with the subtraction of the centre, so the arc network runs on the raw ring:
// dark = (max over the 16 arcs of min over 9 contiguous e) - c; bright = c - (min over arcs of max over 9).
// The network runs on HMNMX2: a 16-bit lane holding 0..255 IS a (subnormal) half with the same order, min / max return one of their
// inputs bit for bit, and HMNMX2 issues at 128 lanes per clock per SM against 64 for VIMNMX3.S16x2 (tools/peaks_int.cu).  With
// two-input operations the cheapest network is the pairwise one: for even k, M8[k] = min e[k+1 .. k+8] from pair and quad minima,
// and the better of the two 9-arcs that contain it is min(M8[k], max(e[k], e[k+9])): 47 operations per polarity.
__device__ __forceinline__ unsigned hmin2u(unsigned a, unsigned b) { unsigned r; asm("min.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned hmax2u(unsigned a, unsigned b) { unsigned r; asm("max.f16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }
__device__ __forceinline__ unsigned fast_score_x2(const unsigned (&e)[16], const unsigned c) {
    unsigned plo[8], phi[8];      // pairs (e[j], e[j+1]), j = 1, 3, .., 15
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        plo[i] = hmin2u(e[(2 * i + 1) & 15], e[(2 * i + 2) & 15]);
        phi[i] = hmax2u(e[(2 * i + 1) & 15], e[(2 * i + 2) & 15]);
    }
    unsigned qlo[8], qhi[8];      // quads e[j .. j+3], j = 1, 3, .., 15
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        qlo[i] = hmin2u(plo[i], plo[(i + 1) & 7]);
        qhi[i] = hmax2u(phi[i], phi[(i + 1) & 7]);
    }
    unsigned dk = 0u, br = 0u;
#pragma unroll
    for (int i = 0; i < 8; ++i) {   // k = 2 i: M8 = e[k+1 .. k+8] = quads at k+1 and k+5; arcs [k, k+8] and [k+1, k+9]
        const unsigned m8lo = hmin2u(qlo[i], qlo[(i + 2) & 7]), m8hi = hmax2u(qhi[i], qhi[(i + 2) & 7]);
        const unsigned ends_hi = hmax2u(e[2 * i], e[(2 * i + 9) & 15]), ends_lo = hmin2u(e[2 * i], e[(2 * i + 9) & 15]);
        const unsigned tlo = hmin2u(m8lo, ends_hi);   // the larger of the two arc minima
        const unsigned thi = hmax2u(m8hi, ends_lo);   // the smaller of the two arc maxima
        dk = i == 0 ? tlo : hmax2u(dk, tlo);
        br = i == 0 ? thi : hmin2u(br, thi);
    }
    // max(dark, bright) - 1 on both lanes
    return __vsub2(__vmaxs2(__vsub2(dk, c), __vsub2(c, br)), 0x00010001u);
}

// g / n == (g * c_div_magic[n]) >> 20 for g < 2^20 / n (words of a cell view: n <= 17, g < 17 * 59)
__constant__ unsigned c_div_magic[20] = {0u, (1u << 20) / 1 + 1, (1u << 20) / 2 + 1, (1u << 20) / 3 + 1, (1u << 20) / 4 + 1, (1u << 20) / 5 + 1,
    (1u << 20) / 6 + 1, (1u << 20) / 7 + 1, (1u << 20) / 8 + 1, (1u << 20) / 9 + 1, (1u << 20) / 10 + 1, (1u << 20) / 11 + 1, (1u << 20) / 12 + 1,
    (1u << 20) / 13 + 1, (1u << 20) / 14 + 1, (1u << 20) / 15 + 1, (1u << 20) / 16 + 1, (1u << 20) / 17 + 1, (1u << 20) / 18 + 1, (1u << 20) / 19 + 1};

// Dynamic shared memory, per warp (warp_bytes apart): tile (rows x FC_TP bytes, 16-byte aligned like the level) |
// score map ((ih + 2) x SP bytes, zero frame) at score_off | list (u16: y << 8 | x) at list_off
struct FcCell { int l, c, ci, cj, iniX, iniY, tw, th; bool valid; };
template <int SP>
__global__ void __launch_bounds__(FC_WARPS * 32, 6)   // shared memory allows six CTAs per SM: keep the registers within that
k_fast_cells(const __grid_constant__ TmaMaps tm, int* __restrict__ cell_cnt, uint32_t* __restrict__ stage, const Plan P,
             const int warp_bytes, const int score_off, const int list_off, const int cells_per_warp, const int box_h, const int b0) {
    ORB_PDL_SYNC();
    constexpr int TP = FC_TP, TPW = TP / 4;
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t s_bar[FC_WARPS];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt = (1u << lane) - 1, full = 0xffffffffu;
    uint8_t* const wt = smem + warp * warp_bytes;
    uint8_t* const sc = wt + score_off;
    uint16_t* const seg = reinterpret_cast<uint16_t*>(wt + list_off);
    const uint32_t* const tw32 = reinterpret_cast<const uint32_t*>(wt);
    const int b = blockIdx.y;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"((unsigned)__cvta_generic_to_shared(&s_bar[warp])) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // make the init visible to the async proxy
    }
    __syncwarp();

    int cg = (blockIdx.x * FC_WARPS + warp) * cells_per_warp;
    const int cg_end = min(cg + cells_per_warp, P.total_cells);
    int lv = 0;
    // cell view [iniX, maxX) x [iniY, maxY) (reference :819-839)
    auto geom = [&](int cgi) {
        FcCell g;
        while (lv + 1 < P.nlevels && cgi >= P.lv[lv + 1].cell_off) ++lv;
        const LevelGeom& G = P.lv[lv];
        g.l = lv;
        g.c = cgi - G.cell_off;
        g.ci = (int)__umulhi((unsigned)g.c, G.ncols_magic); g.cj = g.c - g.ci * G.n_cols;
        g.iniY = ORB_WIN + g.ci * G.h_cell; g.iniX = ORB_WIN + g.cj * G.w_cell;
        g.tw = min(g.iniX + G.w_cell + 6, G.max_bx) - g.iniX; g.th = min(g.iniY + G.h_cell + 6, G.max_by) - g.iniY;
        g.valid = !(g.iniY >= G.max_by - 3 || g.iniX >= G.max_bx - 6 || g.tw - 6 <= 0 || g.th - 6 <= 0);
        return g;
    };
    // (1) the view (80 x box_h bytes from the 16-byte boundary left of it: the tile rows keep the level's alignment) arrives by TMA;
    // the box of the NEXT cell is requested as soon as the last reader of the tile (the score phase) is done, so that it flies
    // under the neighbourhood test and the output phase
    auto request = [&](const FcCell& g) {
        if (lane == 0) {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
            asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(TP * box_h) : "memory");
            asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                         ::"r"((unsigned)__cvta_generic_to_shared(wt)), "l"(&tm.lv[g.l]), "r"(bar), "r"(g.iniX & ~15), "r"(g.iniY), "r"(b0 + b) : "memory");
        }
    };
    if (cg >= cg_end) return;
    FcCell g = geom(cg);
    if (g.valid) request(g);
    unsigned phase = 0;
#pragma unroll 1
    for (; cg < cg_end; ++cg) {
        FcCell gn;
        gn.valid = false;
        int* const out_cnt = cell_cnt + (size_t)b * P.total_cells + cg;
        if (!g.valid) {
            if (lane == 0) *out_cnt = 0;
            if (cg + 1 < cg_end) { gn = geom(cg + 1); if (gn.valid) request(gn); }
            g = gn;
            continue;
        }
        const LevelGeom& G = P.lv[g.l];
        const int c = g.c, ci = g.ci, cj = g.cj, iniX = g.iniX;
        const int iw = g.tw - 6, ih = g.th - 6;  // interior = tested pixels
        const int shift = iniX & 15;
        {
            const unsigned bar = (unsigned)__cvta_generic_to_shared(&s_bar[warp]);
            unsigned ok;
            do {
                asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                             : "=r"(ok) : "r"(bar), "r"(phase) : "memory");
            } while (!ok);
            phase ^= 1u;
        }
        const uint8_t* tile = wt + shift;
        // interior words of a row: word wx holds interior pixels xi = 4 * wx - shift - 3 + k, k = 0..3
        const int w0 = (shift + 3) >> 2, nwi = ((shift + 3 + iw - 1) >> 2) - w0 + 1;

        // The reference runs FAST at iniThFAST and, when a cell comes back empty, again at minThFAST (:843-850).  Same here: the first
        // attempt tests, scores and suppresses at iniTh - far fewer pixels pass the necessary test and reach the arc network than at
        // minTh - and only a cell without a single maximum is done again at minTh (the tile is still in shared memory).
        int n_m = 0;
#pragma unroll 1
        for (int attempt = 0; attempt < 2; ++attempt) {
        const int t = attempt == 0 ? P.ini_th : P.min_th;
        const unsigned kk = (unsigned)(127 - min(t, 126)) * 0x01010101u;
        // (2) necessary test: both ring pixels of the vertical pair (1, 9) and of the horizontal pair
        // (5, 13) within t of the centre => no 9-arc.  Two passes over lane-contiguous runs of word PAIRS (8 pixels, aligned to 8 bytes:
        // three LDS.64 + two LDS.32 per step; lane order == row-major order): pass 1 tests and counts, one warp scan turns the counts
        // into offsets, pass 2 writes the survivors in order.  No ballot and no POPC per word; the masks wait in the (not yet zeroed)
        // score-map region; which bytes of a tile word are interior pixels comes from a 20-entry table at the head of the list.
        int n_a;
        {
            const int xfirst = 4 * w0 - shift - 3;                       // <= 0: interior x of byte 0 of the first word
            const unsigned mfirst = 0x80808080u << (8 * -xfirst);
            const int over = xfirst + 4 * nwi - iw;                      // bytes of the last word beyond the interior (0..3)
            const unsigned mlast = 0x80808080u >> (8 * over);
            const int wl = w0 + nwi - 1;                                 // last interior word of a tile row
            uint32_t* const wmask = reinterpret_cast<uint32_t*>(seg);    // the list is written by pass 2 only
            if (lane < TPW) wmask[lane] = (lane < w0 || lane > wl) ? 0u : ((lane == w0 ? mfirst : 0x80808080u) & (lane == wl ? mlast : 0x80808080u));
            __syncwarp();
            const int p0 = w0 >> 1, npairs = (wl >> 1) - p0 + 1;
            const int np_tot = npairs * ih;
            const int R = (np_tot + 31) >> 5;                            // pairs per lane
            const int f0 = lane * R, nrun = min(R, np_tot - f0);
            uint32_t* const mk = reinterpret_cast<uint32_t*>(sc) + lane;
            unsigned c4 = 0;                                             // survivors of this lane, one counter per byte position
            int nw = 0;                                                  // words of this lane's run that hold a survivor
            if (nrun > 0) {
                // (Starting lane i's run at offset i / 8, cyclically, makes these loads conflict-free for cells up to 32 rows - lanes i
                // and i + 8 otherwise sit 8 rows apart, i.e. in the same banks - but measured no faster: 1.742 vs 1.758 ms.)
                const int y0 = (int)(((unsigned)f0 * c_div_magic[npairs]) >> 20);
                int pi = f0 - y0 * npairs;
                const uint32_t* cw = tw32 + (y0 + 3) * TPW + 2 * (p0 + pi);
                int ib = (y0 << 8) + 4 * (2 * (p0 + pi) - w0);           // y << 8 | 4 * word (bits 0..6, 8..13; the mask uses 7, 15, 23, 31)
#pragma unroll 1
                for (int st = 0; st < nrun; ++st) {
                    const uint2 C = *reinterpret_cast<const uint2*>(cw), N = *reinterpret_cast<const uint2*>(cw - 3 * TPW),
                                S = *reinterpret_cast<const uint2*>(cw + 3 * TPW);
                    const unsigned L = cw[-1], Rr = cw[2];
                    const uint2 wm = *reinterpret_cast<const uint2*>(wmask + 2 * (p0 + pi));
                    const unsigned E0 = __byte_perm(C.x, C.y, 0x6543), W0 = __byte_perm(L, C.x, 0x4321);   // x + 3, x - 3
                    const unsigned E1 = __byte_perm(C.y, Rr, 0x6543), W1 = __byte_perm(C.x, C.y, 0x4321);
                    const unsigned m0 = (gt_bytes(__vabsdiffu4(N.x, C.x), kk) | gt_bytes(__vabsdiffu4(S.x, C.x), kk)) &
                                        (gt_bytes(__vabsdiffu4(E0, C.x), kk) | gt_bytes(__vabsdiffu4(W0, C.x), kk)) & wm.x;
                    const unsigned m1 = (gt_bytes(__vabsdiffu4(N.y, C.y), kk) | gt_bytes(__vabsdiffu4(S.y, C.y), kk)) &
                                        (gt_bytes(__vabsdiffu4(E1, C.y), kk) | gt_bytes(__vabsdiffu4(W1, C.y), kk)) & wm.y;
                    // words without a survivor (most of them at iniTh) are not revisited
                    if (m0) { mk[nw * 32] = m0 | (unsigned)ib; ++nw; }
                    if (m1) { mk[nw * 32] = m1 | (unsigned)(ib + 4); ++nw; }
                    c4 += (m0 >> 7) + (m1 >> 7);
                    ++pi; cw += 2; ib += 8;
                    if (pi == npairs) { pi = 0; cw += TPW - 2 * npairs; ib += 256 - 8 * npairs; }
                }
            }
            const int cnt = (int)((c4 * 0x01010101u) >> 24);             // <= 8 R < 256
            int inc = cnt;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int v = __shfl_up_sync(full, inc, o); if (lane >= o) inc += v; }
            n_a = __shfl_sync(full, inc, 31);
            uint16_t* o = seg + (inc - cnt);
#pragma unroll 1
            for (int st = 0; st < nw; ++st) {
                const unsigned v = mk[st * 32];
                const unsigned i0 = (v & 0x3f7fu) + (unsigned)xfirst;
                if (v & 0x80u) *o++ = (uint16_t)i0;
                if (v & 0x8000u) *o++ = (uint16_t)(i0 + 1);
                if (v & 0x800000u) *o++ = (uint16_t)(i0 + 2);
                if (v & 0x80000000u) *o++ = (uint16_t)(i0 + 3);
            }
            __syncwarp();
            uint4* sc4 = reinterpret_cast<uint4*>(sc);
            for (int i = lane; i < (ih + 2) * (SP / 16); i += 32) sc4[i] = make_uint4(0u, 0u, 0u, 0u);
        }
        __syncwarp();

        // (3) exact score, two survivors per lane; corners (s >= minTh) compacted in place.  Lane i takes survivors k0 + i and
        // k0 + 32 + i (not 2 i and 2 i + 1): the 32 pixels behind one byte gather then span about three tile rows instead of six, i.e.
        // fewer words than banks, and the gathers (34 per pixel pair, the LSU bound of this phase) stay nearly conflict-free.
        int n_c = 0;
#pragma unroll 1
        for (int k0 = 0; k0 < n_a; k0 += 64) {
            const int ka = k0 + lane, kb = ka + 32;
            const bool va = ka < n_a, vb = kb < n_a;
            const int ia = va ? (int)seg[ka] : 0, ib = vb ? (int)seg[kb] : ia;
            const int ya = ia >> 8, xa = ia & 0xff, yb = ib >> 8, xb = ib & 0xff;
            const uint8_t* ca = tile + (ya + 3) * TP + xa + 3;
            const uint8_t* cb = tile + (yb + 3) * TP + xb + 3;
            const unsigned cen = (unsigned)ca[0] | ((unsigned)cb[0] << 16);
            unsigned e[16];
#define RING(k_, off_) e[k_] = (unsigned)ca[off_] | ((unsigned)cb[off_] << 16)
            RING(0, 3 * TP);       RING(1, 3 * TP + 1);   RING(2, 2 * TP + 2);    RING(3, TP + 3);
            RING(4, 3);            RING(5, -TP + 3);      RING(6, -2 * TP + 2);   RING(7, -3 * TP + 1);
            RING(8, -3 * TP);      RING(9, -3 * TP - 1);  RING(10, -2 * TP - 2);  RING(11, -TP - 3);
            RING(12, -3);          RING(13, TP - 3);      RING(14, 2 * TP - 2);   RING(15, 3 * TP - 1);
#undef RING
            const unsigned s2 = fast_score_x2(e, cen);
            const int sa = (int)(short)(s2 & 0xffffu), sb = (int)(short)(s2 >> 16);
            const bool fa = va && sa >= t, fb = vb && sb >= t;
            if (fa) sc[(ya + 1) * SP + xa + 1] = (uint8_t)sa;
            if (fb) sc[(yb + 1) * SP + xb + 1] = (uint8_t)sb;
            const unsigned ba = __ballot_sync(full, fa), bb = __ballot_sync(full, fb);  // also orders the reads above before the writes below
            if (fa) seg[n_c + __popc(ba & lt)] = (uint16_t)ia;                    // list order: the 32 first survivors, then the 32 next
            if (fb) seg[n_c + __popc(ba) + __popc(bb & lt)] = (uint16_t)ib;
            n_c += __popc(ba) + __popc(bb);
            __syncwarp();
        }

        __syncwarp();   // every lane is done reading the tile
        if (attempt == 1 && cg + 1 < cg_end) { gn = geom(cg + 1); if (gn.valid) request(gn); }   // no further attempt: the tile is free

        // (4) strict local maxima over the 8 neighbours (the zero frame stands for "outside the cell"), compacted in place
        n_m = 0;
#pragma unroll 1
        for (int k0 = 0; k0 < n_c; k0 += 32) {
            const int k = k0 + lane;
            bool mx = false;
            int i = 0;
            if (k < n_c) {
                i = seg[k];
                const uint8_t* q = sc + ((i >> 8) + 1) * SP + (i & 0xff) + 1;
                const unsigned r0 = __vimax3_u32(q[-SP - 1], q[-SP], q[-SP + 1]), r1 = __vimax3_u32(q[-1], q[1], q[SP - 1]);
                mx = (unsigned)q[0] > __vimax3_u32(r0, r1, max((unsigned)q[SP], (unsigned)q[SP + 1]));
            }
            const unsigned bm = __ballot_sync(full, mx);   // also orders this iteration's reads before the writes
            if (mx) seg[n_m + __popc(bm & lt)] = (uint16_t)i;
            n_m += __popc(bm);
            __syncwarp();
        }
        if (n_m > 0 || P.ini_th == P.min_th) {
            if (attempt == 0 && cg + 1 < cg_end) { gn = geom(cg + 1); if (gn.valid) request(gn); }
            break;
        }
        }   // attempt

        // (5) ordered write of the maxima (all of them passed the threshold of the attempt that found them)
        if (lane == 0) *out_cnt = n_m;
        const int ky0 = 3 + ci * G.h_cell, kx0 = 3 + cj * G.w_cell;
        uint32_t* out0 = stage + (size_t)b * P.stage_per_image + G.stage_off + (size_t)c * G.cell_cap;
        for (int k = lane; k < n_m; k += 32) {
            const int i = seg[k], y = i >> 8, x = i & 0xff;
            const int s = sc[(y + 1) * SP + x + 1];
            out0[k] = ((uint32_t)s << 24) | ((uint32_t)(y + ky0) << 12) | (uint32_t)(x + kx0);
        }
        __syncwarp();   // the score map and the list are free for the next cell
        g = gn;
    }
}

// ------------------------------------------------------------------------------------------
// Orientation + descriptor + final keypoint record.  One warp per keypoint slot.
__device__ __forceinline__ float fast_atan2_deg(float y, float x) {
    // cv::fastAtan2, float32, every operation individually rounded
    const float k180pi = (float)(180.0 / 3.14159265358979323846);
    const float p1 = 0.9997878412794807f * k180pi, p3 = -0.3258083974640975f * k180pi;
    const float p5 = 0.1555786518463281f * k180pi, p7 = -0.04432655554792128f * k180pi;
    const float ax = fabsf(x), ay = fabsf(y);
    float a, c, c2;
    if (ax >= ay) {
        c = __fdiv_rn(ay, __fadd_rn(ax, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c);
    } else {
        c = __fdiv_rn(ax, __fadd_rn(ay, 2.2204460492503131e-16f));
        c2 = __fmul_rn(c, c);
        a = __fsub_rn(90.f, __fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(__fadd_rn(__fmul_rn(p7, c2), p5), c2), p3), c2), p1), c));
    }
    if (x < 0) a = __fsub_rn(180.f, a);
    if (y < 0) a = __fsub_rn(360.f, a);
    return a;
}

#define OD_WARPS 4         // warps per CTA (independent of each other after the prologue)
#define OD_KPW 8           // keypoints walked by one warp (2 for small batches): the patches of keypoint j + 1 are in flight while j is described
// Both patches of a keypoint arrive by TMA (one cp.async.bulk.tensor per patch, issued by lane 0, SASS UTMALDG): the innermost box
// coordinate of a u8 tensor must be a multiple of 16 bytes (tools/tma_align_test.cu: anything else is an illegal instruction), so
// the boxes are wider than the patches and start at the 16-byte boundary below the patch.
#define OD_BB_W 64         // blurred patch: 37 x 37 around the keypoint (rBRIEF reaches |13| rotated, < 18.4) + up to 15 bytes of alignment
#define OD_BB_H 37
#define OD_RB_W 48         // raw patch: 31 x 31 (IC_Angle disc) + up to 15 bytes of alignment
#define OD_RB_H 31
#define OD_BLUR_BYTES 2432 // 64 * 37 rounded up to the 128-byte destination alignment of TMA
#define OD_RAW_BYTES 1536  // 48 * 31 rounded up (exactly 384 words: 12 per lane)
#define OD_SLOT_BYTES OD_BLUR_BYTES   // a slot holds one box at a time: the raw boxes first (moments), then the blurred ones (rBRIEF)
// pattern copy in global memory: every lane reads the 32 bytes of ITS descriptor byte (16 points) once per warp
__device__ __align__(16) signed char g_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};

// Weights of the intensity-centroid moments for the raw box as it is staged: 31 rows x 12 aligned words, the patch starting at byte
// rsh = 0..15 of a row.  Entry [rsh][row * 12 + word] = (x: the four u = column - 15 as signed bytes, 0 outside the radius-15 disc
// (|u| <= umax[|v|]); y: v = row - 15 inside the disc, 0 outside).  m10 += dp4a(pixels, x); m01 += dp4a(pixels, y): exact integers,
// any summation order.  Filled by orb_upload_moment_table() when an extractor is created (umax comes from the reference's table
// code, :466-489).
__device__ uint2 g_moment_w[16 * 384];
cudaError_t orb_upload_moment_table(const int* umax) {
    std::vector<uint2> h(16 * 384);   // same contents from every caller (the disc is a constant of the algorithm)
    for (int rsh = 0; rsh < 16; ++rsh)
        for (int i = 0; i < 384; ++i) {
            const int r = i / 12, c = i % 12, v = r - ORB_HALF_PATCH;
            uint32_t wu = 0, wv = 0;
            for (int j = 0; j < 4 && r < OD_RB_H; ++j) {
                const int u = 4 * c + j - rsh - ORB_HALF_PATCH;
                if (u >= -ORB_HALF_PATCH && u <= ORB_HALF_PATCH && (u < 0 ? -u : u) <= umax[v < 0 ? -v : v]) {
                    wu |= (uint32_t)(uint8_t)(signed char)u << (8 * j);
                    wv |= (uint32_t)(uint8_t)(signed char)v << (8 * j);
                }
            }
            h[rsh * 384 + i] = make_uint2(wu, wv);
        }
    return cudaMemcpyToSymbol(g_moment_w, h.data(), h.size() * sizeof(uint2));
}
__device__ __forceinline__ int dp4a_u8_s8(unsigned px, unsigned w, int acc) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(w), "r"(acc));
    return d;
}
// mbarrier / TMA PTX (same forms as pyramid_kernels.cu)
__device__ __forceinline__ unsigned od_smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void od_mbar_init(uint64_t* bar) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(od_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void od_mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(od_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void od_mbar_wait(uint64_t* bar, unsigned phase) {
    unsigned ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(od_smem_u32(bar)), "r"(phase) : "memory");
    } while (!ok);
}
__device__ __forceinline__ unsigned od_lds_u8(unsigned addr) {
    unsigned v;
    asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}
// packed single precision (Blackwell FMUL2 / FFMA2 / FADD2): two IEEE operations per instruction, each rounded like the scalar one
__device__ __forceinline__ unsigned long long f2_pack(float lo, float hi) { unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ unsigned long long f2_mul(unsigned long long a, unsigned long long b) { unsigned long long r; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ unsigned long long f2_fma(unsigned long long a, unsigned long long b, unsigned long long c) { unsigned long long r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ unsigned long long f2_add(unsigned long long a, unsigned long long b) { unsigned long long r; asm("add.rn.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ void od_tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(od_smem_u32(smem_dst)), "l"(map), "r"(od_smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}

// One warp walks OD_KPW consecutive keypoints of the image's final (level-concatenated) order, double-buffered: lane 0 issues the
// TMA box of keypoint j + 1, then the warp waits for keypoint j's mbarrier and computes on shared memory only.  The 16 pattern
// points of a lane's descriptor byte live in registers as floats.
__global__ void __launch_bounds__(OD_WARPS * 32, 7)
k_orient_describe(const __grid_constant__ TmaMaps tm_raw, const __grid_constant__ TmaMaps tm_blur, const uint32_t* __restrict__ lvl_kp,
                  const int* __restrict__ lvl_cnt, orb_keypoint* __restrict__ kps, uint8_t* __restrict__ desc,
                  int* __restrict__ counts, const Plan P, const int b0, const int kpw) {
    ORB_PDL_SYNC();
    extern __shared__ __align__(128) uint8_t od_smem[];
    __shared__ uint64_t s_bar[OD_WARPS][2];
    __shared__ int s_pre[ORB_MAX_LEVELS + 1];
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        const int* cnt = lvl_cnt + b * ORB_MAX_LEVELS;
        int t = 0;
        for (int l = 0; l < P.nlevels; ++l) { s_pre[l] = t; t += cnt[l]; }
        for (int l = P.nlevels; l <= ORB_MAX_LEVELS; ++l) s_pre[l] = t;
        if (blockIdx.x == 0) counts[b] = t;
    }
    if (lane == 0) {
        od_mbar_init(&s_bar[warp][0]); od_mbar_init(&s_bar[warp][1]);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // make the init visible to the async proxy
    }
    __syncthreads();
    const int o0 = (blockIdx.x * OD_WARPS + warp) * kpw;
    const int n = min(kpw, s_pre[ORB_MAX_LEVELS] - o0);
    if (n <= 0) return;

    // lane j < n looks up keypoint o0 + j: level, quadtree slot, packed key
    uint32_t my_key = 0;
    int my_l = 0;
    if (lane < n) {
        const int o = o0 + lane;
        while (my_l + 1 < P.nlevels && o >= s_pre[my_l + 1]) ++my_l;
        my_key = lvl_kp[(size_t)b * P.kp_cap + P.lv[my_l].kp_off + (o - s_pre[my_l])];
    }
    // the 16 pattern points of this lane's descriptor byte, as packed pairs (the two points of comparison k): x's and y's
    unsigned long long ptx2[8], pty2[8];
    {
        const uint4 q0 = reinterpret_cast<const uint4*>(g_pattern)[2 * lane], q1 = reinterpret_cast<const uint4*>(g_pattern)[2 * lane + 1];
        const unsigned w[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
#pragma unroll
        for (int i = 0; i < 8; ++i) {   // word i = (x0, y0, x1, y1) of comparison i
            ptx2[i] = f2_pack((float)(signed char)(w[i] & 0xff), (float)(signed char)((w[i] >> 16) & 0xff));
            pty2[i] = f2_pack((float)(signed char)((w[i] >> 8) & 0xff), (float)(signed char)(w[i] >> 24));
        }
    }
    uint8_t* const wbuf = od_smem + warp * (2 * OD_SLOT_BYTES);
    const int img = b0 + b;

    // Three phases per warp, so that the orientation arithmetic (fastAtan2, sincosf: warp-uniform per keypoint, ~85 instructions) is
    // done ONCE for the warp's keypoints, lane j working on keypoint j, instead of once per keypoint by 32 identical lanes:
    //   1. moments: the raw boxes of the keypoints stream through the two slots (TMA, one keypoint ahead); lane j keeps (m10, m01) of
    //      keypoint j;
    //   2. lane j: angle, sine, cosine and the 28-byte keypoint record of keypoint j;
    //   3. rBRIEF: the blurred boxes stream through the same two slots; sine / cosine of keypoint j come by shuffle.
    auto prefetch = [&](int j, bool blurred) {
        const uint32_t key = __shfl_sync(0xffffffffu, my_key, j);
        const int l = __shfl_sync(0xffffffffu, my_l, j);
        if (lane == 0) {
            const int px = ORB_KEY_X(key) + ORB_WIN, py = ORB_KEY_Y(key) + ORB_WIN;
            uint8_t* buf = wbuf + (j & 1) * OD_SLOT_BYTES;
            uint64_t* bar = &s_bar[warp][j & 1];
            if (blurred) {
                od_mbar_expect_tx(bar, OD_BB_W * OD_BB_H);
                od_tma_load_3d(buf, &tm_blur.lv[l], bar, (px - 18) & ~15, py - 18, img);
            } else {
                od_mbar_expect_tx(bar, OD_RB_W * OD_RB_H);
                od_tma_load_3d(buf, &tm_raw.lv[l], bar, (px - ORB_HALF_PATCH) & ~15, py - ORB_HALF_PATCH, img);
            }
        }
    };
    // ---- 1. IC_Angle moments (reference :77-113): integer sums over the radius-15 disc, two byte dot products per staged word
    int my_m10 = 0, my_m01 = 0;
    prefetch(0, false);
#pragma unroll 1
    for (int j = 0; j < n; ++j) {
        if (j + 1 < n) prefetch(j + 1, false);   // its slot was last read by keypoint j - 1, which ended with __syncwarp
        const uint32_t key = __shfl_sync(0xffffffffu, my_key, j);
        const int px = ORB_KEY_X(key) + ORB_WIN;
        const uint32_t* rw = reinterpret_cast<const uint32_t*>(wbuf + (j & 1) * OD_SLOT_BYTES) + lane;
        const uint2* wt = g_moment_w + ((px - ORB_HALF_PATCH) & 15) * 384 + lane;
        od_mbar_wait(&s_bar[warp][j & 1], (unsigned)((j >> 1) & 1));
        int m10 = 0, m01 = 0;
#pragma unroll
        for (int k = 0; k < 12; ++k) {
            const uint2 wq = __ldg(wt + k * 32);
            const unsigned v = rw[k * 32];
            m10 = dp4a_u8_s8(v, wq.x, m10);
            m01 = dp4a_u8_s8(v, wq.y, m01);
        }
        m10 = __reduce_add_sync(0xffffffffu, m10);   // REDUX.SUM: one instruction instead of five shuffle + add steps
        m01 = __reduce_add_sync(0xffffffffu, m01);
        if (lane == j) { my_m10 = m10; my_m01 = m01; }
        __syncwarp();   // every lane is done with this slot before lane 0 re-arms it
    }
    const int used0 = (n + 1) >> 1, used1 = n >> 1;   // completed phases of the two slots' barriers so far
    prefetch(0, true);

    // ---- 2. orientation and keypoint record, lane j for keypoint j
    float my_sin = 0.f, my_cos = 1.f;
    if (lane < n) {
        const float angle = fast_atan2_deg((float)my_m01, (float)my_m10);
        const float rad = __fmul_rn(angle, 0.017453292519943295f);  // factorPI = (float)(CV_PI/180.f)
        orb_sincosf(rad, &my_sin, &my_cos);
        const LevelGeom& G = P.lv[my_l];
        const int px = ORB_KEY_X(my_key) + ORB_WIN, py = ORB_KEY_Y(my_key) + ORB_WIN;
        float fx = (float)px, fy = (float)py;
        if (my_l != 0) { fx = __fmul_rn(fx, G.sf); fy = __fmul_rn(fy, G.sf); }
        uint32_t* rec = reinterpret_cast<uint32_t*>(kps + (size_t)b * P.kp_cap + o0 + lane);   // 28 bytes = cv::KeyPoint
        rec[0] = __float_as_uint(fx); rec[1] = __float_as_uint(fy); rec[2] = __float_as_uint(G.patch_size); rec[3] = __float_as_uint(angle);
        rec[4] = __float_as_uint((float)ORB_KEY_S(my_key)); rec[5] = (uint32_t)my_l; rec[6] = 0xffffffffu;   // class_id = -1
    }

    // ---- 3. rBRIEF (reference :119-164): lane = descriptor byte, 8 comparisons on the staged blurred patch.  cvRound of the rotated
    // coordinates as "add 1.5 * 2^23, read the mantissa" (round half to even like cvRound's cvtss2si / F2I.RN; |value| < 2^22)
    const unsigned long long magic2 = f2_pack(12582912.0f, 12582912.0f);
#pragma unroll 1
    for (int j = 0; j < n; ++j) {
        if (j + 1 < n) prefetch(j + 1, true);
        const uint32_t key = __shfl_sync(0xffffffffu, my_key, j);
        const float bsin = __shfl_sync(0xffffffffu, my_sin, j), a = __shfl_sync(0xffffffffu, my_cos, j);
        const int px = ORB_KEY_X(key) + ORB_WIN;
        const uint8_t* buf = wbuf + (j & 1) * OD_SLOT_BYTES;
        const unsigned long long sin2 = f2_pack(bsin, bsin), cos2 = f2_pack(a, a), nsin2 = f2_pack(-bsin, -bsin);
        // byte (row r, column q) of the patch sits at 64 * (18 + r) + 18 + dsh + q; the two mantissa offsets 0x4B400000 are folded
        // into the base (32-bit shared-window address arithmetic, modulo 2^32)
        const unsigned spc = od_smem_u32(buf) + 18u * OD_BB_W + 18u + (unsigned)((px - 18) & 15) - 0x4B400000u * (unsigned)(OD_BB_W + 1);
        od_mbar_wait(&s_bar[warp][j & 1], (unsigned)(((j >> 1) + ((j & 1) ? used1 : used0)) & 1));
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            // row = cvRound(x sin + y cos) = fma(x, sin, rn(y cos)), column = cvRound(x cos - y sin) = fma(x, cos, -rn(y sin)) (reference :129-131
            // as GCC contracts it); rn(y * -sin) == -rn(y * sin).  Both points of the comparison in one packed instruction each.
            const unsigned long long r2 = f2_add(f2_fma(ptx2[k], sin2, f2_mul(pty2[k], cos2)), magic2);
            const unsigned long long q2 = f2_add(f2_fma(ptx2[k], cos2, f2_mul(pty2[k], nsin2)), magic2);
            const unsigned r0 = (unsigned)r2, r1 = (unsigned)(r2 >> 32), q0 = (unsigned)q2, q1 = (unsigned)(q2 >> 32);
            const unsigned t0 = od_lds_u8(spc + r0 * OD_BB_W + q0), t1 = od_lds_u8(spc + r1 * OD_BB_W + q1);
            val |= (t0 < t1) << k;
        }
        desc[((size_t)b * P.kp_cap + o0 + j) * 32 + lane] = (uint8_t)val;
        __syncwarp();   // every lane is done with this slot before lane 0 re-arms it
    }
}

// ------------------------------------------------------------------------------------------
// host-callable launchers (used by extractor.cu)
void launch_resize_untiled(cudaStream_t s, uint8_t* pyr, const LevelGeom& src, const LevelGeom& dst, int batch) {
    dim3 block(64, 4), grid((dst.w + 255) / 256, (dst.h + 4 * RESIZE_ROWS - 1) / (4 * RESIZE_ROWS), batch);
    k_resize_level<<<grid, block, 0, s>>>(pyr, src, dst);
}
// Per-warp shared-memory layout of k_fast_cells for the largest cell of any level.
static bool fast_smem(const Plan& P, int* sp, int* warp_bytes, int* score_off, int* list_off) {
    int mh = 0, mw = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelGeom& G = P.lv[l];
        if (G.n_cols == 0) continue;
        mh = max(mh, G.h_cell); mw = max(mw, G.w_cell);
    }
    if (mw > 59 || mh > 59) return false;       // cannot happen: a cell is ceil(window / floor(window / 30)) <= 59 pixels
    *sp = mw + 2 <= 48 ? 48 : 64;               // score pitch: interior + zero frame
    *score_off = FC_TP * (mh + 6);
    // the score-map region first holds the masks of the necessary test: 32 lanes x (2 words per pair of a lane's run) u32
    const int np_max = (((mw + 6) / 4 + 1) / 2 + 1) * mh, scratch = 128 * 2 * ((np_max + 31) / 32);
    *list_off = *score_off + max(*sp * (mh + 2), scratch);
    *warp_bytes = (*list_off + 2 * mw * mh + 4 + 127) & ~127;   // the tile is a TMA destination: 128-byte aligned
    return true;
}
int fast_box_rows(const Plan& P) {   // rows of the TMA box of k_fast_cells: the tallest cell view of any level
    int mh = 0;
    for (int l = 0; l < P.nlevels; ++l) if (P.lv[l].n_cols > 0) mh = max(mh, P.lv[l].h_cell);
    return mh + 6;
}
cudaError_t launch_fast(cudaStream_t s, const TmaMaps& tm, int* cell_cnt, uint32_t* stage, const Plan& P, int b0, int batch) {
    int sp, wb, so, lo;
    if (!fast_smem(P, &sp, &wb, &so, &lo)) return cudaErrorInvalidConfiguration;
    const int tot = wb * FC_WARPS;
    auto kern = sp == 48 ? k_fast_cells<48> : k_fast_cells<64>;
    // always: static (mbarriers) + dynamic shared memory can cross the 48 KB default even when the dynamic part alone does not
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tot);
    if (e != cudaSuccess) return e;
    // a small batch cannot fill every SM's 24 resident warps with 4 cells per warp: one cell per warp then (per-frame latency)
    const int cpw = (long long)batch * P.total_cells >= 4LL * orb_sm_count() * 24 * FC_CELLS ? FC_CELLS : 1;
    dim3 grid((P.total_cells + FC_WARPS * cpw - 1) / (FC_WARPS * cpw), batch);
    e = orb_launch(orb_pdl(batch), kern, grid, dim3(FC_WARPS * 32), (size_t)tot, s, tm, cell_cnt, stage, P, wb, so, lo, cpw, fast_box_rows(P), b0);
    return e != cudaSuccess ? e : cudaGetLastError();   // a refused launch must fail loudly, not leave stale candidate lists behind
}
cudaError_t launch_orient_describe(cudaStream_t s, const TmaMaps& tm_raw, const TmaMaps& tm_blur, const uint32_t* lvl_kp, const int* lvl_cnt,
                                   orb_keypoint* kps, uint8_t* desc, int* counts, const Plan& P, int b0, int batch, bool after_event_wait) {
    const int smem = OD_WARPS * 2 * OD_SLOT_BYTES;
    cudaError_t e = cudaFuncSetAttribute(k_orient_describe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);   // per device
    if (e != cudaSuccess) return e;
    // small batches: fewer keypoints per warp, so that the frame's ~2000 keypoints spread over all SMs (per-frame latency)
    const int kpw = (long long)batch * P.kp_cap >= 2LL * orb_sm_count() * 28 * OD_KPW ? OD_KPW : 2;
    dim3 grid((P.kp_cap + OD_WARPS * kpw - 1) / (OD_WARPS * kpw), batch);
    // behind a cross-stream event wait (the side-stream blur) the launch is a plain one
    e = orb_launch(orb_pdl(batch) && !after_event_wait, k_orient_describe, grid, dim3(OD_WARPS * 32), (size_t)smem, s, tm_raw, tm_blur, lvl_kp, lvl_cnt, kps, desc, counts, P, b0, kpw);
    return e != cudaSuccess ? e : cudaGetLastError();
}

// ------------------------------------------------------------------------------------------
// Exhaustive self-check of orb_sincosf on the device (include/orb_b200.h: orb_sincosf_sweep_device).
__global__ void __launch_bounds__(256) k_sincosf_sweep(const uint32_t first, const unsigned long long count, unsigned long long* __restrict__ sum) {
    unsigned long long acc = 0ull;
    for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < count; i += (unsigned long long)gridDim.x * 256) {
        const uint32_t u = first + (uint32_t)i;
        float s, c;
        orb_sincosf(__uint_as_float(u), &s, &c);
        unsigned long long v = ((unsigned long long)__float_as_uint(s) << 32) | __float_as_uint(c);
        v ^= (unsigned long long)u * 0x9E3779B97F4A7C15ull;
        v *= 0xD6E8FEB86659FD93ull;
        v ^= v >> 32;
        acc += v;
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) atomicAdd(sum, acc);
}
extern "C" int orb_sincosf_sweep_device(uint32_t first_bits, uint32_t last_bits, int device, uint64_t* checksum) {
    if (!checksum || last_bits < first_bits) return ORB_ERR_BAD_ARG;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); orb_set_error("no CUDA device: this library has no CPU fallback"); return ORB_ERR_CUDA; }
    if (device < 0 || device >= ndev) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(device));
    unsigned long long* d = nullptr;
    ORB_CUDA(orb_dev_alloc((void**)&d, 8));
    ORB_CUDA(cudaMemset(d, 0, 8));
    k_sincosf_sweep<<<148 * 16, 256>>>(first_bits, (unsigned long long)last_bits - first_bits + 1ull, d);
    unsigned long long h = 0;
    cudaError_t e = cudaMemcpy(&h, d, 8, cudaMemcpyDeviceToHost);
    orb_dev_free(d);
    if (e != cudaSuccess) { orb_set_error(cudaGetErrorString(e)); return ORB_ERR_CUDA; }
    *checksum = h;
    return ORB_OK;
}

// Self-check of the device fastAtan2 used by the orientation stage (include/orb_b200.h: orb_fast_atan2_device).
__global__ void __launch_bounds__(256) k_fast_atan2(const float* __restrict__ y, const float* __restrict__ x, const int n, float* __restrict__ out) {
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < n) out[i] = fast_atan2_deg(y[i], x[i]);
}
extern "C" int orb_fast_atan2_device(const float* y, const float* x, int n, int device, float* out) {
    if (!y || !x || !out || n < 0) return ORB_ERR_BAD_ARG;
    if (n == 0) return ORB_OK;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); orb_set_error("no CUDA device: this library has no CPU fallback"); return ORB_ERR_CUDA; }
    if (device < 0 || device >= ndev) return ORB_ERR_BAD_ARG;
    ORB_CUDA(cudaSetDevice(device));
    float* d = nullptr;
    ORB_CUDA(orb_dev_alloc((void**)&d, (size_t)n * 12));
    cudaError_t e = cudaMemcpy(d, y, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(d + n, x, (size_t)n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) { k_fast_atan2<<<(n + 255) / 256, 256>>>(d, d + n, n, d + 2 * (size_t)n); e = cudaMemcpy(out, d + 2 * (size_t)n, (size_t)n * 4, cudaMemcpyDeviceToHost); }
    orb_dev_free(d);
    if (e != cudaSuccess) { orb_set_error(cudaGetErrorString(e)); return ORB_ERR_CUDA; }
    return ORB_OK;
}
