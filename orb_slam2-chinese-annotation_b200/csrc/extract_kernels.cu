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
    k_ingest<<<grid, 256, 0, s>>>(src, dst, w, h, stride, img_stride, dst_pitch, dst_img_stride, nch, magic);
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
// One CTA per group of ORB_FAST_G consecutive cells of a cell row (their interiors tile a
// contiguous strip).  The strip's 4-pixel words are split into 8 contiguous row-major chunks, one
// per warp, and every list below is a per-warp segment that stays in row-major order, so the
// final write needs no sort and no atomics:
//   (1) stage the strip in shared memory with 16-byte loads;
//   (2) byte-SIMD necessary test at minTh on 4 pixels per thread, survivors appended in order;
//   (3) exact score of the survivors, two pixels per thread on packed 16-bit lanes (VIADD.16x2 /
//       VIMNMX3.S16x2), into a shared score map; corners compacted in place, in order;
//   (4) strict local maxima inside the own cell, compacted in place; per-warp per-cell counts;
//   (5) per-cell threshold choice, prefix over warps, ordered (row-major) write.
// The tile pitch TP and score pitch TP - 32 are compile-time so that every ring / neighbour
// offset is an immediate.
#define FAST_THREADS 256
#define FAST_WARPS 8

__device__ __forceinline__ unsigned gt_bytes(unsigned ad, unsigned k) {  // bit 7 of byte j set <=> ad.j > 127 - k.j  (k.j <= 127)
    return ad | ((ad & 0x7f7f7f7fu) + k);
}

// FAST score of two pixels at once.  e[k] = ring[k] (raw bytes) on packed 16-bit lanes, c = the two centres.
// min / max commute with the subtraction of the centre, so the arc network runs on the raw ring:
// dark = (max over the 16 arcs of min over 9 contiguous e) - c; bright = c - (min over arcs of max over 9).
__device__ __forceinline__ unsigned fast_score_x2(const unsigned (&e)[16], const unsigned c) {
    unsigned lo3[16], hi3[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        lo3[k] = __vimin3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
        hi3[k] = __vimax3_s16x2(e[k], e[(k + 1) & 15], e[(k + 2) & 15]);
    }
    unsigned a9[16], z9[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) {
        a9[k] = __vimin3_s16x2(lo3[k], lo3[(k + 3) & 15], lo3[(k + 6) & 15]);
        z9[k] = __vimax3_s16x2(hi3[k], hi3[(k + 3) & 15], hi3[(k + 6) & 15]);
    }
    unsigned dk = __vimax3_s16x2(a9[0], a9[1], a9[2]), br = __vimin3_s16x2(z9[0], z9[1], z9[2]);
#pragma unroll
    for (int k = 3; k < 15; k += 2) {
        dk = __vimax3_s16x2(dk, a9[k], a9[k + 1]);
        br = __vimin3_s16x2(br, z9[k], z9[k + 1]);
    }
    // max(dark, bright) - 1 on both lanes
    return __vsub2(__vmaxs2(__vsub2(__vmaxs2(dk, a9[15]), c), __vsub2(c, __vmins2(br, z9[15]))), 0x00010001u);
}

// Dynamic shared memory: tile (rows x TP bytes, 16-byte aligned like the level) | score map
// ((ih + 2) x SP bytes, zero frame) | lists (u16: y << 8 | x)
template <int TP>
__global__ void __launch_bounds__(FAST_THREADS)
k_fast_cells(const uint8_t* __restrict__ pyr, int* __restrict__ cell_cnt, uint32_t* __restrict__ stage, const Plan P,
             const int smem_tile_bytes, const int smem_score_bytes) {
    constexpr int SP = TP - 32, TPW = TP / 4;
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ int s_wc[FAST_WARPS][ORB_FAST_G];   // per warp, per cell: maxima | maxima with s >= iniTh << 16

    int l = 0;
    while (l + 1 < P.nlevels && (int)blockIdx.x >= P.lv[l + 1].grp_off) ++l;
    const LevelGeom& G = P.lv[l];
    const int g = blockIdx.x - G.grp_off;
    const int ci = (int)(((unsigned)g * G.grp_magic) >> 20), cj0 = (g - ci * G.n_grp_x) * ORB_FAST_G;
    const int nc = min(ORB_FAST_G, G.n_cols - cj0);
    const int b = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned lt = (1u << lane) - 1;
    const int cell0 = ci * G.n_cols + cj0;
    int* out_cnt = cell_cnt + (size_t)b * P.total_cells + G.cell_off + cell0;
    const int w_cell = G.w_cell;

    // strip view [iniX, maxX) x [iniY, maxY) (reference :819-839).  A cell is skipped when
    // iniX >= maxBorderX - 6; only the last processed cell of a row can be clipped.
    const int iniY = ORB_WIN + ci * G.h_cell, iniX = ORB_WIN + cj0 * w_cell;
    int nv = 0;
    if (iniY < G.max_by - 3)
        for (int j = 0; j < nc; ++j) nv += (iniX + j * w_cell < G.max_bx - 6);
    const int maxY = min(iniY + G.h_cell + 6, G.max_by), maxX = min(iniX + nv * w_cell + 6, G.max_bx);
    const int tw = maxX - iniX, th = maxY - iniY;
    const int iw = tw - 6, ih = th - 6;  // interior = tested pixels
    if (nv == 0 || iw <= 0 || ih <= 0) { if (tid < nc) out_cnt[tid] = 0; return; }
    if (tid >= nv && tid < nc) out_cnt[tid] = 0;

    uint8_t* sc = smem + smem_tile_bytes;
    uint16_t* list = reinterpret_cast<uint16_t*>(sc + smem_score_bytes);

    // (1) stage with aligned 16-byte loads (view starts at byte `shift` of the tile rows)
    const int shift = iniX & 15;
    {
        const uint8_t* S = pyr + G.off + (size_t)b * G.img_stride + (size_t)iniY * G.pitch + (iniX - shift);
        const int nq = (tw + shift + 15) >> 4;
        const unsigned qmagic = (1u << 16) / (unsigned)nq + 1u;   // i / nq for i < 2^16 / nq
        uint4* t4 = reinterpret_cast<uint4*>(smem);
        for (int i = tid; i < nq * th; i += FAST_THREADS) {
            const int y = (int)(((unsigned)i * qmagic) >> 16), x = i - y * nq;
            cp_async16(&t4[y * (TP / 16) + x], reinterpret_cast<const uint4*>(S + (size_t)y * G.pitch) + x);
        }
        uint32_t* sc32 = reinterpret_cast<uint32_t*>(sc);
        for (int i = tid; i < (ih + 2) * (SP / 4); i += FAST_THREADS) sc32[i] = 0u;
        cp_async_wait_all();
    }
    __syncthreads();
    const uint32_t* tw32 = reinterpret_cast<const uint32_t*>(smem);
    const uint8_t* tile = smem + shift;
    // interior words of a row: word wx holds interior pixels xi = 4 * wx - shift - 3 + k, k = 0..3
    const int w0 = (shift + 3) >> 2, nwi = ((shift + 3 + iw - 1) >> 2) - w0 + 1;
    const unsigned gmagic = (1u << 20) / (unsigned)nwi + 1u;
    const int ngroups = nwi * ih;
    const int chunk = (ngroups + FAST_WARPS - 1) / FAST_WARPS;
    uint16_t* seg = list + warp * chunk * 4;
    const int t = P.min_th, ini_th = P.ini_th;

    // (2) necessary test: both ring pixels of the vertical pair (1, 9) and of the horizontal pair
    // (5, 13) within t of the centre => no 9-arc
    int n_a = 0;
    {
        const unsigned kk = (unsigned)(127 - min(t, 126)) * 0x01010101u;
        const int xfirst = 4 * w0 - shift - 3;                       // <= 0: interior x of byte 0 of the first word
        const unsigned mfirst = 0x80808080u << (8 * -xfirst);
        const int over = xfirst + 4 * nwi - iw;                      // bytes of the last word beyond the interior (0..3)
        const unsigned mlast = 0x80808080u >> (8 * over);
        const int gend = min(ngroups, (warp + 1) * chunk);
        for (int gi = warp * chunk + lane; gi - lane < gend; gi += 32) {
            unsigned m = 0, i0 = 0;
            if (gi < gend) {
                const int y = (int)(((unsigned)gi * gmagic) >> 20), wxi = gi - y * nwi;
                const uint32_t* c = tw32 + (y + 3) * TPW + w0 + wxi;
                const unsigned C = c[0], N = c[-3 * TPW], S = c[3 * TPW], L = c[-1], R = c[1];
                const unsigned E = __byte_perm(C, R, 0x6543), Wp = __byte_perm(L, C, 0x4321);  // x + 3, x - 3
                const unsigned pv = gt_bytes(__vabsdiffu4(N, C), kk) | gt_bytes(__vabsdiffu4(S, C), kk);
                const unsigned ph = gt_bytes(__vabsdiffu4(E, C), kk) | gt_bytes(__vabsdiffu4(Wp, C), kk);
                m = pv & ph & 0x80808080u;
                if (wxi == 0) m &= mfirst;
                if (wxi == nwi - 1) m &= mlast;
                i0 = (unsigned)((y << 8) + xfirst + 4 * wxi);
            }
            const int cnt = __popc(m);
            const unsigned b1 = __ballot_sync(0xffffffffu, cnt >= 1);
            if (b1) {
                const unsigned b2 = __ballot_sync(0xffffffffu, cnt >= 2), b3 = __ballot_sync(0xffffffffu, cnt >= 3),
                               b4 = __ballot_sync(0xffffffffu, cnt >= 4);
                int o = n_a + __popc(b1 & lt) + __popc(b2 & lt) + __popc(b3 & lt) + __popc(b4 & lt);
                if (m & 0x80u) seg[o++] = (uint16_t)i0;
                if (m & 0x8000u) seg[o++] = (uint16_t)(i0 + 1);
                if (m & 0x800000u) seg[o++] = (uint16_t)(i0 + 2);
                if (m & 0x80000000u) seg[o++] = (uint16_t)(i0 + 3);
                n_a += __popc(b1) + __popc(b2) + __popc(b3) + __popc(b4);
            }
        }
    }
    __syncwarp();

    // (3) exact score, two survivors per thread; corners (s >= minTh) compacted in place
    int n_c = 0;
    for (int k0 = 0; k0 < n_a; k0 += 64) {
        const int k = k0 + 2 * lane;
        const bool va = k < n_a, vb = k + 1 < n_a;
        const unsigned pr = va ? reinterpret_cast<const uint32_t*>(seg)[k >> 1] : 0u;  // segments start 8-byte aligned
        const int ia = (int)(pr & 0xffffu), ib = vb ? (int)(pr >> 16) : ia;
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
        const unsigned ba = __ballot_sync(0xffffffffu, fa), bb = __ballot_sync(0xffffffffu, fb);  // also orders the reads above before the writes below
        int o = n_c + __popc(ba & lt) + __popc(bb & lt);
        if (fa) seg[o++] = (uint16_t)ia;
        if (fb) seg[o] = (uint16_t)ib;
        n_c += __popc(ba) + __popc(bb);
        __syncwarp();
    }
    __syncthreads();   // score map complete

    // (4) strict local maxima inside the own cell's interior, compacted in place (in order); per-lane
    // counts per cell in 16-bit fields (cells 0,1 | cells 2,3), reduced once per warp
    unsigned ca01 = 0, ca23 = 0, cs01 = 0, cs23 = 0;
    int n_m = 0;
    for (int k0 = 0; k0 < n_c; k0 += 32) {
        const int k = k0 + lane;
        bool mx = false;
        int i = 0;
        if (k < n_c) {
            i = seg[k];
            const int y = i >> 8, x = i & 0xff;
            const uint8_t* q = sc + (y + 1) * SP + x + 1;
            const int s = q[0];
            const int c = (int)(((unsigned)x * G.wcell_magic) >> 20);
            const int lx = x - c * w_cell;
            const int cw = (c == nv - 1) ? iw - c * w_cell : w_cell;
            const int nb0 = max(q[-SP], q[SP]);
            const int nbl = max(max(q[-1], q[-SP - 1]), q[SP - 1]);
            const int nbr = max(max(q[1], q[-SP + 1]), q[SP + 1]);
            mx = s > nb0 && (lx == 0 || s > nbl) && (lx == cw - 1 || s > nbr);
            if (mx) {
                const unsigned inc = 1u << ((c & 1) << 4), incs = s >= ini_th ? inc : 0u;
                if (c < 2) { ca01 += inc; cs01 += incs; } else { ca23 += inc; cs23 += incs; }
            }
        }
        const unsigned bm = __ballot_sync(0xffffffffu, mx);   // also orders this iteration's reads before the writes
        if (mx) seg[n_m + __popc(bm & lt)] = (uint16_t)i;
        n_m += __popc(bm);
        __syncwarp();
    }
#pragma unroll
    for (int o = 16; o; o >>= 1) {
        ca01 += __shfl_xor_sync(0xffffffffu, ca01, o); ca23 += __shfl_xor_sync(0xffffffffu, ca23, o);
        cs01 += __shfl_xor_sync(0xffffffffu, cs01, o); cs23 += __shfl_xor_sync(0xffffffffu, cs23, o);
    }
    if (lane < ORB_FAST_G) {
        const unsigned a = lane < 2 ? ca01 : ca23, st = lane < 2 ? cs01 : cs23;
        const int sh = (lane & 1) << 4;
        s_wc[warp][lane] = (int)(((a >> sh) & 0xffffu) | (((st >> sh) & 0xffffu) << 16));
    }
    __syncthreads();

    // (5) threshold per cell (reference :843-850), offsets of this warp's segment, ordered write.
    // lane = 4 * w' + c holds the counts of warp w' for cell c (FAST_WARPS * ORB_FAST_G == 32)
    int tot = s_wc[lane >> 2][lane & 3];
    int before = (lane >> 2) < warp ? tot : 0;
#pragma unroll
    for (int o = 4; o < 32; o <<= 1) {
        tot += __shfl_xor_sync(0xffffffffu, tot, o);
        before += __shfl_xor_sync(0xffffffffu, before, o);
    }
    const bool use_ini = (tot >> 16) > 0;
    if (warp == 0 && lane < nv) out_cnt[lane] = use_ini ? (tot >> 16) : (tot & 0xffff);
    const int my_base = use_ini ? (before >> 16) : (before & 0xffff), my_th = use_ini ? ini_th : 0;
    int base_c[ORB_FAST_G], th_c[ORB_FAST_G];
#pragma unroll
    for (int c = 0; c < ORB_FAST_G; ++c) {
        base_c[c] = __shfl_sync(0xffffffffu, my_base, c);
        th_c[c] = __shfl_sync(0xffffffffu, my_th, c);
    }
    const int ky0 = 3 + ci * G.h_cell, kx00 = 3 + cj0 * w_cell;
    uint32_t* out0 = stage + (size_t)b * P.stage_per_image + G.stage_off + (size_t)cell0 * G.cell_cap;
    for (int k0 = 0; k0 < n_m; k0 += 32) {
        const int k = k0 + lane;
        int c = -1, s = 0, y = 0, x = 0;
        if (k < n_m) {
            const int i = seg[k];
            y = i >> 8; x = i & 0xff;
            s = sc[(y + 1) * SP + x + 1];
            c = (int)(((unsigned)x * G.wcell_magic) >> 20);
        }
#pragma unroll
        for (int cc = 0; cc < ORB_FAST_G; ++cc) {
            const bool mine = c == cc && s >= th_c[cc];
            const unsigned bm = __ballot_sync(0xffffffffu, mine);
            if (mine) out0[cc * G.cell_cap + base_c[cc] + __popc(bm & lt)] = ((uint32_t)s << 24) | ((uint32_t)(y + ky0) << 12) | (uint32_t)(x + kx00);
            base_c[cc] += __popc(bm);
        }
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

#define OD_WARPS 8
#define OD_KP_PER_CTA 32   // keypoint slots per CTA: each warp walks OD_KP_PER_CTA / OD_WARPS of them
#define OD_DW 11           // words per staged row of the 37 x 37 blurred patch (37 + 3 bytes of alignment slack)
#define OD_RW 9            // words per staged row of the 31 x 31 raw patch
// pattern copy in global memory: the CTA prologue reads it coalesced (constant memory would
// serialise the per-thread indices)
__device__ __align__(16) signed char g_pattern[1024] = {
#include "../../include/orb_pattern_31.inc"
};

// Weights of the intensity-centroid moments for the raw patch as it is fetched: 31 rows x OD_RW aligned words, the patch starting
// at byte rsh = 0..3 of the first word.  Entry [rsh][row * OD_RW + word] = (x: the four u = column - 15 as signed bytes, 0 outside
// the radius-15 disc (|u| <= umax[|v|]); y: 1 inside the disc, 0 outside).  m10 += dp4a(pixels, x); m01 += v * dp4a(pixels, y).
// Filled by orb_upload_moment_table() when an extractor is created (umax comes from the reference's table code, :466-489).
__device__ uint2 g_moment_w[4 * 32 * OD_RW];
cudaError_t orb_upload_moment_table(const int* umax) {
    uint2 h[4 * 32 * OD_RW];   // same contents from every caller (the disc is a constant of the algorithm)
    for (int rsh = 0; rsh < 4; ++rsh)
        for (int i = 0; i < 32 * OD_RW; ++i) {
            const int r = i / OD_RW, c = i % OD_RW, v = r - ORB_HALF_PATCH;
            uint32_t wu = 0, wm = 0;
            for (int j = 0; j < 4 && r < 31; ++j) {
                const int u = 4 * c + j - rsh - ORB_HALF_PATCH;
                if (u >= -ORB_HALF_PATCH && u <= ORB_HALF_PATCH && (u < 0 ? -u : u) <= umax[v < 0 ? -v : v]) {
                    wu |= (uint32_t)(uint8_t)(signed char)u << (8 * j);
                    wm |= 1u << (8 * j);
                }
            }
            h[rsh * 32 * OD_RW + i] = make_uint2(wu, wm);
        }
    return cudaMemcpyToSymbol(g_moment_w, h, sizeof(h));
}
__device__ __forceinline__ int dp4a_u8_s8(unsigned px, unsigned w, int acc) {
    int d;
    asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(px), "r"(w), "r"(acc));
    return d;
}

// Both patches are fetched with aligned 32-bit loads (all of them in flight together) and keep
// the level's word alignment in shared memory; the rBRIEF gathers and the moment columns then
// read bytes from shared memory only.
__global__ void __launch_bounds__(OD_WARPS * 32, 4)
k_orient_describe(const uint8_t* __restrict__ pyr, const uint8_t* __restrict__ blur, const uint32_t* __restrict__ lvl_kp,
                  const int* __restrict__ lvl_cnt, orb_keypoint* __restrict__ kps, uint8_t* __restrict__ desc,
                  int* __restrict__ counts, const Plan P) {
    __shared__ float2 s_pat[512];
    __shared__ uint32_t s_desc[OD_WARPS][37 * OD_DW];
    __shared__ int s_cnt[ORB_MAX_LEVELS], s_pre[ORB_MAX_LEVELS];
    const int b = blockIdx.y;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int* cnt = lvl_cnt + b * ORB_MAX_LEVELS;
    for (int i = threadIdx.x; i < 512; i += OD_WARPS * 32) {  // point i = byte (i / 16), sample (i % 16); stored [sample][byte]: conflict-free
        const char2 p = reinterpret_cast<const char2*>(g_pattern)[i];
        s_pat[(i & 15) * 32 + (i >> 4)] = make_float2((float)p.x, (float)p.y);
    }
    if (threadIdx.x == 0) {
        int t = 0;
        for (int l = 0; l < P.nlevels; ++l) { const int c = cnt[l]; s_cnt[l] = c; s_pre[l] = t; t += c; }
        if (blockIdx.x == 0) counts[b] = t;
    }
    __syncthreads();

    uint32_t* sd = s_desc[warp];
#pragma unroll 1
    for (int it = 0; it < OD_KP_PER_CTA / OD_WARPS; ++it) {
        const int slot = blockIdx.x * OD_KP_PER_CTA + it * OD_WARPS + warp;
        if (slot >= P.kp_cap) break;
        int l = 0;
        while (l + 1 < P.nlevels && slot >= P.lv[l + 1].kp_off) ++l;
        const LevelGeom& G = P.lv[l];
        const int idx = slot - G.kp_off;
        if (idx >= s_cnt[l]) continue;
        const int out_idx = idx + s_pre[l];

        const uint32_t key = lvl_kp[(size_t)b * P.kp_cap + slot];
        const int px = ORB_KEY_X(key) + ORB_WIN, py = ORB_KEY_Y(key) + ORB_WIN;
        const int pitch = G.pitch;
        const size_t img_off = G.off + (size_t)b * G.img_stride;
        // rows are 64-byte aligned, so the alignment of a patch row is that of its first column
        const int dsh = (px - 18) & 3, rsh = (px - 15) & 3;
        const uint8_t* dbase = blur + img_off + (size_t)(py - 18) * pitch + (px - 18 - dsh);
        const uint8_t* rbase = pyr + img_off + (size_t)(py - 15) * pitch + (px - 15 - rsh);
        __syncwarp();  // the previous keypoint's readers are done with sd / sr
        uint32_t vr[9], vd[13];
#pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int i = k * 32 + lane;
            const int r = (i * 57) >> 9, c = i - r * OD_RW;   // i / 9 for i < 288
            vr[k] = i < 31 * OD_RW ? __ldg(reinterpret_cast<const uint32_t*>(rbase + r * pitch) + c) : 0u;
        }
#pragma unroll
        for (int k = 0; k < 13; ++k) {
            const int i = k * 32 + lane;
            const int r = (i * 745) >> 13, c = i - r * OD_DW;  // i / 11 for i < 416
            vd[k] = i < 37 * OD_DW ? __ldg(reinterpret_cast<const uint32_t*>(dbase + r * pitch) + c) : 0u;
        }
#pragma unroll
        for (int k = 0; k < 13; ++k) { const int i = k * 32 + lane; if (i < 37 * OD_DW) sd[i] = vd[k]; }
        __syncwarp();

        // IC_Angle (reference :77-113): integer moments over the radius-15 disc, straight from the fetched words (exact integers,
        // any summation order): per word one byte dot product with the u weights and one with the disc mask
        int m10 = 0, m01 = 0;
        {
            const uint2* wt = g_moment_w + rsh * (32 * OD_RW) + lane;
#pragma unroll
            for (int k = 0; k < 9; ++k) {
                const uint2 wq = __ldg(wt + k * 32);
                const int v = (((k * 32 + lane) * 57) >> 9) - ORB_HALF_PATCH;
                m10 = dp4a_u8_s8(vr[k], wq.x, m10);
                m01 += v * dp4a_u8_s8(vr[k], wq.y, 0);
            }
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            m10 += __shfl_xor_sync(0xffffffffu, m10, o);
            m01 += __shfl_xor_sync(0xffffffffu, m01, o);
        }
        const float angle = fast_atan2_deg((float)m01, (float)m10);

        // rBRIEF (reference :119-164): lane = descriptor byte, 8 comparisons on the staged patch
        const float rad = __fmul_rn(angle, 0.017453292519943295f);  // factorPI = (float)(CV_PI/180.f)
        float a, bsin;
        orb_sincosf(rad, &bsin, &a);
        const uint8_t* spc = reinterpret_cast<const uint8_t*>(sd) + 18 * (OD_DW * 4) + 18 + dsh;
        int val = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const float2 p0 = s_pat[(2 * k) * 32 + lane], p1 = s_pat[(2 * k + 1) * 32 + lane];
            const int r0 = __float2int_rn(__fmaf_rn(p0.x, bsin, __fmul_rn(p0.y, a)));
            const int q0 = __float2int_rn(__fmaf_rn(p0.x, a, -__fmul_rn(p0.y, bsin)));
            const int r1 = __float2int_rn(__fmaf_rn(p1.x, bsin, __fmul_rn(p1.y, a)));
            const int q1 = __float2int_rn(__fmaf_rn(p1.x, a, -__fmul_rn(p1.y, bsin)));
            const int t0 = spc[r0 * (OD_DW * 4) + q0], t1 = spc[r1 * (OD_DW * 4) + q1];
            val |= (t0 < t1) << k;
        }
        desc[((size_t)b * P.kp_cap + out_idx) * 32 + lane] = (uint8_t)val;
        if (lane < 7) {  // one 4-byte field of the 28-byte record per lane
            float fx = (float)px, fy = (float)py;
            if (l != 0) { fx = __fmul_rn(fx, G.sf); fy = __fmul_rn(fy, G.sf); }
            uint32_t f;
            switch (lane) {
                case 0: f = __float_as_uint(fx); break;
                case 1: f = __float_as_uint(fy); break;
                case 2: f = __float_as_uint(G.patch_size); break;
                case 3: f = __float_as_uint(angle); break;
                case 4: f = __float_as_uint((float)ORB_KEY_S(key)); break;
                case 5: f = (uint32_t)l; break;
                default: f = 0xffffffffu; break;  // class_id = -1
            }
            reinterpret_cast<uint32_t*>(kps + (size_t)b * P.kp_cap + out_idx)[lane] = f;
        }
    }
}

// ------------------------------------------------------------------------------------------
// host-callable launchers (used by extractor.cu)
void launch_resize_untiled(cudaStream_t s, uint8_t* pyr, const LevelGeom& src, const LevelGeom& dst, int batch) {
    dim3 block(64, 4), grid((dst.w + 255) / 256, (dst.h + 4 * RESIZE_ROWS - 1) / (4 * RESIZE_ROWS), batch);
    k_resize_level<<<grid, block, 0, s>>>(pyr, src, dst);
}
// Tile pitch variant: 192 bytes covers w_cell <= 39 (every level with >= 4 cell columns), 272 the rest (w_cell <= 59).
static int fast_tile_pitch(const Plan& P) {
    int wc = 0;
    for (int l = 0; l < P.nlevels; ++l) if (P.lv[l].n_cols > 0) wc = max(wc, P.lv[l].w_cell);
    return (ORB_FAST_G * wc + 6 + 15 + 15 <= 192 + 15) && (ORB_FAST_G * wc + 2 + 3 <= 160 + 3) ? 192 : 272;
}
static void fast_smem(const Plan& P, int tp, int* tile_bytes, int* score_bytes, int* total) {
    int mh = 0, mv = 0;
    for (int l = 0; l < P.nlevels; ++l) {
        const LevelGeom& G = P.lv[l];
        if (G.n_cols == 0) continue;
        mh = max(mh, G.h_cell);
        // 4-pixel words per strip row (+2 for alignment) x rows, rounded up per warp, 4 entries each
        mv = max(mv, 4 * ((ORB_FAST_G * G.w_cell / 4 + 2) * G.h_cell + FAST_WARPS));
    }
    *tile_bytes = tp * (mh + 6);
    *score_bytes = (tp - 32) * (mh + 2);
    *total = *tile_bytes + *score_bytes + 2 * mv + 16;
}
cudaError_t launch_fast(cudaStream_t s, const uint8_t* pyr, int* cell_cnt, uint32_t* stage, const Plan& P, int batch) {
    const int tp = fast_tile_pitch(P);
    int tb, sb, tot;
    fast_smem(P, tp, &tb, &sb, &tot);
    if (tot > 200 * 1024) return cudaErrorInvalidConfiguration;
    auto kern = tp == 192 ? k_fast_cells<192> : k_fast_cells<272>;
    if (tot > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, tot);
        if (e != cudaSuccess) return e;
    }
    dim3 grid(P.total_groups, batch);
    kern<<<grid, FAST_THREADS, tot, s>>>(pyr, cell_cnt, stage, P, tb, sb);
    return cudaSuccess;
}
void launch_orient_describe(cudaStream_t s, const uint8_t* pyr, const uint8_t* blur, const uint32_t* lvl_kp, const int* lvl_cnt,
                            orb_keypoint* kps, uint8_t* desc, int* counts, const Plan& P, int batch) {
    dim3 grid((P.kp_cap + OD_KP_PER_CTA - 1) / OD_KP_PER_CTA, batch);
    k_orient_describe<<<grid, OD_WARPS * 32, 0, s>>>(pyr, blur, lvl_kp, lvl_cnt, kps, desc, counts, P);
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
