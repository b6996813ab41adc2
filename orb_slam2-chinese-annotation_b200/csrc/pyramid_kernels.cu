// sm_100a kernels of ComputePyramid and the descriptor blur (reference src/ORBextractor.cc:1172-1207, 1143-1145),
// staged through shared memory by the Tensor Memory Accelerator.
//
//   k_resize_tma   cv::resize(INTER_LINEAR) 8u fixed point, level l <- level l-1          :1195
//   k_gauss7_tma   GaussianBlur(7x7, sigma 2, BORDER_REFLECT_101) 8u fixed point           :1143-1145
//
// Both read their source tile with ONE cp.async.bulk.tensor (SASS: UTMALDG) issued by one thread: the hardware does
// the 2-D addressing, the halo and the out-of-image zero fill, so the kernels spend no instructions on staging.  The
// arithmetic runs on two different issue pipes (profiles/MEASURED_INT_PEAKS.json): IDP.4A / IDP.2A / IMAD on the FMA
// pipe, IADD3 / PRMT / SHF on the ALU pipe, 64 lanes per clock per SM each.
#include "orb_internal.h"

// ------------------------------------------------------------------------------------------ TMA / mbarrier PTX
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");   // make the init visible to the async proxy
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned phase) {
    unsigned ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(phase) : "memory");
    } while (!ok);
}
// box (x, y, image) of a 3-D u8 tensor -> shared memory; completes on the mbarrier with the box's byte count
__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int x, int y, int z) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(smem_u32(smem_dst)), "l"(map), "r"(smem_u32(bar)), "r"(x), "r"(y), "r"(z) : "memory");
}

__device__ __forceinline__ int reflect101(int p, int len) {
    if (len == 1) return 0;
    while (p < 0 || p >= len) p = p < 0 ? -p : 2 * (len - 1) - p;
    return p;
}

// ------------------------------------------------------------------------------------------
// Pyramid: level l <- level l-1, OpenCV 8u INTER_LINEAR (11-bit coefficients).
// One CTA (4 warps) produces a 128 x 64 destination tile from one TMA-staged source box.  The box origin is a lower
// bound of (xofs[dx0], yofs[dy0]) computed from the tile index alone ((dx0 * floor(2^16 sw / dw)) >> 16, rounded down
// to 16 bytes: the innermost TMA coordinate must be a multiple of 16 bytes), so thread 0 issues the load before any
// table lookup.  Scales up to 1.35 (x) / 1.33 (y) fit the 192 x 88 box (host-checked; larger scales take
// k_resize_level in extract_kernels.cu).  A lane owns 4 adjacent destination columns and walks down 16 rows; the
// horizontal pass of a source row ((pixel, right neighbour) byte pair cut out of two aligned shared-memory words with
// one PRMT, weighted with one IDP.2A) is kept and reused by the next destination row, which shares that source row
// in 5 of 6 cases at scale 1.2.  The vertical blend is two IMAD.HI.
#define RZ_WARPS 4
#define RZ_ROWS (ORB_RZ_DH / RZ_WARPS)   // destination rows per warp
__device__ __forceinline__ unsigned mad_u32(unsigned a, unsigned b, unsigned c) {   // stays on the FMA pipe (the ALU pipe is the busy one here)
    unsigned r;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(c));
    return r;
}
struct RzCols {   // per-lane column state: byte offsets of the two 8-byte windows, PRMT selectors, packed weights
    int wa, wb;
    unsigned sel[4], al[4];
};
__device__ __forceinline__ void rz_row(const uint8_t* __restrict__ row, const RzCols& c, unsigned (&h)[4]) {
    const unsigned* pa = reinterpret_cast<const unsigned*>(row + c.wa);
    const unsigned* pb = reinterpret_cast<const unsigned*>(row + c.wb);
    const unsigned a0 = pa[0], a1 = pa[1], b0 = pb[0], b1 = pb[1];
    h[0] = __dp2a_lo(c.al[0], __byte_perm(a0, a1, c.sel[0]), 0u) >> 4;
    h[1] = __dp2a_lo(c.al[1], __byte_perm(a0, a1, c.sel[1]), 0u) >> 4;
    h[2] = __dp2a_lo(c.al[2], __byte_perm(b0, b1, c.sel[2]), 0u) >> 4;
    h[3] = __dp2a_lo(c.al[3], __byte_perm(b0, b1, c.sel[3]), 0u) >> 4;
}

__global__ void __launch_bounds__(RZ_WARPS * 32) k_resize_tma(const __grid_constant__ CUtensorMap src_map, uint8_t* __restrict__ pyr,
                                                              const LevelGeom src, const LevelGeom dst, const int b0) {
    __shared__ __align__(128) uint8_t tile[ORB_RZ_BOXH * ORB_RZ_BOXW + 16];   // +16: the last lane's 8-byte window may overhang
    __shared__ uint2 rows[ORB_RZ_DH];   // per destination row: byte offsets of its two source rows in the tile | packed vertical weights
    __shared__ uint64_t bar;
    ORB_PDL_SYNC();
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int dx0 = blockIdx.x * ORB_RZ_DW, dy0 = blockIdx.y * ORB_RZ_DH;
    const int sx0 = (int)(((unsigned)dx0 * dst.rz_qx) >> 16) & ~15, sy0 = (int)(((unsigned)dy0 * dst.rz_qy) >> 16);
    if (tid == 0) {
        mbar_init(&bar, 1);
        mbar_expect_tx(&bar, ORB_RZ_BOXH * ORB_RZ_BOXW);
        tma_load_3d(tile, &src_map, &bar, sx0, sy0, b0 + (int)blockIdx.z);
    }
    // coefficient tables while the box is in flight
    if (tid < ORB_RZ_DH) {
        const int y = min(dy0 + tid, dst.h - 1);
        const int r0 = dst.yofs[y] - sy0;                   // 0 <= yofs <= src.h - 1 for scales >= 1
        const int r1 = min(r0 + 1, src.h - 1 - sy0);
        rows[tid] = make_uint2((unsigned)(r0 * ORB_RZ_BOXW) | ((unsigned)(r1 * ORB_RZ_BOXW) << 16), reinterpret_cast<const unsigned*>(dst.beta)[y]);
    }
    const int x4 = dx0 + 4 * lane;
    const int y0 = dy0 + RZ_ROWS * warp;
    const bool live = x4 < dst.w && y0 < dst.h;
    RzCols c;
    if (live) {   // tables are padded to 4 entries and 16-byte aligned
        const int4 xo = reinterpret_cast<const int4*>(dst.xofs)[x4 >> 2];
        const uint4 al = reinterpret_cast<const uint4*>(dst.alpha)[x4 >> 2];
        // source columns step by 1 or 2 per destination column (scale <= 1.35): pixels 0,1 and pixels 2,3 each fit an 8-byte window
        const int sa = xo.x - sx0, sb = (x4 + 2 < dst.w ? xo.z : xo.x) - sx0;   // table entries past the last column are padding
        c.wa = sa & ~3; c.wb = sb & ~3;
        const unsigned j0 = sa & 3, j1 = (xo.y - sx0) - c.wa, j2 = sb & 3, j3 = (xo.w - sx0) - c.wb;
        c.sel[0] = j0 | ((j0 + 1) << 4); c.sel[1] = j1 | ((j1 + 1) << 4); c.sel[2] = j2 | ((j2 + 1) << 4); c.sel[3] = j3 | ((j3 + 1) << 4);
        c.al[0] = al.x; c.al[1] = al.y; c.al[2] = al.z; c.al[3] = al.w;
    }
    __syncthreads();   // row table and barrier initialisation visible to everyone
    mbar_wait(&bar, 0);
    if (!live) return;
    uint8_t* D = pyr + dst.off + (size_t)(b0 + blockIdx.z) * dst.img_stride + (size_t)y0 * dst.pitch + x4;
    const int nrows = min(RZ_ROWS, dst.h - y0);
    unsigned cur = 0xffffffffu;   // tile offset of the source row whose horizontal pass is in hc
    unsigned hc[4];
#pragma unroll
    for (int r = 0; r < RZ_ROWS; ++r) {
        if (r >= nrows) break;
        const uint2 rw = rows[RZ_ROWS * warp + r];
        const unsigned o0 = rw.x & 0xffffu, o1 = rw.x >> 16;
        unsigned ha[4];
        if (o0 == cur) {   // warp-uniform
#pragma unroll
            for (int i = 0; i < 4; ++i) ha[i] = hc[i];
        } else {
            rz_row(tile + o0, c, ha);
        }
        if (o1 != o0) rz_row(tile + o1, c, hc);
        else {
#pragma unroll
            for (int i = 0; i < 4; ++i) hc[i] = ha[i];
        }
        cur = o1;
        const unsigned B0 = rw.y << 16, B1 = rw.y & 0xffff0000u;   // weights pre-shifted: umulhi(v, b << 16) == (b * v) >> 16
        unsigned v[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) v[i] = (__umulhi(ha[i], B0) + __umulhi(hc[i], B1) + 2u) >> 2;   // convex combination of bytes: no clamp
        // rows are pitch-padded to a multiple of 64 bytes, so the full word is always in bounds
        *reinterpret_cast<uint32_t*>(D) = mad_u32(mad_u32(v[3], 256u, v[2]), 65536u, mad_u32(v[1], 256u, v[0]));
        D += dst.pitch;
    }
}

// ------------------------------------------------------------------------------------------
// 7x7 Gaussian, sigma 2, BORDER_REFLECT_101 at the level's own edge, OpenCV >= 4 fixed point:
// Q0.8 taps {18,34,48,56,48,34,18}; rows Q8.8, columns Q16.16, (v + 2^15) >> 16.  All arithmetic is exact integer,
// so the result does not depend on the order of the two passes.
// CTA = 2 warps over a 128 x 64 tile; the (128 + 32) x (64 + 6) box arrives by TMA (origin x - 16: 16-byte rule).  A lane owns 4 adjacent
// columns and walks down 32 rows with the last 7 row sums in registers.  Row sums: ten IDP.4A per 4 pixels on the
// three ALIGNED words that cover them, with the taps shifted inside the weight words instead of the bytes (no PRMT).
#define B2_ROWS 32   // rows per warp
#define B2_WARPS (ORB_BLUR_TH / B2_ROWS)
#define B2_RW (ORB_BLUR_BOXW / 4)
#define TAPS(a, b, c, d) ((unsigned)(a) | ((unsigned)(b) << 8) | ((unsigned)(c) << 16) | ((unsigned)(d) << 24))

__global__ void __launch_bounds__(B2_WARPS * 32) k_gauss7_tma(const __grid_constant__ TmaMaps maps, const uint8_t* __restrict__ pyr,
                                                              uint8_t* __restrict__ blur, const Plan P, const int b0) {
    __shared__ __align__(128) uint32_t raw[ORB_BLUR_BOXH][B2_RW];
    __shared__ uint64_t bar;
    int l = 0;
    while (l + 1 < P.nlevels && (int)blockIdx.x >= P.lv[l + 1].tile_off) ++l;
    const LevelGeom& G = P.lv[l];
    const int t = blockIdx.x - G.tile_off;
    const int tyi = t / G.tiles_x;
    const int tx = (t - tyi * G.tiles_x) * ORB_BLUR_TW, ty = tyi * ORB_BLUR_TH;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int W = G.w, H = G.h;
    const int img = b0 + (int)blockIdx.y;
    if (tid == 0) mbar_init(&bar, 1);
    __syncthreads();
    if (tid == 0) {
        mbar_expect_tx(&bar, ORB_BLUR_BOXH * ORB_BLUR_BOXW);
        tma_load_3d(raw, &maps.lv[l], &bar, tx - 16, ty - 3, img);
    }
    mbar_wait(&bar, 0);

    // edge tiles: the zero fill outside the image becomes BORDER_REFLECT_101 (columns of the in-image rows first,
    // then whole rows; the sources are always in-image elements of the same box)
    if (tx == 0 || tx + ORB_BLUR_TW + 3 > W || ty == 0 || ty + ORB_BLUR_TH + 3 > H) {
        uint8_t* rb = reinterpret_cast<uint8_t*>(&raw[0][0]);
        if (W < 8 || H < 8) {   // tiny levels: several reflections possible; fetch from the level itself
            const uint8_t* S = pyr + G.off + (size_t)img * G.img_stride;
            for (int i = tid; i < ORB_BLUR_BOXH * ORB_BLUR_BOXW; i += B2_WARPS * 32) {
                const int ry = i / ORB_BLUR_BOXW, rx = i - ry * ORB_BLUR_BOXW;
                const int gy = ty - 3 + ry, gx = tx - 16 + rx;
                if (gx >= 0 && gx < W && gy >= 0 && gy < H) continue;
                rb[i] = S[(size_t)reflect101(min(max(gy, -3), H + 2), H) * G.pitch + reflect101(min(max(gx, -3), W + 2), W)];
            }
        } else {
            for (int ry = tid; ry < ORB_BLUR_BOXH; ry += B2_WARPS * 32) {
                const int gy = ty - 3 + ry;
                if (gy < 0 || gy >= H) continue;
                uint8_t* row = rb + ry * ORB_BLUR_BOXW;
                if (tx == 0) { row[15] = row[17]; row[14] = row[18]; row[13] = row[19]; }
#pragma unroll
                for (int k = 0; k < 3; ++k) {
                    const int rx = W + k - (tx - 16);
                    if (rx >= 0 && rx < ORB_BLUR_BOXW) row[rx] = row[rx - 2 * (k + 1)];   // W-1+(k+1) <- W-1-(k+1)
                }
            }
            __syncthreads();
            for (int i = tid; i < 6 * B2_RW; i += B2_WARPS * 32) {
                const int k = i / B2_RW, j = i - k * B2_RW;
                const int gy = k < 3 ? -1 - k : H + (k - 3);
                const int ry = gy - (ty - 3);
                if (ry < 0 || ry >= ORB_BLUR_BOXH) continue;
                const int sy = (gy < 0 ? -gy : 2 * (H - 1) - gy) - (ty - 3);
                raw[ry][j] = raw[sy][j];
            }
        }
        __syncthreads();
    }

    const int ox = tx + 4 * lane, r0 = warp * B2_ROWS;
    if (ox >= W || ty + r0 >= H) return;
    uint8_t* D = blur + G.off + (size_t)img * G.img_stride + ox;
    // Column pass on vertical PAIRS of row sums (each < 2^16): q[j] = (row a + j, row a + j + 1) of the 7-row window starting at row a,
    // one PRMT per new row, so that a column costs three IDP.2A + one IMAD instead of three adds + four IMAD.
    uint32_t q[6][4], hl[4] = {0u, 0u, 0u, 0u};
#pragma unroll
    for (int i = 0; i < B2_ROWS + 6; ++i) {
        const uint32_t w0 = raw[r0 + i][lane + 3], w1 = raw[r0 + i][lane + 4], w2 = raw[r0 + i][lane + 5];   // pixels ox-4.., ox.., ox+4..
        uint32_t hn[4];
        hn[0] = __dp4a(w0, TAPS(0, 18, 34, 48), __dp4a(w1, TAPS(56, 48, 34, 18), 0u));
        hn[1] = __dp4a(w0, TAPS(0, 0, 18, 34), __dp4a(w1, TAPS(48, 56, 48, 34), __dp4a(w2, TAPS(18, 0, 0, 0), 0u)));
        hn[2] = __dp4a(w0, TAPS(0, 0, 0, 18), __dp4a(w1, TAPS(34, 48, 56, 48), __dp4a(w2, TAPS(34, 18, 0, 0), 0u)));
        hn[3] = __dp4a(w1, TAPS(18, 34, 48, 56), __dp4a(w2, TAPS(48, 34, 18, 0), 0u));
#pragma unroll
        for (int k = 0; k < 4; ++k) {
#pragma unroll
            for (int j = 0; j < 5; ++j) q[j][k] = q[j + 1][k];
            q[5][k] = __byte_perm(hl[k], hn[k], 0x5410);   // (row i - 1, row i)
            hl[k] = hn[k];
        }
        if (i >= 6) {
            const int oy = ty + r0 + i - 6;
            if (oy >= H) break;
            uint32_t v[4];
#pragma unroll
            for (int k = 0; k < 4; ++k)   // rows i-6 .. i with taps 18 34 48 56 48 34 18
                v[k] = __dp2a_lo(q[0][k], 18u | (34u << 8), __dp2a_lo(q[2][k], 48u | (56u << 8), __dp2a_lo(q[4][k], 48u | (34u << 8), 18u * hn[k] + 32768u)));
            // byte 2 of each 24-bit sum is the rounded result
            const uint32_t t01 = __byte_perm(v[0], v[1], 0x0062), t23 = __byte_perm(v[2], v[3], 0x0062);
            *reinterpret_cast<uint32_t*>(D + (size_t)oy * G.pitch) = __byte_perm(t01, t23, 0x5410);
        }
    }
}

// ------------------------------------------------------------------------------------------ launchers
void launch_resize_untiled(cudaStream_t s, uint8_t* pyr, const LevelGeom& src, const LevelGeom& dst, int batch);   // extract_kernels.cu

// `src` / `dst` carry the level offsets of image 0 of the batch layout; b0 = first image slot of this run
void launch_resize(cudaStream_t s, uint8_t* pyr, const CUtensorMap* src_map, bool tma, const LevelGeom& src, const LevelGeom& dst, int b0, int batch) {
    if (!tma) {
        LevelGeom a = src, b = dst;
        a.off += (unsigned long long)b0 * a.img_stride;
        b.off += (unsigned long long)b0 * b.img_stride;
        launch_resize_untiled(s, pyr, a, b, batch);
        return;
    }
    dim3 grid((dst.w + ORB_RZ_DW - 1) / ORB_RZ_DW, (dst.h + ORB_RZ_DH - 1) / ORB_RZ_DH, batch);
    orb_launch(orb_pdl(batch), k_resize_tma, grid, dim3(RZ_WARPS * 32), 0, s, *src_map, pyr, src, dst, b0);
}
void launch_gauss(cudaStream_t s, const TmaMaps& maps, const uint8_t* pyr, uint8_t* blur, const Plan& P, int b0, int batch) {
    dim3 grid(P.total_tiles, batch);
    k_gauss7_tma<<<grid, B2_WARPS * 32, 0, s>>>(maps, pyr, blur, P, b0);
}
