// sm_100a kernel of ORBextractor::DistributeOctTree / ExtractorNode::DivideNode (reference src/ORBextractor.cc:500-782).
//
// One CTA per (level, image).  The reference's std::list<ExtractorNode> is emulated with an array R holding the list
// in REVERSE order (push_front == append), compacted after every pass.  Keys are never moved: every key carries the
// index of the list node that currently holds it ("leaf id"), and one coalesced sweep per pass, one thread per key,
//   - renames the leaf id through the old-node -> new-node table of the pass just committed, and
//   - counts the key into the quadrant it would fall into if its new node is divided next (warp-aggregated
//     shared-memory atomics),
// so that the sequential part of a pass only touches node records (at most nfeatures-per-level + 3 of them), never
// key lists.  A node's box fixes the quadrant of every key geometrically (DivideNode splits at the box centre), which
// is why the counts can be taken before the reference would have built the child key vectors.
//
// Order contract.  The reference keeps keys inside a node in their original order (cell row-major, pixel row-major
// within a cell) and picks the FIRST key of maximal response (:761-779).  That order is a function of the key's
// coordinates alone (the FAST cell interiors tile the window), so the selection is an atomic max on the response
// followed by an atomic min on (original order, key) among the keys that reach it.
// Equal-size ties of the sorted expansion phase (reference :703 sorts pair<int, ExtractorNode*>, i.e. by heap
// address) use the canonical rule "later-created node first" == higher index in R.
#include "orb_internal.h"

// CTA size NT is a template parameter: 256 threads when the batch fills the GPU anyway (throughput: fewest barrier stalls per unit of
// work), 512 for very large levels (3840 x 2160 / 8000 features: level 0 sweeps ~10^5 candidates per pass), 1024 when only a few CTAs
// exist at all (a live camera, batch 1: the level-0 CTA IS the latency of the stage).
#define QT_MAX_WARPS 32

struct QNode {
    int count;
    short x0, x1, y0, y1;  // UL.x, UR.x, UL.y, BL.y
    int split;             // where DivideNode would cut this box: sx | sy << 16 (set when the node is created)
};
__device__ __forceinline__ int node_split(int x0, int x1, int y0, int y1) {   // ceil(half), reference :502-503
    return (x0 + ((x1 - x0 + 1) >> 1)) | ((y0 + ((y1 - y0 + 1) >> 1)) << 16);
}

struct QtShared {
    int warp_sums[2][QT_MAX_WARPS + 1];
    int scal[4];
};

// exclusive scan over the CTA of one value per thread; returns the exclusive prefix, total in *total.  `par`
// alternates between consecutive calls (two copies of the warp totals: no trailing barrier needed).
template <int NT>
__device__ __forceinline__ int block_excl_scan(int v, int* total, QtShared& sh, int& par) {
    constexpr int QT_WARPS = NT / 32;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int* ws = sh.warp_sums[par];
    par ^= 1;
    int inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += n;
    }
    if (lane == 31) ws[warp] = inc;
    __syncthreads();
    int w = lane < QT_WARPS ? ws[lane] : 0;   // every warp scans the warp totals itself
    int winc = w;
#pragma unroll
    for (int o = 1; o < QT_WARPS; o <<= 1) {
        const int n = __shfl_up_sync(0xffffffffu, winc, o);
        if (lane >= o) winc += n;
    }
    *total = __shfl_sync(0xffffffffu, winc, QT_WARPS - 1);
    return inc - v + __shfl_sync(0xffffffffu, winc - w, warp);
}

__device__ __forceinline__ int node_quadrant(const QNode& nd, uint32_t k) {
    return (ORB_KEY_X(k) < (nd.split & 0xffff) ? 0 : 1) + (ORB_KEY_Y(k) < (nd.split >> 16) ? 0 : 2);
}

// counter[target] += 1 for every lane with target >= 0; lanes of a warp with the same target add once (all 32 lanes call)
__device__ __forceinline__ void warp_count(int* counter, int target) {
    const unsigned m = __match_any_sync(0xffffffffu, target);
    if (target >= 0 && (int)(threadIdx.x & 31) == __ffs(m) - 1) atomicAdd(counter + target, __popc(m));
}

// bitonic sort, descending, of n (power of two) 64-bit keys in shared memory
template <int QT_THREADS>
__device__ void block_sort_desc(unsigned long long* a, int n) {
    for (int k = 2; k <= n; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < n; i += QT_THREADS) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const unsigned long long x = a[i], y = a[ixj];
                    const bool up = (i & k) == 0;  // descending overall: "up" blocks keep larger first
                    if (up ? (x < y) : (x > y)) { a[i] = y; a[ixj] = x; }
                }
            }
            __syncthreads();
        }
    }
}

// Dynamic shared memory layout (cap = node_cap_max, p2 = next pow2 >= cap):
//   QNode nodeA[cap] | QNode nodeB[cap] | int cnt4[4 cap] | int nid[4 cap] | int par[cap] | int pre_ne[cap + 1] | int pre_g[cap] |
//   int flag[cap] | u64 sortk[p2]
template <int QT_THREADS>
__global__ void __launch_bounds__(QT_THREADS)
k_quadtree(const int* __restrict__ cell_cnt, int* __restrict__ cell_prefix, const uint32_t* __restrict__ stage,
           uint32_t* __restrict__ keys, uint32_t* __restrict__ lvl_kp, int* __restrict__ lvl_cnt, int* __restrict__ lvl_ncand,
           int* __restrict__ status, const Plan P, const int cap, const int p2) {
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ QtShared sh;
    ORB_PDL_SYNC();
    // blocks are dispatched in linear order (x fastest): image = x, level = y starts every image's level 0 (the longest CTA by far:
    // it sweeps the most candidates per pass) before any level 1, i.e. longest-processing-time-first over the whole batch
    const int b = blockIdx.x, l = blockIdx.y;
    const LevelGeom& G = P.lv[l];
    const int tid = threadIdx.x;
    int sp = 0;   // scan parity

    QNode* cur = reinterpret_cast<QNode*>(smem);
    QNode* nxt = cur + cap;
    int* cnt4 = reinterpret_cast<int*>(nxt + cap);   // occupancy of the four quadrants of every node of the current list
    int* nid = cnt4 + 4 * cap;                        // old node, quadrant -> node of the new list
    int* par = nid + 4 * cap;
    int* pre_ne = par + cap;
    int* pre_g = pre_ne + cap + 1;
    int* flag = pre_g + cap;
    unsigned long long* sortk = reinterpret_cast<unsigned long long*>(smem + (((size_t)((uint8_t*)(flag + cap) - smem) + 7) & ~(size_t)7));

    const int n_cells = G.n_cols * G.n_rows;
    const int* ccnt = cell_cnt + (size_t)b * P.total_cells + G.cell_off;
    int* cpre = cell_prefix + (size_t)b * P.total_cells + G.cell_off;
    const uint32_t* stg = stage + (size_t)b * P.stage_per_image + G.stage_off;
    uint32_t* kbuf = keys + (size_t)b * 2 * P.stage_per_image + G.stage_off;   // the level's candidates, original order
    uint32_t* leaf = kbuf + P.stage_per_image;                                  // list node holding each candidate
    uint32_t* out = lvl_kp + (size_t)b * P.kp_cap + G.kp_off;

    // ---- 0. ordered gather of the per-cell FAST lists (cells row-major)
    int carry = 0;
    for (int c0 = 0; c0 < n_cells; c0 += QT_THREADS) {
        const int c = c0 + tid;
        const int v = c < n_cells ? ccnt[c] : 0;
        int tot;
        const int ex = block_excl_scan<QT_THREADS>(v, &tot, sh, sp);
        if (c < n_cells) cpre[c] = carry + ex;
        carry += tot;
    }
    const int n = carry;
    __syncthreads();
    // one thread per cell (lists are short; a warp per cell would serialise n_cells / 8 dependent round trips)
    for (int c = tid; c < n_cells; c += QT_THREADS) {
        const int cnt = ccnt[c], o = cpre[c];
        const uint32_t* s = stg + (size_t)c * G.cell_cap;
#pragma unroll 4
        for (int i = 0; i < cnt; ++i) kbuf[o + i] = s[i];
    }
    if (tid == 0) lvl_ncand[b * ORB_MAX_LEVELS + l] = n;
    if (n == 0) { if (tid == 0) lvl_cnt[b * ORB_MAX_LEVELS + l] = 0; return; }

    // ---- 1. root nodes (reference :562-607): key -> root (int)(x / hX)
    const int N = G.n_feat;
    const int n_ini = G.n_ini;
    const float hX = G.h_x;
    for (int r = tid; r < n_ini; r += QT_THREADS) nid[r] = 0;   // keys per root
    __syncthreads();   // also: kbuf complete
    for (int i0 = 0; i0 < n; i0 += QT_THREADS) {
        const int i = i0 + tid;
        int r = -1;
        if (i < n) {
            r = min((int)__fdiv_rn((float)ORB_KEY_X(kbuf[i]), hX), n_ini - 1);
            leaf[i] = (uint32_t)r;
        }
        warp_count(nid, r);
    }
    __syncthreads();
    if (tid == 0) {   // drop empty roots; list order r = 0 .. n_ini-1 == array order reversed
        int m = 0;
        for (int r = n_ini - 1; r >= 0; --r) {
            const int c = nid[r];
            if (c > 0) {
                QNode nd;
                nd.count = c;
                nd.x0 = (short)(int)(hX * (float)r);
                nd.x1 = (short)(int)(hX * (float)(r + 1));
                nd.y0 = 0;
                nd.y1 = (short)G.win_h;
                nd.split = node_split(nd.x0, nd.x1, nd.y0, nd.y1);
                par[r] = m;
                cur[m++] = nd;
            }
        }
        sh.scal[0] = m;
    }
    __syncthreads();
    int size = sh.scal[0];
    for (int i = tid; i < 4 * size; i += QT_THREADS) cnt4[i] = 0;
    __syncthreads();
    for (int i0 = 0; i0 < n; i0 += QT_THREADS) {
        const int i = i0 + tid;
        int t = -1;
        if (i < n) {
            const int nl = par[leaf[i]];
            leaf[i] = (uint32_t)nl;
            const QNode nd = cur[nl];
            if (nd.count > 1) t = 4 * nl + node_quadrant(nd, kbuf[i]);
        }
        warp_count(cnt4, t);
    }
    __syncthreads();

    // ---- 2. expansion passes
    bool sorted_phase = false;
    bool finish = false;
    while (!finish) {
        // a. parents in processing order (list order, or by descending (count, index) in the sorted phase); the same scan
        // ranks the nodes that survive an unsorted pass (every node that is not a parent)
        int np = 0, n_keep = 0;
        {
            int c_par = 0, c_keep = 0;
            for (int i0 = 0; i0 < size; i0 += QT_THREADS) {
                const int i = i0 + tid;
                const int f = (i < size && cur[i].count > 1) ? 1 : 0;
                const int g = (i < size && !f) ? 1 : 0;
                int tot;
                const int ex = block_excl_scan<QT_THREADS>(f | (g << 16), &tot, sh, sp);
                if (f) {
                    if (sorted_phase) sortk[c_par + (ex & 0xffff)] = ((unsigned long long)(unsigned)cur[i].count << 32) | (unsigned)i;
                    else flag[i] = c_par + (ex & 0xffff);  // ascending rank, flipped below
                } else if (g) flag[i] = c_keep + (ex >> 16);   // position among the survivors (unsorted pass)
                c_par += tot & 0xffff;
                c_keep += tot >> 16;
            }
            np = c_par; n_keep = c_keep;
        }
        if (np == 0) break;  // every node holds one key: list cannot grow (uniform)
        if (!sorted_phase) {
            __syncthreads();
            for (int i = tid; i < size; i += QT_THREADS)
                if (cur[i].count > 1) par[np - 1 - flag[i]] = i;
        } else {
            int np2 = 1;
            while (np2 < np) np2 <<= 1;
            for (int i = np + tid; i < np2; i += QT_THREADS) sortk[i] = 0ull;
            __syncthreads();
            block_sort_desc<QT_THREADS>(sortk, np2);
            for (int i = tid; i < np; i += QT_THREADS) par[i] = (int)(sortk[i] & 0xffffffffu);
        }
        __syncthreads();

        // c. prefix of non-empty children and of list growth, in processing order
        int carry_ne = 0;
        for (int k0 = 0; k0 < np; k0 += QT_THREADS) {
            const int k = k0 + tid;
            int ne = 0;
            if (k < np) {
                const int4 c = reinterpret_cast<const int4*>(cnt4)[par[k]];
                ne = (c.x > 0) + (c.y > 0) + (c.z > 0) + (c.w > 0);
            }
            int tot;
            const int ex = block_excl_scan<QT_THREADS>(ne, &tot, sh, sp);
            if (k < np) {
                pre_ne[k] = carry_ne + ex;
                pre_g[k] = carry_ne + ex + ne - (k + 1);   // size growth after dividing parents 0..k (inclusive)
            }
            carry_ne += tot;
        }
        if (tid == 0) { pre_ne[np] = carry_ne; sh.scal[1] = np; }
        __syncthreads();
        int ncom = np;
        if (sorted_phase) {   // the reference stops dividing as soon as the list reaches N (:749-750)
            for (int k = tid; k < np; k += QT_THREADS)
                if (size + pre_g[k] >= N && (k == 0 || size + pre_g[k - 1] < N)) atomicMin(&sh.scal[1], k + 1);
            __syncthreads();
            ncom = sh.scal[1];
        }

        // d. next list = surviving nodes (same relative order) ++ children in creation order; nid = old -> new
        int n_surv;
        if (!sorted_phase || ncom == np) {
            n_surv = n_keep;   // every parent is divided: the survivors were ranked by the scan of step a
            for (int i = tid; i < size; i += QT_THREADS) {
                if (cur[i].count > 1) continue;
                const int pos = flag[i];
                nxt[pos] = cur[i];
                reinterpret_cast<int4*>(nid)[i] = make_int4(pos, pos, pos, pos);
            }
        } else {
            // only the first ncom parents are divided: everything else survives
            for (int i = tid; i < size; i += QT_THREADS) flag[i] = 1;
            __syncthreads();
            for (int k = tid; k < ncom; k += QT_THREADS) flag[par[k]] = 0;
            __syncthreads();
            int c_keep = 0;
            for (int i0 = 0; i0 < size; i0 += QT_THREADS) {
                const int i = i0 + tid;
                const int g = i < size ? flag[i] : 0;
                int tot;
                const int ex = block_excl_scan<QT_THREADS>(g, &tot, sh, sp);
                if (g) {
                    const int pos = c_keep + ex;
                    nxt[pos] = cur[i];
                    reinterpret_cast<int4*>(nid)[i] = make_int4(pos, pos, pos, pos);
                }
                c_keep += tot;
            }
            n_surv = c_keep;
        }
        const int n_children = pre_ne[ncom];
        const int new_size = n_surv + n_children;
        if (new_size > cap) { if (tid == 0) atomicExch(status, ORB_ERR_INTERNAL); break; }
        int my_expand = 0;
        for (int k = tid; k < ncom; k += QT_THREADS) {
            const int pi = par[k];
            const QNode nd = cur[pi];
            const int hx = (nd.x1 - nd.x0 + 1) >> 1, hy = (nd.y1 - nd.y0 + 1) >> 1;
            const int sx = nd.x0 + hx, sy = nd.y0 + hy;
            const int4 c4 = reinterpret_cast<const int4*>(cnt4)[pi];
            const int cc[4] = {c4.x, c4.y, c4.z, c4.w};
            int ids[4];
            int pos = n_surv + pre_ne[k];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int c = cc[q];
                ids[q] = pos;
                if (c > 0) {
                    QNode ch;
                    ch.count = c;
                    ch.x0 = (short)((q & 1) ? sx : nd.x0);
                    ch.x1 = (short)((q & 1) ? nd.x1 : sx);
                    ch.y0 = (short)((q & 2) ? sy : nd.y0);
                    ch.y1 = (short)((q & 2) ? nd.y1 : sy);
                    ch.split = node_split(ch.x0, ch.x1, ch.y0, ch.y1);
                    nxt[pos++] = ch;
                    my_expand += c > 1;
                }
            }
            reinterpret_cast<int4*>(nid)[pi] = make_int4(ids[0], ids[1], ids[2], ids[3]);
        }
        int n_expand;
        block_excl_scan<QT_THREADS>(my_expand, &n_expand, sh, sp);
        { QNode* t = cur; cur = nxt; nxt = t; }   // nxt = the list before this pass (boxes still needed by the sweep)

        // e. termination (reference :688-757)
        if (new_size >= N || new_size == size) finish = true;
        else if (!sorted_phase && new_size + 3 * n_expand > N) sorted_phase = true;
        size = new_size;
        if (!finish)
            for (int i = tid; i < 4 * size; i += QT_THREADS) cnt4[i] = 0;
        __syncthreads();

        // f. key sweep: rename the leaf ids, count the quadrant occupancy of every node that may be divided next
        for (int i0 = 0; i0 < n; i0 += QT_THREADS) {
            const int i = i0 + tid;
            int t = -1;
            if (i < n) {
                const uint32_t k = kbuf[i];
                const int ol = (int)leaf[i];
                const int nl = nid[4 * ol + node_quadrant(nxt[ol], k)];
                leaf[i] = (uint32_t)nl;
                if (!finish) {
                    const QNode nd = cur[nl];
                    if (nd.count > 1) t = 4 * nl + node_quadrant(nd, k);
                }
            }
            if (!finish) warp_count(cnt4, t);
        }
        __syncthreads();
    }

    // ---- 3. best response per node, first in original order among equals (reference :761-779), emitted in list order
    int* best_s = pre_g;
    unsigned long long* best_k = reinterpret_cast<unsigned long long*>(cnt4);
    __syncthreads();
    for (int i = tid; i < size; i += QT_THREADS) { best_s[i] = -1; best_k[i] = ~0ull; }
    __syncthreads();
    for (int i = tid; i < n; i += QT_THREADS) atomicMax(&best_s[leaf[i]], ORB_KEY_S(kbuf[i]));
    __syncthreads();
    const unsigned mw = 0xffffffffu / (unsigned)G.w_cell + 1u, mh = 0xffffffffu / (unsigned)G.h_cell + 1u;   // exact quotients for x < 4096, d < 64
    for (int i = tid; i < n; i += QT_THREADS) {
        const uint32_t k = kbuf[i];
        const int nl = (int)leaf[i];
        if (ORB_KEY_S(k) != best_s[nl]) continue;
        // original order = (cell row, cell column, row in cell, column in cell); cell interiors start 3 pixels into the window
        const int xr = ORB_KEY_X(k) - 3, yr = ORB_KEY_Y(k) - 3;
        const int cj = (int)__umulhi((unsigned)xr, mw), ci = (int)__umulhi((unsigned)yr, mh);
        const unsigned ord = ((unsigned)ci << 20) | ((unsigned)cj << 12) | ((unsigned)(yr - ci * G.h_cell) << 6) | (unsigned)(xr - cj * G.w_cell);
        atomicMin(&best_k[nl], ((unsigned long long)ord << 32) | k);
    }
    __syncthreads();
    for (int i = tid; i < size; i += QT_THREADS) out[size - 1 - i] = (uint32_t)best_k[i];
    if (tid == 0) lvl_cnt[b * ORB_MAX_LEVELS + l] = size;
}

// ------------------------------------------------------------------------------------------
size_t quadtree_smem(int cap, int* p2_out) {
    int p2 = 1;
    while (p2 < cap) p2 <<= 1;
    *p2_out = p2;
    size_t bytes = (size_t)cap * (2 * sizeof(QNode) + 16 + 16 + 4 + 4 + 4 + 4) + 4 + 8;
    bytes = (bytes + 7) & ~(size_t)7;
    return bytes + (size_t)p2 * 8;
}
cudaError_t launch_quadtree(cudaStream_t s, const int* cell_cnt, int* cell_prefix, const uint32_t* stage, uint32_t* keys,
                            uint32_t* lvl_kp, int* lvl_cnt, int* lvl_ncand, int* status, const Plan& P, int batch) {
    int cap = 0;
    for (int l = 0; l < P.nlevels; ++l) cap = max(cap, P.lv[l].node_cap);
    int p2;
    const size_t bytes = quadtree_smem(cap, &p2);
    if (bytes > 200 * 1024) return cudaErrorInvalidConfiguration;
    int nfeat = 0;
    for (int l = 0; l < P.nlevels; ++l) nfeat = max(nfeat, P.lv[l].n_feat);
    const int nt = batch * P.nlevels <= 2 * orb_sm_count() ? 1024 : (nfeat >= 1200 ? 512 : 256);
    auto kern = nt == 1024 ? k_quadtree<1024> : (nt == 512 ? k_quadtree<512> : k_quadtree<256>);
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
    if (e != cudaSuccess) return e;
    dim3 grid(batch, P.nlevels);
    e = orb_launch(orb_pdl(batch), kern, grid, dim3(nt), bytes, s, cell_cnt, cell_prefix, stage, keys, lvl_kp, lvl_cnt, lvl_ncand, status, P, cap, p2);
    return e != cudaSuccess ? e : cudaGetLastError();
}
