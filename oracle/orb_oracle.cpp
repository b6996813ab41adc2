// TEST INFRASTRUCTURE — CPU oracle for the ORB extractor (not product code; see orb_oracle.h).
//
// Restates /root/reference/src/ORBextractor.cc with OpenCV 4.13 8-bit arithmetic (prim.hpp):
//   tables            <- ORBextractor::ORBextractor            :428-489
//   build_pyramid     <- ORBextractor::ComputePyramid          :1172-1207
//   level_candidates  <- ComputeKeyPointsOctTree cell loop     :785-865
//   distribute        <- DistributeOctTree + DivideNode        :500-782
//   ic_angle          <- IC_Angle / computeOrientation         :77-113, 491-498
//   brief             <- computeOrbDescriptor                  :119-164 (pattern :168-426)
//   extract           <- ORBextractor::operator()              :1091-1169
//
// Two choices the reference leaves open are FIXED here (and stated in DESIGN.md):
//  * quadtree tie-break: the reference sorts pair<int, ExtractorNode*> (:703), i.e. equal key
//    counts are ordered by heap address.  Canonical rule: later-created node first (what a
//    monotonically growing heap gives).  oracle/_ref runs the verbatim reference under such a heap.
//  * rBRIEF rotation uses the FMA contraction GCC emits under the reference's own flags
//    (-O3 -march=native, CMakeLists.txt:10-11) on FMA hardware: fma(x,b,rn(y*a)), fma(x,a,-rn(y*b)),
//    and cos/sin come from one glibc sincosf call.
#include "orb_oracle.h"
#include "prim.hpp"
#include <list>
#include <utility>

namespace {

using namespace orc;

const int kEdge = 19;       // EDGE_THRESHOLD, reference :74
const int kHalfPatch = 15;  // HALF_PATCH_SIZE, reference :73
const int kPatch = 31;      // PATCH_SIZE, reference :72

const int8_t kPattern[1024] = {
#include "../include/orb_pattern_31.inc"
};

struct Level {
    int w = 0, h = 0;
    std::vector<uint8_t> img, blur;
    bool blurred = false;
    std::vector<FastPt> cand;        // FAST list fed to the quadtree (window-relative coords)
    std::vector<orc_keypoint> kps;   // after quadtree + angle, level coordinates
};

}  // namespace

struct orc_extractor {
    int nfeatures, nlevels, ini_th, min_th;
    double scale_factor;  // the reference stores the float argument in a double member (ORBextractor.h:98)
    std::vector<float> sf, inv_sf, sigma2, inv_sigma2;
    std::vector<int> n_per_level, umax;
    std::vector<Level> lv;
};

namespace {

void make_tables(orc_extractor& e) {
    const int L = e.nlevels;
    e.sf.assign(L, 1.f); e.sigma2.assign(L, 1.f); e.inv_sf.resize(L); e.inv_sigma2.resize(L);
    for (int i = 1; i < L; ++i) {
        e.sf[i] = (float)(e.sf[i - 1] * e.scale_factor);  // float * double -> double -> float (:439)
        e.sigma2[i] = e.sf[i] * e.sf[i];
    }
    for (int i = 0; i < L; ++i) { e.inv_sf[i] = 1.0f / e.sf[i]; e.inv_sigma2[i] = 1.0f / e.sigma2[i]; }

    e.n_per_level.assign(L, 0);
    const float factor = (float)(1.0f / e.scale_factor);
    float desired = e.nfeatures * (1 - factor) / (1 - (float)std::pow((double)factor, (double)L));
    int sum = 0;
    for (int l = 0; l < L - 1; ++l) {
        e.n_per_level[l] = cv_round(desired);
        sum += e.n_per_level[l];
        desired *= factor;
    }
    e.n_per_level[L - 1] = std::max(e.nfeatures - sum, 0);

    // circular patch row extents (:473-488)
    e.umax.assign(kHalfPatch + 1, 0);
    const int vmax = cv_floor(kHalfPatch * std::sqrt(2.f) / 2 + 1);
    const int vmin = cv_ceil(kHalfPatch * std::sqrt(2.f) / 2);
    const double hp2 = kHalfPatch * kHalfPatch;
    for (int v = 0; v <= vmax; ++v) e.umax[v] = cv_round(std::sqrt(hp2 - v * v));
    for (int v = kHalfPatch, v0 = 0; v >= vmin; --v) {
        while (e.umax[v0] == e.umax[v0 + 1]) ++v0;
        e.umax[v] = v0;
        ++v0;
    }
}

void build_pyramid(orc_extractor& e, const uint8_t* img, int w, int h, int stride) {
    e.lv.assign(e.nlevels, Level());
    for (int l = 0; l < e.nlevels; ++l) {
        Level& L = e.lv[l];
        const float s = e.inv_sf[l];
        L.w = cv_round((float)w * s);
        L.h = cv_round((float)h * s);
        L.img.resize((size_t)L.w * L.h);
        if (l == 0) {
            for (int y = 0; y < h; ++y) std::memcpy(&L.img[(size_t)y * w], img + (size_t)y * stride, w);
        } else {
            const Level& P = e.lv[l - 1];
            resize_linear_u8(P.img.data(), P.w, P.h, P.w, L.img.data(), L.w, L.h, L.w);
        }
    }
}

// Per-cell FAST with the ini->min threshold retry; output order: cells row-major, pixels row-major.
void level_candidates(const orc_extractor& e, Level& L) {
    L.cand.clear();
    const int minBX = kEdge - 3, minBY = minBX;
    const int maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
    const float width = (float)(maxBX - minBX), height = (float)(maxBY - minBY);
    const int nCols = (int)(width / 30.f), nRows = (int)(height / 30.f);
    if (nCols <= 0 || nRows <= 0) return;  // reference loops do not execute
    const int wCell = (int)std::ceil(width / nCols), hCell = (int)std::ceil(height / nRows);
    std::vector<FastPt> cell;
    for (int i = 0; i < nRows; ++i) {
        const int iniY = minBY + i * hCell;
        int maxY = iniY + hCell + 6;
        if (iniY >= maxBY - 3) continue;
        if (maxY > maxBY) maxY = maxBY;
        for (int j = 0; j < nCols; ++j) {
            const int iniX = minBX + j * wCell;
            int maxX = iniX + wCell + 6;
            if (iniX >= maxBX - 6) continue;
            if (maxX > maxBX) maxX = maxBX;
            const uint8_t* view = &L.img[(size_t)iniY * L.w + iniX];
            fast9_16(view, maxX - iniX, maxY - iniY, L.w, e.ini_th, cell);
            if (cell.empty()) fast9_16(view, maxX - iniX, maxY - iniY, L.w, e.min_th, cell);
            for (const FastPt& p : cell) L.cand.push_back({p.x + j * wCell, p.y + i * hCell, p.score});
        }
    }
}

struct QNode {
    std::vector<int> keys;  // indices into the candidate list, original order preserved
    int x0, x1, y0, y1;     // UL.x, UR.x, UL.y, BL.y
    bool no_more = false;
    long seq = 0;           // creation order: the canonical stand-in for the heap address
    std::list<QNode>::iterator self;
};

void divide(const QNode& n, const FastPt* pts, QNode out[4]) {
    const int hx = (int)std::ceil((float)(n.x1 - n.x0) / 2), hy = (int)std::ceil((float)(n.y1 - n.y0) / 2);
    const int sx = n.x0 + hx, sy = n.y0 + hy;
    out[0].x0 = n.x0; out[0].x1 = sx;   out[0].y0 = n.y0; out[0].y1 = sy;
    out[1].x0 = sx;   out[1].x1 = n.x1; out[1].y0 = n.y0; out[1].y1 = sy;
    out[2].x0 = n.x0; out[2].x1 = sx;   out[2].y0 = sy;   out[2].y1 = n.y1;
    out[3].x0 = sx;   out[3].x1 = n.x1; out[3].y0 = sy;   out[3].y1 = n.y1;
    for (int k : n.keys) {
        const float x = (float)pts[k].x, y = (float)pts[k].y;
        int q = (x < sx) ? ((y < sy) ? 0 : 2) : ((y < sy) ? 1 : 3);
        out[q].keys.push_back(k);
    }
    for (int q = 0; q < 4; ++q) out[q].no_more = out[q].keys.size() == 1;
}

// DistributeOctTree: returns indices of the kept candidates in the reference's output order.
// stats (optional, test witness): [0] rounds of the sorted phase (:692-757), [1] adjacent equal counts in its sorted lists,
// [2] early breaks (:745) that fell BETWEEN two nodes of equal count, i.e. where the (count, pointer) tie-break decided the result.
std::vector<int> distribute(const FastPt* pts, int n, int minX, int maxX, int minY, int maxY, int N, int* stats = nullptr) {
    std::vector<int> result;
    const int nIni = (int)std::round((float)(maxX - minX) / (maxY - minY));
    const float hX = (float)(maxX - minX) / nIni;
    std::list<QNode> nodes;
    long seq = 0;
    std::vector<QNode*> roots(std::max(nIni, 0));
    for (int i = 0; i < nIni; ++i) {
        QNode r;
        r.x0 = (int)(hX * (float)i); r.x1 = (int)(hX * (float)(i + 1));
        r.y0 = 0; r.y1 = maxY - minY;
        r.seq = seq++;
        nodes.push_back(r);
        roots[i] = &nodes.back();
    }
    if (n > 0 && nIni <= 0) return result;  // reference: undefined (aspect < 0.5); callers reject it
    for (int i = 0; i < n; ++i) roots[(size_t)((float)pts[i].x / hX)]->keys.push_back(i);
    for (auto it = nodes.begin(); it != nodes.end();) {
        if (it->keys.size() == 1) { it->no_more = true; ++it; }
        else if (it->keys.empty()) it = nodes.erase(it);
        else ++it;
    }

    typedef std::pair<int, QNode*> SizedNode;
    auto by_size_then_age = [](const SizedNode& a, const SizedNode& b) {
        return a.first != b.first ? a.first < b.first : a.second->seq < b.second->seq;
    };
    auto push_children = [&](QNode ch[4], std::vector<SizedNode>& expandable, int* n_expand) {
        for (int q = 0; q < 4; ++q) {
            if (ch[q].keys.empty()) continue;
            ch[q].seq = seq++;
            nodes.push_front(ch[q]);
            nodes.front().self = nodes.begin();
            if (nodes.front().keys.size() > 1) {
                if (n_expand) ++*n_expand;
                expandable.push_back(SizedNode((int)nodes.front().keys.size(), &nodes.front()));
            }
        }
    };

    std::vector<SizedNode> expandable;
    bool finish = false;
    while (!finish) {
        const int prev = (int)nodes.size();
        int nToExpand = 0;
        expandable.clear();
        for (auto it = nodes.begin(); it != nodes.end();) {
            if (it->no_more) { ++it; continue; }
            QNode ch[4];
            divide(*it, pts, ch);
            push_children(ch, expandable, &nToExpand);
            it = nodes.erase(it);
        }
        if ((int)nodes.size() >= N || (int)nodes.size() == prev) {
            finish = true;
        } else if ((int)nodes.size() + nToExpand * 3 > N) {
            while (!finish) {
                const int prev2 = (int)nodes.size();
                std::vector<SizedNode> todo = expandable;
                expandable.clear();
                std::sort(todo.begin(), todo.end(), by_size_then_age);
                if (stats) {
                    stats[0]++;
                    for (size_t j = 1; j < todo.size(); ++j) stats[1] += todo[j].first == todo[j - 1].first;
                }
                for (int j = (int)todo.size() - 1; j >= 0; --j) {
                    QNode ch[4];
                    divide(*todo[j].second, pts, ch);
                    push_children(ch, expandable, nullptr);
                    nodes.erase(todo[j].second->self);
                    if ((int)nodes.size() >= N) {
                        if (stats && j > 0 && todo[j - 1].first == todo[j].first) stats[2]++;
                        break;
                    }
                }
                if ((int)nodes.size() >= N || (int)nodes.size() == prev2) finish = true;
            }
        }
    }
    result.reserve(nodes.size());
    for (const QNode& nd : nodes) {
        int best = nd.keys[0];
        for (size_t k = 1; k < nd.keys.size(); ++k)
            if ((float)pts[nd.keys[k]].score > (float)pts[best].score) best = nd.keys[k];
        result.push_back(best);
    }
    return result;
}

float ic_angle(const Level& L, int x, int y, const std::vector<int>& umax) {
    const uint8_t* c = &L.img[(size_t)y * L.w + x];
    int m01 = 0, m10 = 0;
    for (int u = -kHalfPatch; u <= kHalfPatch; ++u) m10 += u * c[u];
    for (int v = 1; v <= kHalfPatch; ++v) {
        int vsum = 0;
        const int d = umax[v];
        for (int u = -d; u <= d; ++u) {
            const int lo = c[u + v * L.w], up = c[u - v * L.w];
            vsum += lo - up;
            m10 += u * (lo + up);
        }
        m01 += v * vsum;
    }
    return fast_atan2((float)m01, (float)m10);
}

void brief(const Level& L, int x, int y, float angle_deg, uint8_t* desc) {
    const float factorPI = (float)(M_PI / 180.f);
    const float angle = angle_deg * factorPI;
    float a, b;
    sincosf(angle, &b, &a);  // a = cos, b = sin
    const uint8_t* c = &L.blur[(size_t)y * L.w + x];
    auto sample = [&](int i) -> int {
        const float px = (float)kPattern[2 * i], py = (float)kPattern[2 * i + 1];
        volatile float ya = py * a, yb = py * b;
        const int r = cv_round(fmaf(px, b, ya));
        const int col = cv_round(fmaf(px, a, -yb));
        return c[r * L.w + col];
    };
    for (int i = 0; i < 32; ++i) {
        int val = 0;
        for (int k = 0; k < 8; ++k) {
            const int t0 = sample(16 * i + 2 * k), t1 = sample(16 * i + 2 * k + 1);
            val |= (t0 < t1) << k;
        }
        desc[i] = (uint8_t)val;
    }
}

}  // namespace

extern "C" {

void orc_resize_linear_u8(const uint8_t* s, int sw, int sh, int ss, uint8_t* d, int dw, int dh, int ds) {
    orc::resize_linear_u8(s, sw, sh, ss, d, dw, dh, ds);
}
void orc_gauss7_u8(const uint8_t* s, int w, int h, int ss, uint8_t* d, int ds) { orc::gauss7_u8(s, w, h, ss, d, ds); }
void orc_remap_linear_u8(const uint8_t* s, int sw, int sh, int ss, const float* mx, const float* my, uint8_t* d, int dw, int dh, int ds) {
    orc::remap_linear_u8(s, sw, sh, ss, mx, my, d, dw, dh, ds);
}
void orc_gray_u8(const uint8_t* s, int w, int h, int ss, int channels, int rgb_order, uint8_t* d, int ds) { orc::gray_u8(s, w, h, ss, channels, rgb_order, d, ds); }
void orc_border101(const uint8_t* s, int w, int h, int ss, uint8_t* d, int ds, int b) { orc::border101(s, w, h, ss, d, ds, b); }
int orc_fast9_16(const uint8_t* img, int w, int h, int stride, int th, int32_t* xys, int cap) {
    std::vector<FastPt> v;
    orc::fast9_16(img, w, h, stride, th, v);
    for (size_t i = 0; i < v.size() && (int)i < cap; ++i) { xys[3 * i] = v[i].x; xys[3 * i + 1] = v[i].y; xys[3 * i + 2] = v[i].score; }
    return (int)v.size();
}
float orc_fast_atan2(float y, float x) { return orc::fast_atan2(y, x); }
void orc_fast_score_map(const uint8_t* img, int w, int h, int stride, int32_t* score) {
    int ofs[16];
    for (int k = 0; k < 16; ++k) ofs[k] = kRingDy[k] * stride + kRingDx[k];
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x)
            score[(size_t)y * w + x] = (x < 3 || y < 3 || x >= w - 3 || y >= h - 3) ? 0 : orc::fast_score(img + (size_t)y * stride + x, ofs);
}
void orc_sincosf(float x, float* s, float* c) { sincosf(x, s, c); }

orc_extractor* orc_extractor_create(int nfeatures, float scale_factor, int nlevels, int ini_th, int min_th) {
    if (nlevels < 1 || nfeatures < 0) return nullptr;
    orc_extractor* e = new orc_extractor();
    e->nfeatures = nfeatures; e->scale_factor = scale_factor; e->nlevels = nlevels; e->ini_th = ini_th; e->min_th = min_th;
    make_tables(*e);
    return e;
}
void orc_extractor_destroy(orc_extractor* e) { delete e; }

void orc_extractor_tables(const orc_extractor* e, float* sf, float* inv_sf, float* s2, float* inv_s2, int32_t* npl, int32_t* umax) {
    for (int i = 0; i < e->nlevels; ++i) {
        if (sf) sf[i] = e->sf[i];
        if (inv_sf) inv_sf[i] = e->inv_sf[i];
        if (s2) s2[i] = e->sigma2[i];
        if (inv_s2) inv_s2[i] = e->inv_sigma2[i];
        if (npl) npl[i] = e->n_per_level[i];
    }
    if (umax) for (int i = 0; i < 16; ++i) umax[i] = e->umax[i];
}

int orc_extract(orc_extractor* e, const uint8_t* img, int w, int h, int stride, orc_keypoint* kps, uint8_t* desc, int cap) {
    if (!img || w <= 0 || h <= 0) return 0;  // _image.empty() -> return (:1095)
    build_pyramid(*e, img, w, h, stride);
    int total = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        level_candidates(*e, L);
        const int minBX = kEdge - 3, maxBX = L.w - kEdge + 3, maxBY = L.h - kEdge + 3;
        std::vector<int> kept = distribute(L.cand.data(), (int)L.cand.size(), minBX, maxBX, minBX, maxBY, e->n_per_level[l]);
        const int patch = (int)(kPatch * e->sf[l]);
        L.kps.clear();
        for (int k : kept) {
            orc_keypoint kp;
            kp.x = (float)L.cand[k].x + minBX; kp.y = (float)L.cand[k].y + minBX;
            kp.size = (float)patch; kp.angle = -1.f; kp.response = (float)L.cand[k].score; kp.octave = l;
            L.kps.push_back(kp);
        }
        total += (int)L.kps.size();
    }
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        for (orc_keypoint& kp : L.kps) kp.angle = ic_angle(L, cv_round(kp.x), cv_round(kp.y), e->umax);
    }
    if (total > cap) return -1;
    int off = 0;
    for (int l = 0; l < e->nlevels; ++l) {
        Level& L = e->lv[l];
        if (L.kps.empty()) continue;
        L.blur.resize(L.img.size());
        gauss7_u8(L.img.data(), L.w, L.h, L.w, L.blur.data(), L.w);
        L.blurred = true;
        const float scale = e->sf[l];
        for (const orc_keypoint& kp : L.kps) {
            brief(L, cv_round(kp.x), cv_round(kp.y), kp.angle, desc + (size_t)off * 32);
            orc_keypoint o = kp;
            if (l != 0) { o.x = kp.x * scale; o.y = kp.y * scale; }
            kps[off++] = o;
        }
    }
    return off;
}

const uint8_t* orc_level(const orc_extractor* e, int l, int* w, int* h, int* stride) {
    if (l < 0 || l >= (int)e->lv.size()) return nullptr;
    *w = e->lv[l].w; *h = e->lv[l].h; *stride = e->lv[l].w;
    return e->lv[l].img.data();
}
const uint8_t* orc_blurred_level(const orc_extractor* e, int l, int* w, int* h, int* stride) {
    if (l < 0 || l >= (int)e->lv.size() || !e->lv[l].blurred) return nullptr;
    *w = e->lv[l].w; *h = e->lv[l].h; *stride = e->lv[l].w;
    return e->lv[l].blur.data();
}
int orc_level_candidates(const orc_extractor* e, int l, int32_t* xys, int cap) {
    const auto& c = e->lv[l].cand;
    for (size_t i = 0; i < c.size() && (int)i < cap; ++i) { xys[3 * i] = c[i].x; xys[3 * i + 1] = c[i].y; xys[3 * i + 2] = c[i].score; }
    return (int)c.size();
}
int orc_level_keypoints(const orc_extractor* e, int l, orc_keypoint* kps, int cap) {
    const auto& k = e->lv[l].kps;
    for (size_t i = 0; i < k.size() && (int)i < cap; ++i) kps[i] = k[i];
    return (int)k.size();
}
int orc_distribute_stats(const int32_t* xys, int n, int minX, int maxX, int minY, int maxY, int N, int32_t* kept_idx, int cap, int32_t* stats) {
    std::vector<FastPt> pts(n);
    for (int i = 0; i < n; ++i) pts[i] = {xys[3 * i], xys[3 * i + 1], xys[3 * i + 2]};
    int st[3] = {0, 0, 0};
    std::vector<int> kept = distribute(pts.data(), n, minX, maxX, minY, maxY, N, st);
    for (size_t i = 0; i < kept.size() && (int)i < cap; ++i) kept_idx[i] = kept[i];
    for (int k = 0; k < 3; ++k) stats[k] = st[k];
    return (int)kept.size();
}
int orc_distribute(const int32_t* xys, int n, int minX, int maxX, int minY, int maxY, int N, int32_t* kept_idx, int cap) {
    std::vector<FastPt> pts(n);
    for (int i = 0; i < n; ++i) pts[i] = {xys[3 * i], xys[3 * i + 1], xys[3 * i + 2]};
    std::vector<int> kept = distribute(pts.data(), n, minX, maxX, minY, maxY, N);
    for (size_t i = 0; i < kept.size() && (int)i < cap; ++i) kept_idx[i] = kept[i];
    return (int)kept.size();
}

int orc_descriptor_distance(const uint8_t* a, const uint8_t* b) {
    // reference src/ORBmatcher.cc:1814-1830: eight 32-bit XOR + SWAR popcount
    int dist = 0;
    for (int i = 0; i < 8; ++i) {
        uint32_t x, y;
        std::memcpy(&x, a + 4 * i, 4); std::memcpy(&y, b + 4 * i, 4);
        uint32_t v = x ^ y;
        v = v - ((v >> 1) & 0x55555555u);
        v = (v & 0x33333333u) + ((v >> 2) & 0x33333333u);
        dist += (int)((((v + (v >> 4)) & 0x0F0F0F0Fu) * 0x01010101u) >> 24);
    }
    return dist;
}

}  // extern "C"
