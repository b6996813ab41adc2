// TEST INFRASTRUCTURE — CPU oracle for stereo matching and the ORBmatcher searches (not product
// code; see orb_oracle.h).  Line-by-line restatements over plain arrays.  PINNED against the reference itself:
// src/ORBmatcher.cc, Frame.cc, KeyFrame.cc, MapPoint.cc, Map.cc, KeyFrameDatabase.cc and the vendored DBoW2 are
// compiled verbatim (oracle/Makefile `refmatch`, against oracle/cvshim) into oracle/_ref/ref_match_driver and
// run on real Frame / KeyFrame / MapPoint objects; tests/test_oracle_refmatch.py compares every function below
// with it, bit for bit, on the scenarios the GPU tests use (the reference has no tests or golden vectors of
// its own for this path):
//
//   orc_stereo_match                 <- Frame::ComputeStereoMatches            src/Frame.cc:516-704
//   grid / area query                <- Frame::AssignFeaturesToGrid, PosInGrid,
//                                       GetFeaturesInArea                      src/Frame.cc:261-276, 368-436
//   orc_search_by_projection_mps     <- ORBmatcher::SearchByProjection(F,MPs)  src/ORBmatcher.cc:47-141
//   orc_search_by_projection_last    <- ORBmatcher::SearchByProjection(C,L)    src/ORBmatcher.cc:1460-1619
//   orc_search_by_bow                <- ORBmatcher::SearchByBoW(KF,F)          src/ORBmatcher.cc:164-306
//   orc_search_for_triangulation     <- ORBmatcher::SearchForTriangulation     src/ORBmatcher.cc:718-900, 144-161
//   orc_search_by_bow_keyframes      <- ORBmatcher::SearchByBoW(KF,KF)         src/ORBmatcher.cc:581-716
//   orc_search_by_projection_claiming<- SearchByProjection(F,KF,found,th,dist) src/ORBmatcher.cc:1622-1759 and (KF,Scw,..) :311-427
//   orc_search_projected_best        <- candidate loops of Fuse x2             src/ORBmatcher.cc:903-1207
//   orc_search_by_sim3               <- ORBmatcher::SearchBySim3               src/ORBmatcher.cc:1212-1455
//   orc_search_for_initialization    <- ORBmatcher::SearchForInitialization    src/ORBmatcher.cc:429-578
//   three_maxima                     <- ORBmatcher::ComputeThreeMaxima         src/ORBmatcher.cc:1765-1809
//
// Float expressions of the form x*a + y*b are evaluated as fmaf(x, a, y*b): that is the code GCC
// emits for the reference under its own flags (-O3 -march=native, CMakeLists.txt:10-11) on FMA
// hardware (first product fused, second rounded).  This file is compiled with -ffp-contract=off,
// so every fused operation is explicit.
#include "orb_oracle.h"
#include "../include/orb_b200.h"  // plain view structs of the C ABI (interface types only)
#include <algorithm>
#include <climits>
#include <cmath>
#include <cstring>
#include <utility>
#include <vector>

namespace {

const int TH_HIGH = 100, TH_LOW = 50, HISTO_LENGTH = 30;
const int GRID_COLS = 64, GRID_ROWS = 48;

inline int ddist(const uint8_t* a, const uint8_t* b) { return orc_descriptor_distance(a, b); }

struct Grid {
    std::vector<std::vector<int>> cell;  // [ix * GRID_ROWS + iy]
};

void build_grid(const orb_frame_view& f, Grid& g) {
    g.cell.assign(GRID_COLS * GRID_ROWS, std::vector<int>());
    for (int i = 0; i < f.n; ++i) {
        const int px = (int)std::round((f.x[i] - f.grid.min_x) * f.grid.inv_w);
        const int py = (int)std::round((f.y[i] - f.grid.min_y) * f.grid.inv_h);
        if (px < 0 || px >= GRID_COLS || py < 0 || py >= GRID_ROWS) continue;
        g.cell[px * GRID_ROWS + py].push_back(i);
    }
}

void features_in_area(const orb_frame_view& f, const Grid& g, float x, float y, float r, int minLevel, int maxLevel, std::vector<int>& out) {
    out.clear();
    const int nMinCellX = std::max(0, (int)std::floor((x - f.grid.min_x - r) * f.grid.inv_w));
    if (nMinCellX >= GRID_COLS) return;
    const int nMaxCellX = std::min(GRID_COLS - 1, (int)std::ceil((x - f.grid.min_x + r) * f.grid.inv_w));
    if (nMaxCellX < 0) return;
    const int nMinCellY = std::max(0, (int)std::floor((y - f.grid.min_y - r) * f.grid.inv_h));
    if (nMinCellY >= GRID_ROWS) return;
    const int nMaxCellY = std::min(GRID_ROWS - 1, (int)std::ceil((y - f.grid.min_y + r) * f.grid.inv_h));
    if (nMaxCellY < 0) return;
    const bool bCheckLevels = (minLevel > 0) || (maxLevel >= 0);
    for (int ix = nMinCellX; ix <= nMaxCellX; ++ix)
        for (int iy = nMinCellY; iy <= nMaxCellY; ++iy)
            for (int idx : g.cell[ix * GRID_ROWS + iy]) {
                if (bCheckLevels) {
                    if (f.octave[idx] < minLevel) continue;
                    if (maxLevel >= 0 && f.octave[idx] > maxLevel) continue;
                }
                const float dx = f.x[idx] - x, dy = f.y[idx] - y;
                if (std::fabs(dx) < r && std::fabs(dy) < r) out.push_back(idx);
            }
}

void three_maxima(const std::vector<int>* histo, int L, int& ind1, int& ind2, int& ind3) {
    int max1 = 0, max2 = 0, max3 = 0;
    for (int i = 0; i < L; ++i) {
        const int s = (int)histo[i].size();
        if (s > max1) { max3 = max2; max2 = max1; max1 = s; ind3 = ind2; ind2 = ind1; ind1 = i; }
        else if (s > max2) { max3 = max2; max2 = s; ind3 = ind2; ind2 = i; }
        else if (s > max3) { max3 = s; ind3 = i; }
    }
    if (max2 < 0.1f * (float)max1) { ind2 = -1; ind3 = -1; }
    else if (max3 < 0.1f * (float)max1) ind3 = -1;
}

inline int rot_bin(float a1, float a2) {
    float rot = a1 - a2;
    if (rot < 0.0) rot += 360.0f;
    int bin = (int)std::round(rot * (1.0f / HISTO_LENGTH));
    if (bin == HISTO_LENGTH) bin = 0;
    return bin;
}

}  // namespace

extern "C" {

// Pyramid levels are passed as plain views so that the same code serves the oracle extractor
// (orc_stereo_match below) and the verbatim reference extractor of oracle/_ref.
int orc_stereo_match_levels(int nlevels, const uint8_t* const* lvL, const int32_t* lwA, const int32_t* lhA, const int32_t* lsA,
                            const uint8_t* const* lvR, const int32_t* rwA, const int32_t* rhA, const int32_t* rsA,
                            const orc_keypoint* kpL, const uint8_t* dL, int nL, const orc_keypoint* kpR, const uint8_t* dR, int nR,
                            const float* sf, const float* inv_sf, float mbf, float minD, float maxD, float* uRight, float* depth,
                            int32_t* best_sad) {
    for (int i = 0; i < nL; ++i) { uRight[i] = -1.0f; depth[i] = -1.0f; if (best_sad) best_sad[i] = -1; }
    const int thOrbDist = (TH_HIGH + TH_LOW) / 2;
    const int nRows = lhA[0];
    std::vector<std::vector<int>> vRowIndices(nRows);
    for (int iR = 0; iR < nR; ++iR) {
        const float kpY = kpR[iR].y;
        const float r = 2.0f * sf[kpR[iR].octave];
        const int maxr = (int)std::ceil(kpY + r), minr = (int)std::floor(kpY - r);
        for (int yi = minr; yi <= maxr; ++yi)
            if (yi >= 0 && yi < nRows) vRowIndices[yi].push_back(iR);  // reference indexes unchecked; in range for extractor output
    }
    std::vector<std::pair<int, int>> vDistIdx;
    for (int iL = 0; iL < nL; ++iL) {
        const orc_keypoint& kl = kpL[iL];
        const int levelL = kl.octave;
        const float vL = kl.y, uL = kl.x;
        if ((int)vL < 0 || (int)vL >= nRows) continue;
        const std::vector<int>& cand = vRowIndices[(int)vL];
        if (cand.empty()) continue;
        const float minU = uL - maxD, maxU = uL - minD;
        if (maxU < 0) continue;
        int bestDist = TH_HIGH;
        int bestIdxR = 0;
        for (int iR : cand) {
            const orc_keypoint& kr = kpR[iR];
            if (kr.octave < levelL - 1 || kr.octave > levelL + 1) continue;
            if (kr.x >= minU && kr.x <= maxU) {
                const int dist = ddist(dL + 32 * (size_t)iL, dR + 32 * (size_t)iR);
                if (dist < bestDist) { bestDist = dist; bestIdxR = iR; }
            }
        }
        if (bestDist >= thOrbDist) continue;
        const float uR0 = kpR[bestIdxR].x;
        const float scaleFactor = inv_sf[levelL];
        const float scaleduL = std::round(kl.x * scaleFactor), scaledvL = std::round(kl.y * scaleFactor), scaleduR0 = std::round(uR0 * scaleFactor);
        const int w = 5, LL = 5;
        if (levelL < 0 || levelL >= nlevels) continue;
        const int lw = lwA[levelL], lh = lhA[levelL], ls = lsA[levelL], rw = rwA[levelL], rh = rhA[levelL], rs = rsA[levelL];
        const uint8_t* imL = lvL[levelL];
        const uint8_t* imR = lvR[levelL];
        const float iniu = scaleduR0 + LL - w, endu = scaleduR0 + LL + w + 1;
        if (iniu < 0 || endu >= rw) continue;
        const int su = (int)scaleduL, sv = (int)scaledvL, sr = (int)scaleduR0;
        if (sv - w < 0 || sv + w >= lh || sv + w >= rh || su - w < 0 || su + w >= lw || sr - LL - w < 0 || sr + LL + w >= rw) continue;  // cv::Mat range asserts
        float IL[11][11];
        const float cl = (float)imL[(size_t)sv * ls + su];
        for (int y = 0; y < 11; ++y) for (int x = 0; x < 11; ++x) IL[y][x] = (float)imL[(size_t)(sv - w + y) * ls + (su - w + x)] - cl;
        int bestSad = INT_MAX, bestincR = 0;
        float vDists[11];
        for (int incR = -LL; incR <= LL; ++incR) {
            const float cr = (float)imR[(size_t)sv * rs + (sr + incR)];
            double acc = 0;  // cv::norm(NORM_L1) on CV_32F accumulates in double
            for (int y = 0; y < 11; ++y)
                for (int x = 0; x < 11; ++x) {
                    const float ir = (float)imR[(size_t)(sv - w + y) * rs + (sr + incR - w + x)] - cr;
                    acc += std::fabs(IL[y][x] - ir);
                }
            const float dist = (float)acc;
            if (dist < bestSad) { bestSad = (int)dist; bestincR = incR; }
            vDists[LL + incR] = dist;
        }
        if (bestincR == -LL || bestincR == LL) continue;
        const float dist1 = vDists[LL + bestincR - 1], dist2 = vDists[LL + bestincR], dist3 = vDists[LL + bestincR + 1];
        const float deltaR = (dist1 - dist3) / (2.0f * (dist1 + dist3 - 2.0f * dist2));
        if (deltaR < -1 || deltaR > 1) continue;
        float bestuR = sf[levelL] * ((float)scaleduR0 + (float)bestincR + deltaR);
        float disparity = uL - bestuR;
        if (disparity >= minD && disparity < maxD) {
            if (disparity <= 0) { disparity = 0.01; bestuR = uL - 0.01; }
            depth[iL] = mbf / disparity;
            uRight[iL] = bestuR;
            if (best_sad) best_sad[iL] = bestSad;
            vDistIdx.push_back(std::make_pair(bestSad, iL));
        }
    }
    if (vDistIdx.empty()) return 0;  // reference reads vDistIdx[0] of an empty vector here (undefined)
    std::sort(vDistIdx.begin(), vDistIdx.end());
    const float median = vDistIdx[vDistIdx.size() / 2].first;
    const float thDist = 1.5f * 1.4f * median;
    int kept = (int)vDistIdx.size();
    for (int i = (int)vDistIdx.size() - 1; i >= 0; --i) {
        if (vDistIdx[i].first < thDist) break;
        uRight[vDistIdx[i].second] = -1; depth[vDistIdx[i].second] = -1; --kept;
    }
    return kept;
}

int orc_stereo_match(const orc_extractor* L, const orc_extractor* R, const orc_keypoint* kpL, const uint8_t* dL, int nL,
                     const orc_keypoint* kpR, const uint8_t* dR, int nR, const float* sf, const float* inv_sf, float mbf, float minD,
                     float maxD, float* uRight, float* depth, int32_t* best_sad) {
    const uint8_t *pl[32], *pr[32];
    int32_t lw[32], lh[32], ls[32], rw[32], rh[32], rs[32];
    int n = 0;
    while (n < 32 && (pl[n] = orc_level(L, n, &lw[n], &lh[n], &ls[n])) != nullptr && (pr[n] = orc_level(R, n, &rw[n], &rh[n], &rs[n])) != nullptr) ++n;
    return orc_stereo_match_levels(n, pl, lw, lh, ls, pr, rw, rh, rs, kpL, dL, nL, kpR, dR, nR, sf, inv_sf, mbf, minD, maxD, uRight, depth, best_sad);
}

void orc_match_bruteforce(const uint8_t* a, const uint8_t* b, int n_a, int n_b, int32_t* best_idx, int32_t* best_dist, int32_t* second_dist) {
    for (int i = 0; i < n_a; ++i) {
        int b1 = 256, b2 = 256, bi = -1;
        for (int j = 0; j < n_b; ++j) {
            const int d = ddist(a + 32 * (size_t)i, b + 32 * (size_t)j);
            if (d < b1) { b2 = b1; b1 = d; bi = j; } else if (d < b2) b2 = d;
        }
        best_idx[i] = bi; best_dist[i] = b1; second_dist[i] = b2;
    }
}

void orc_assign_features_to_grid(const float* x, const float* y, int n, const orb_grid_params* gp, int32_t* cell_start, int32_t* cell_items, int32_t* n_items) {
    orb_frame_view f; std::memset(&f, 0, sizeof(f));
    f.n = n; f.x = x; f.y = y; f.grid = *gp;
    Grid g; build_grid(f, g);
    int m = 0;
    for (int c = 0; c < GRID_COLS * GRID_ROWS; ++c) { cell_start[c] = m; for (int i : g.cell[c]) cell_items[m++] = i; }
    cell_start[GRID_COLS * GRID_ROWS] = m;
    *n_items = m;
}

/* Frame::GetFeaturesInArea (src/Frame.cc:368-424) / KeyFrame::GetFeaturesInArea (src/KeyFrame.cc:594-633, min = max = -1)
 * for nq query windows; q_start[nq + 1], q_items[cap]; returns the total or -1 when cap is too small. */
int orc_features_in_area(const orb_frame_view* f, int nq, const float* qx, const float* qy, const float* qr, const int32_t* qmin,
                         const int32_t* qmax, int32_t* q_start, int32_t* q_items, int cap) {
    Grid g; build_grid(*f, g);
    std::vector<int> v;
    int m = 0;
    for (int q = 0; q < nq; ++q) {
        q_start[q] = m;
        features_in_area(*f, g, qx[q], qy[q], qr[q], qmin[q], qmax[q], v);
        if (m + (int)v.size() > cap) return -1;
        for (int i : v) q_items[m++] = i;
    }
    q_start[nq] = m;
    return m;
}

int orc_search_by_projection_mps(const orb_frame_view* f, const orb_mappoint_view* mps, float th, float nn_ratio, const uint8_t* occupied,
                                 int32_t* match) {
    Grid g; build_grid(*f, g);
    std::vector<uint8_t> occ(f->n, 0);
    for (int i = 0; i < f->n; ++i) { match[i] = -1; if (occupied) occ[i] = occupied[i]; }
    int nmatches = 0;
    const bool bFactor = th != 1.0;
    std::vector<int> vIndices;
    for (int iMP = 0; iMP < mps->n; ++iMP) {
        if (!mps->valid[iMP]) continue;
        const int nPredictedLevel = mps->level[iMP];
        float r = ((double)mps->view_cos[iMP] > 0.998) ? 2.5f : 4.0f;
        if (bFactor) r *= th;
        features_in_area(*f, g, mps->proj_x[iMP], mps->proj_y[iMP], r * f->scale_factors[nPredictedLevel], nPredictedLevel - 1, nPredictedLevel, vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* MPdesc = mps->desc + 32 * (size_t)iMP;
        int bestDist = 256, bestLevel = -1, bestDist2 = 256, bestLevel2 = -1, bestIdx = -1;
        for (int idx : vIndices) {
            if (occ[idx]) continue;
            if (f->u_right && f->u_right[idx] > 0) {
                const float er = std::fabs(mps->proj_xr[iMP] - f->u_right[idx]);
                if (er > r * f->scale_factors[nPredictedLevel]) continue;
            }
            const int dist = ddist(MPdesc, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestLevel2 = bestLevel; bestLevel = f->octave[idx]; bestIdx = idx; }
            else if (dist < bestDist2) { bestLevel2 = f->octave[idx]; bestDist2 = dist; }
        }
        if (bestDist <= TH_HIGH) {
            if (bestLevel == bestLevel2 && bestDist > nn_ratio * bestDist2) continue;
            match[bestIdx] = iMP;
            occ[bestIdx] = mps->blocks[iMP];  // F.mvpMapPoints[idx]->Observations() > 0 for later points
            ++nmatches;
        }
    }
    return nmatches;
}

int orc_search_by_projection_last(const orb_frame_view* cur, const orb_lastframe_view* last, const orb_camera* cam, float th, int forward,
                                  int backward, int check_orientation, const uint8_t* occupied, int32_t* match) {
    Grid g; build_grid(*cur, g);
    std::vector<uint8_t> occ(cur->n, 0);
    for (int i = 0; i < cur->n; ++i) { match[i] = -1; if (occupied) occ[i] = occupied[i]; }
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    std::vector<int> vIndices2;
    for (int i = 0; i < last->n; ++i) {
        if (!last->valid[i]) continue;
        const float xc = last->xc[i], yc = last->yc[i];
        const float invzc = 1.0 / last->zc[i];
        if (invzc < 0) continue;
        const float u = fmaf(cam->fx * xc, invzc, cam->cx);
        const float v = fmaf(cam->fy * yc, invzc, cam->cy);
        if (u < cur->grid.min_x || u > cur->grid.max_x) continue;
        if (v < cur->grid.min_y || v > cur->grid.max_y) continue;
        const int nLastOctave = last->octave[i];
        const float radius = th * cur->scale_factors[nLastOctave];
        if (forward) features_in_area(*cur, g, u, v, radius, nLastOctave, -1, vIndices2);
        else if (backward) features_in_area(*cur, g, u, v, radius, 0, nLastOctave, vIndices2);
        else features_in_area(*cur, g, u, v, radius, nLastOctave - 1, nLastOctave + 1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* dMP = last->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            if (occ[i2]) continue;
            if (cur->u_right && cur->u_right[i2] > 0) {
                const float ur = fmaf(-cam->mbf, invzc, u);
                const float er = std::fabs(ur - cur->u_right[i2]);
                if (er > radius) continue;
            }
            const int dist = ddist(dMP, cur->desc + 32 * (size_t)i2);
            if (dist < bestDist) { bestDist = dist; bestIdx2 = i2; }
        }
        if (bestDist <= TH_HIGH) {
            match[bestIdx2] = i;
            occ[bestIdx2] = last->blocks[i];
            ++nmatches;
            if (check_orientation) rotHist[rot_bin(last->angle[i], cur->angle[bestIdx2])].push_back(bestIdx2);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { match[idx] = -1; --nmatches; }
    }
    return nmatches;
}

int orc_search_by_bow(const orb_frame_view* kf, const orb_featvec_view* kfv, const uint8_t* kf_has_mp, const orb_frame_view* f,
                      const orb_featvec_view* fv, float nn_ratio, int check_orientation, int32_t* match_f) {
    for (int i = 0; i < f->n; ++i) match_f[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    int a = 0, b = 0;
    while (a < kfv->n_nodes && b < fv->n_nodes) {
        if (kfv->node_ids[a] == fv->node_ids[b]) {
            for (int p = kfv->start[a]; p < kfv->start[a + 1]; ++p) {
                const unsigned realIdxKF = kfv->items[p];
                if (!kf_has_mp[realIdxKF]) continue;
                const uint8_t* dKF = kf->desc + 32 * (size_t)realIdxKF;
                int bestDist1 = 256, bestIdxF = -1, bestDist2 = 256;
                for (int q = fv->start[b]; q < fv->start[b + 1]; ++q) {
                    const unsigned realIdxF = fv->items[q];
                    if (match_f[realIdxF] >= 0) continue;
                    const int dist = ddist(dKF, f->desc + 32 * (size_t)realIdxF);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdxF = (int)realIdxF; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 <= TH_LOW && (float)bestDist1 < nn_ratio * (float)bestDist2) {
                    match_f[bestIdxF] = (int)realIdxKF;
                    if (check_orientation) rotHist[rot_bin(kf->angle[realIdxKF], f->angle[bestIdxF])].push_back(bestIdxF);
                    ++nmatches;
                }
            }
            ++a; ++b;
        } else if (kfv->node_ids[a] < fv->node_ids[b]) {
            a = (int)(std::lower_bound(kfv->node_ids, kfv->node_ids + kfv->n_nodes, fv->node_ids[b]) - kfv->node_ids);
        } else {
            b = (int)(std::lower_bound(fv->node_ids, fv->node_ids + fv->n_nodes, kfv->node_ids[a]) - fv->node_ids);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match_f[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}

static bool check_dist_epipolar(float x1, float y1, float x2, float y2, int oct2, const float* F12, const float* sigma2) {
    const float a = fmaf(x1, F12[0], y1 * F12[3]) + F12[6];
    const float b = fmaf(x1, F12[1], y1 * F12[4]) + F12[7];
    const float c = fmaf(x1, F12[2], y1 * F12[5]) + F12[8];
    const float num = fmaf(a, x2, b * y2) + c;
    const float den = fmaf(a, a, b * b);
    if (den == 0) return false;
    const float dsqr = num * num / den;
    return dsqr < 3.84 * sigma2[oct2];
}

int orc_search_for_triangulation(const orb_frame_view* kf1, const orb_featvec_view* fv1, const uint8_t* has_mp1, const orb_frame_view* kf2,
                                 const orb_featvec_view* fv2, const uint8_t* has_mp2, const float* F12, float ex, float ey,
                                 const float* sigma2, int only_stereo, int check_orientation, int32_t* match12) {
    for (int i = 0; i < kf1->n; ++i) match12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    int a = 0, b = 0;
    while (a < fv1->n_nodes && b < fv2->n_nodes) {
        if (fv1->node_ids[a] == fv2->node_ids[b]) {
            for (int p = fv1->start[a]; p < fv1->start[a + 1]; ++p) {
                const int idx1 = (int)fv1->items[p];
                if (has_mp1[idx1]) continue;
                const bool bStereo1 = kf1->u_right && kf1->u_right[idx1] >= 0;
                if (only_stereo && !bStereo1) continue;
                const uint8_t* d1 = kf1->desc + 32 * (size_t)idx1;
                int bestDist = TH_LOW, bestIdx2 = -1;
                for (int q = fv2->start[b]; q < fv2->start[b + 1]; ++q) {
                    const int idx2 = (int)fv2->items[q];
                    if (has_mp2[idx2]) continue;  // vbMatched2 is never set in the reference (:744, 796)
                    const bool bStereo2 = kf2->u_right && kf2->u_right[idx2] >= 0;
                    if (only_stereo && !bStereo2) continue;
                    const int dist = ddist(d1, kf2->desc + 32 * (size_t)idx2);
                    if (dist > TH_LOW || dist > bestDist) continue;
                    if (!bStereo1 && !bStereo2) {
                        const float distex = ex - kf2->x[idx2], distey = ey - kf2->y[idx2];
                        if (fmaf(distex, distex, distey * distey) < 100 * kf2->scale_factors[kf2->octave[idx2]]) continue;
                    }
                    if (check_dist_epipolar(kf1->x[idx1], kf1->y[idx1], kf2->x[idx2], kf2->y[idx2], kf2->octave[idx2], F12, sigma2)) {
                        bestIdx2 = idx2; bestDist = dist;
                    }
                }
                if (bestIdx2 >= 0) {
                    match12[idx1] = bestIdx2;
                    ++nmatches;
                    if (check_orientation) rotHist[rot_bin(kf1->angle[idx1], kf2->angle[bestIdx2])].push_back(idx1);
                }
            }
            ++a; ++b;
        } else if (fv1->node_ids[a] < fv2->node_ids[b]) {
            a = (int)(std::lower_bound(fv1->node_ids, fv1->node_ids + fv1->n_nodes, fv2->node_ids[b]) - fv1->node_ids);
        } else {
            b = (int)(std::lower_bound(fv2->node_ids, fv2->node_ids + fv2->n_nodes, fv1->node_ids[a]) - fv2->node_ids);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match12[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}


// ORBmatcher::SearchByBoW(KeyFrame*, KeyFrame*, vpMatches12)   src/ORBmatcher.cc:581-716
int orc_search_by_bow_keyframes(const orb_frame_view* kf1, const orb_featvec_view* fv1, const uint8_t* has_mp1, const orb_frame_view* kf2,
                                const orb_featvec_view* fv2, const uint8_t* has_mp2, float nn_ratio, int check_orientation, int32_t* match12) {
    for (int i = 0; i < kf1->n; ++i) match12[i] = -1;
    std::vector<bool> vbMatched2(kf2->n, false);
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    int a = 0, b = 0;
    while (a < fv1->n_nodes && b < fv2->n_nodes) {
        if (fv1->node_ids[a] == fv2->node_ids[b]) {
            for (int p = fv1->start[a]; p < fv1->start[a + 1]; ++p) {
                const int idx1 = (int)fv1->items[p];
                if (!has_mp1[idx1]) continue;                       // !pMP1 || pMP1->isBad()
                const uint8_t* d1 = kf1->desc + 32 * (size_t)idx1;
                int bestDist1 = 256, bestIdx2 = -1, bestDist2 = 256;
                for (int q = fv2->start[b]; q < fv2->start[b + 1]; ++q) {
                    const int idx2 = (int)fv2->items[q];
                    if (vbMatched2[idx2] || !has_mp2[idx2]) continue;
                    const int dist = ddist(d1, kf2->desc + 32 * (size_t)idx2);
                    if (dist < bestDist1) { bestDist2 = bestDist1; bestDist1 = dist; bestIdx2 = idx2; }
                    else if (dist < bestDist2) bestDist2 = dist;
                }
                if (bestDist1 < TH_LOW) {
                    if ((float)bestDist1 < nn_ratio * (float)bestDist2) {
                        match12[idx1] = bestIdx2;
                        vbMatched2[bestIdx2] = true;
                        if (check_orientation) rotHist[rot_bin(kf1->angle[idx1], kf2->angle[bestIdx2])].push_back(idx1);
                        ++nmatches;
                    }
                }
            }
            ++a; ++b;
        } else if (fv1->node_ids[a] < fv2->node_ids[b]) {
            a = (int)(std::lower_bound(fv1->node_ids, fv1->node_ids + fv1->n_nodes, fv2->node_ids[b]) - fv1->node_ids);
        } else {
            b = (int)(std::lower_bound(fv2->node_ids, fv2->node_ids + fv2->n_nodes, fv1->node_ids[a]) - fv2->node_ids);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx : rotHist[i]) { match12[idx] = -1; --nmatches; }
        }
    }
    return nmatches;
}

// The search loops of
//   SearchByProjection(Frame&, KeyFrame*, sAlreadyFound, th, ORBdist)   src/ORBmatcher.cc:1622-1759 (loop :1666-1712)
//   SearchByProjection(KeyFrame*, Scw, vpPoints, vpMatched, th)         src/ORBmatcher.cc:311-427   (loop :371-420)
// over points the caller has already projected and gated (orb_projected_view).  occupied =
// CurrentFrame.mvpMapPoints[i2] != NULL / vpMatched[idx] != NULL; an accepted keypoint becomes occupied.
int orc_search_by_projection_claiming(const orb_frame_view* f, const orb_projected_view* pts, const uint8_t* occupied, int max_dist,
                                      int check_orientation, int32_t* match, int32_t* point_best_idx) {
    Grid g; build_grid(*f, g);
    std::vector<uint8_t> occ(f->n, 0);
    for (int i = 0; i < f->n; ++i) { match[i] = -1; if (occupied) occ[i] = occupied[i]; }
    std::vector<int> rotHist[HISTO_LENGTH];
    int nmatches = 0;
    std::vector<int> vIndices;
    for (int i = 0; i < pts->n; ++i) {
        if (point_best_idx) point_best_idx[i] = -1;
        if (!pts->valid[i]) continue;
        features_in_area(*f, g, pts->u[i], pts->v[i], pts->radius[i], pts->min_level[i], pts->max_level[i], vIndices);
        if (vIndices.empty()) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            if (occ[idx]) continue;
            const int dist = ddist(dMP, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (bestDist <= max_dist && bestIdx >= 0) {
            match[bestIdx] = i;
            occ[bestIdx] = 1;
            if (point_best_idx) point_best_idx[i] = bestIdx;
            ++nmatches;
            if (check_orientation) rotHist[rot_bin(pts->angle[i], f->angle[bestIdx])].push_back(bestIdx);
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i)
            if (i != ind1 && i != ind2 && i != ind3)
                for (int idx : rotHist[i]) { match[idx] = -1; --nmatches; }
    }
    return nmatches;
}

// The candidate loop shared by Fuse(KF, MapPoints, th) src/ORBmatcher.cc:903-1054 (loop :960-1017, chi-square gate
// when inv_level_sigma2 != NULL), Fuse(KF, Scw, ...) :1056-1207 (loop :1146-1174) and each direction of
// SearchBySim3 :1212-1455 (loops :1299-1331, :1379-1411).
int orc_search_projected_best(const orb_frame_view* f, const orb_projected_view* pts, const float* inv_level_sigma2, int max_dist,
                              int32_t* best_idx, int32_t* best_dist) {
    Grid g; build_grid(*f, g);
    int nfound = 0;
    std::vector<int> vIndices;
    for (int i = 0; i < pts->n; ++i) {
        best_idx[i] = -1;
        if (best_dist) best_dist[i] = 256;
        if (!pts->valid[i]) continue;
        const float u = pts->u[i], v = pts->v[i];
        features_in_area(*f, g, u, v, pts->radius[i], -1, -1, vIndices);     // KeyFrame::GetFeaturesInArea: no level filter
        if (vIndices.empty()) continue;
        const uint8_t* dMP = pts->desc + 32 * (size_t)i;
        int bestDist = 256, bestIdx = -1;
        for (int idx : vIndices) {
            const int kpLevel = f->octave[idx];
            if (kpLevel < pts->min_level[i] || kpLevel > pts->max_level[i]) continue;
            if (inv_level_sigma2) {
                const float kpx = f->x[idx], kpy = f->y[idx];
                const float ex = u - kpx, ey = v - kpy;
                if (f->u_right && f->u_right[idx] >= 0) {
                    const float er = pts->ur[i] - f->u_right[idx];
                    const float e2 = fmaf(er, er, fmaf(ex, ex, ey * ey));
                    if (e2 * inv_level_sigma2[kpLevel] > 7.8) continue;
                } else {
                    const float e2 = fmaf(ex, ex, ey * ey);
                    if (e2 * inv_level_sigma2[kpLevel] > 5.99) continue;
                }
            }
            const int dist = ddist(dMP, f->desc + 32 * (size_t)idx);
            if (dist < bestDist) { bestDist = dist; bestIdx = idx; }
        }
        if (best_dist) best_dist[i] = bestDist;
        if (bestDist <= max_dist && bestIdx >= 0) { best_idx[i] = bestIdx; ++nfound; }
    }
    return nfound;
}

// ORBmatcher::SearchBySim3   src/ORBmatcher.cc:1212-1455 (agreement step :1423-1439)
int orc_search_by_sim3(const orb_frame_view* kf1, const orb_projected_view* pts1_in_kf2, const orb_frame_view* kf2,
                       const orb_projected_view* pts2_in_kf1, int32_t* match12) {
    std::vector<int32_t> vnMatch1(std::max(kf1->n, 1), -1), vnMatch2(std::max(kf2->n, 1), -1);
    orc_search_projected_best(kf2, pts1_in_kf2, nullptr, TH_HIGH, vnMatch1.data(), nullptr);
    orc_search_projected_best(kf1, pts2_in_kf1, nullptr, TH_HIGH, vnMatch2.data(), nullptr);
    int nFound = 0;
    for (int i1 = 0; i1 < kf1->n; ++i1) {
        match12[i1] = -1;
        const int idx2 = vnMatch1[i1];
        if (idx2 >= 0) {
            const int idx1 = vnMatch2[idx2];
            if (idx1 == i1) { match12[i1] = idx2; ++nFound; }
        }
    }
    return nFound;
}

// ORBmatcher::SearchForInitialization   src/ORBmatcher.cc:429-578
int orc_search_for_initialization(const orb_frame_view* f1, const orb_frame_view* f2, const float* prev_x, const float* prev_y, int windowSize,
                                  float nn_ratio, int check_orientation, int32_t* vnMatches12) {
    Grid g; build_grid(*f2, g);
    int nmatches = 0;
    for (int i = 0; i < f1->n; ++i) vnMatches12[i] = -1;
    std::vector<int> rotHist[HISTO_LENGTH];
    std::vector<int> vMatchedDistance(f2->n, INT_MAX), vnMatches21(f2->n, -1);
    std::vector<int> vIndices2;
    for (int i1 = 0; i1 < f1->n; ++i1) {
        const int level1 = f1->octave[i1];
        if (level1 > 0) continue;
        features_in_area(*f2, g, prev_x[i1], prev_y[i1], (float)windowSize, level1, level1, vIndices2);
        if (vIndices2.empty()) continue;
        const uint8_t* d1 = f1->desc + 32 * (size_t)i1;
        int bestDist = INT_MAX, bestDist2 = INT_MAX, bestIdx2 = -1;
        for (int i2 : vIndices2) {
            const int dist = ddist(d1, f2->desc + 32 * (size_t)i2);
            if (vMatchedDistance[i2] <= dist) continue;
            if (dist < bestDist) { bestDist2 = bestDist; bestDist = dist; bestIdx2 = i2; }
            else if (dist < bestDist2) bestDist2 = dist;
        }
        if (bestDist <= TH_LOW) {
            if (bestDist < (float)bestDist2 * nn_ratio) {
                if (vnMatches21[bestIdx2] >= 0) { vnMatches12[vnMatches21[bestIdx2]] = -1; --nmatches; }
                vnMatches12[i1] = bestIdx2;
                vnMatches21[bestIdx2] = i1;
                vMatchedDistance[bestIdx2] = bestDist;
                ++nmatches;
                if (check_orientation) rotHist[rot_bin(f1->angle[i1], f2->angle[bestIdx2])].push_back(i1);
            }
        }
    }
    if (check_orientation) {
        int ind1 = -1, ind2 = -1, ind3 = -1;
        three_maxima(rotHist, HISTO_LENGTH, ind1, ind2, ind3);
        for (int i = 0; i < HISTO_LENGTH; ++i) {
            if (i == ind1 || i == ind2 || i == ind3) continue;
            for (int idx1 : rotHist[i])
                if (vnMatches12[idx1] >= 0) { vnMatches12[idx1] = -1; --nmatches; }
        }
    }
    return nmatches;
}
}  // extern "C"

/* Frame::isInFrustum (src/Frame.cc:303-366) without the PredictScale step (dist is returned for it).  cv::Mat algebra as
 * OpenCV 4.13 evaluates it (R*P+t: float, left to right; norm / dot: double); scalar float expressions with the fused
 * operations GCC emits for the reference's -O3 -march=native build, written out (this file is built with -ffp-contract=off).
 * Pinned against the verbatim reference build: tests/test_oracle_refmatch.py::test_is_in_frustum. */
extern "C" void orc_is_in_frustum(const float* R, const float* t, const float* Ow, const orb_camera* cam, const orb_grid_params* b, float cos_limit, int n,
                                  const float* xw, const float* nrm, const float* mind, const float* maxd, uint8_t* in_view, float* u_out, float* v_out,
                                  float* ur_out, float* cos_out, float* dist_out) {
    for (int i = 0; i < n; ++i) {
        const float X = xw[3 * i], Y = xw[3 * i + 1], Z = xw[3 * i + 2];
        float u = 0.f, v = 0.f, ur = 0.f, vc = 0.f, dist = 0.f; uint8_t ok = 0;
        const float pcx = ((R[0] * X + R[1] * Y) + R[2] * Z) + t[0];
        const float pcy = ((R[3] * X + R[4] * Y) + R[5] * Z) + t[1];
        const float pcz = ((R[6] * X + R[7] * Y) + R[8] * Z) + t[2];
        do {
            if (pcz < 0.0f) break;
            const float invz = 1.0f / pcz;
            u = fmaf(cam->fx * pcx, invz, cam->cx);
            v = fmaf(cam->fy * pcy, invz, cam->cy);
            if (u < b->min_x || u > b->max_x) break;
            if (v < b->min_y || v > b->max_y) break;
            const float px = X - Ow[0], py = Y - Ow[1], pz = Z - Ow[2];
            dist = (float)std::sqrt(((double)px * px + (double)py * py) + (double)pz * pz);
            if (dist < 0.8f * mind[i] || dist > 1.2f * maxd[i]) break;
            const double dot = ((double)px * nrm[3 * i] + (double)py * nrm[3 * i + 1]) + (double)pz * nrm[3 * i + 2];
            vc = (float)(dot / (double)dist);
            if (vc < cos_limit) break;
            ok = 1;
            ur = fmaf(-cam->mbf, invz, u);
        } while (0);
        in_view[i] = ok; u_out[i] = u; v_out[i] = v; ur_out[i] = ur; cos_out[i] = vc; dist_out[i] = dist;
    }
}

/* MapPoint::ComputeDistinctiveDescriptors (src/MapPoint.cc:270-320) on gathered descriptor rows. */
extern "C" void orc_distinctive_descriptors(int n_points, const int32_t* start, const uint8_t* desc, int32_t* best_idx) {
    for (int p = 0; p < n_points; ++p) {
        const int s = start[p], N = start[p + 1] - s;
        int BestMedian = INT_MAX, BestIdx = -1;
        std::vector<std::vector<int>> D(N, std::vector<int>(N, 0));
        for (int i = 0; i < N; ++i)
            for (int j = i + 1; j < N; ++j) D[i][j] = D[j][i] = ddist(desc + (size_t)(s + i) * 32, desc + (size_t)(s + j) * 32);
        for (int i = 0; i < N; ++i) {
            std::vector<int> vDists(D[i]);
            std::sort(vDists.begin(), vDists.end());
            const int median = vDists[0.5 * (N - 1)];
            if (median < BestMedian) { BestMedian = median; BestIdx = i; }
        }
        best_idx[p] = BestIdx;
    }
}
