// TEST INFRASTRUCTURE — CPU oracle primitives (not product code).
//
// Plain-pointer restatements of the four OpenCV primitives whose arithmetic the ORB front-end
// depends on.  OpenCV is an un-vendored, un-pinned dependency of the reference
// (/root/reference/CMakeLists.txt:31-37); these follow OpenCV 4.13.0 8-bit semantics and are
// pinned against the real cv2 4.13.0 by tests/test_oracle_primitives.py and the committed
// vectors in tests/golden/ (generator: tests/golden/make_golden.py).
//
//   resize_linear_u8   cv::resize(..., INTER_LINEAR)        call site src/ORBextractor.cc:1195
//   gauss7_u8          cv::GaussianBlur(7x7, sigma 2,
//                                       BORDER_REFLECT_101) call site src/ORBextractor.cc:1145
//   fast9_16           cv::FAST(img, kps, th, true)         call site src/ORBextractor.cc:843,848
//   fast_atan2         cv::fastAtan2                        call site src/ORBextractor.cc:112
//   border101          cv::copyMakeBorder(REFLECT_101)      call site src/ORBextractor.cc:1197,1202
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
// use anything under oracle/.
#pragma once
#include <climits>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>

namespace orc {

// cvRound: round-half-to-even under the default rounding mode (SSE cvtss2si / lrint).
static inline int cv_round(float v) { return (int)lrintf(v); }
static inline int cv_round(double v) { return (int)lrint(v); }
static inline int cv_floor(double v) { int i = (int)v; return i - (i > v); }
static inline int cv_ceil(double v) { int i = (int)v; return i + (i < v); }

static inline int reflect101(int p, int len) {
    // BORDER_REFLECT_101: gfedcb|abcdefgh|gfedcba ; len==1 -> 0
    if (len == 1) return 0;
    while (p < 0 || p >= len) {
        if (p < 0) p = -p;
        else p = 2 * (len - 1) - p;
    }
    return p;
}

// ---------------------------------------------------------------------------------------------
// INTER_LINEAR, 8UC1.  11-bit fixed-point coefficients (INTER_RESIZE_COEF_SCALE = 2048);
// horizontal pass keeps int32, vertical pass is ((b0*(r0>>4))>>16 + (b1*(r1>>4))>>16 + 2)>>2.
// x clamps the fractional weight at the borders, y clamps the row index (OpenCV resizeGeneric_).
static inline void resize_linear_u8(const uint8_t* src, int sw, int sh, int sstride,
                                    uint8_t* dst, int dw, int dh, int dstride) {
    const double scale_x = (double)sw / dw, scale_y = (double)sh / dh;
    std::vector<int> xofs(dw), yofs(dh);
    std::vector<short> alpha(2 * dw), beta(2 * dh);
    for (int dx = 0; dx < dw; ++dx) {
        float fx = (float)((dx + 0.5) * scale_x - 0.5);
        int sx = cv_floor(fx);
        fx -= sx;
        if (sx < 0) { fx = 0; sx = 0; }
        if (sx >= sw - 1) { fx = 0; sx = sw - 1; }
        xofs[dx] = sx;
        alpha[2 * dx] = (short)cv_round((1.f - fx) * 2048);
        alpha[2 * dx + 1] = (short)cv_round(fx * 2048);
    }
    for (int dy = 0; dy < dh; ++dy) {
        float fy = (float)((dy + 0.5) * scale_y - 0.5);
        int sy = cv_floor(fy);
        fy -= sy;
        yofs[dy] = sy;
        beta[2 * dy] = (short)cv_round((1.f - fy) * 2048);
        beta[2 * dy + 1] = (short)cv_round(fy * 2048);
    }
    std::vector<int> r0(dw), r1(dw);
    int have0 = INT32_MIN, have1 = INT32_MIN;
    auto hrow = [&](int sy, std::vector<int>& out) {
        const uint8_t* S = src + (size_t)sy * sstride;
        for (int dx = 0; dx < dw; ++dx) {
            int sx = xofs[dx];
            int sx1 = sx + 1 < sw ? sx + 1 : sw - 1;
            out[dx] = S[sx] * alpha[2 * dx] + S[sx1] * alpha[2 * dx + 1];
        }
    };
    for (int dy = 0; dy < dh; ++dy) {
        int sy0 = std::min(std::max(yofs[dy], 0), sh - 1);
        int sy1 = std::min(std::max(yofs[dy] + 1, 0), sh - 1);
        if (have1 == sy0) { r0.swap(r1); have0 = sy0; have1 = INT32_MIN; }
        if (have0 != sy0) { hrow(sy0, r0); have0 = sy0; }
        if (have1 != sy1) {
            if (sy1 == sy0) r1 = r0; else hrow(sy1, r1);
            have1 = sy1;
        }
        const int b0 = beta[2 * dy], b1 = beta[2 * dy + 1];
        uint8_t* D = dst + (size_t)dy * dstride;
        for (int dx = 0; dx < dw; ++dx) {
            int v = (((b0 * (r0[dx] >> 4)) >> 16) + ((b1 * (r1[dx] >> 4)) >> 16) + 2) >> 2;
            D[dx] = (uint8_t)std::min(std::max(v, 0), 255);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// GaussianBlur 7x7, sigma 2, BORDER_REFLECT_101, 8UC1 — OpenCV >= 4 fixed-point path:
// Q0.8 kernel {18,34,48,56,48,34,18}; horizontal Q8.8 in u16, vertical Q16.16 in u32,
// result (v + 2^15) >> 16.  src and dst may alias only if dst == src exactly (a copy is taken).
static const int kGauss7[7] = {18, 34, 48, 56, 48, 34, 18};

static inline void gauss7_u8(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride) {
    std::vector<uint16_t> H((size_t)w * h);
    for (int y = 0; y < h; ++y) {
        const uint8_t* S = src + (size_t)y * sstride;
        uint16_t* Hr = &H[(size_t)y * w];
        for (int x = 0; x < w; ++x) {
            unsigned acc = 0;
            if (x >= 3 && x + 3 < w) {
                for (int i = 0; i < 7; ++i) acc += kGauss7[i] * S[x + i - 3];
            } else {
                for (int i = 0; i < 7; ++i) acc += kGauss7[i] * S[reflect101(x + i - 3, w)];
            }
            Hr[x] = (uint16_t)acc;
        }
    }
    for (int y = 0; y < h; ++y) {
        const uint16_t* rows[7];
        for (int j = 0; j < 7; ++j) rows[j] = &H[(size_t)reflect101(y + j - 3, h) * w];
        uint8_t* D = dst + (size_t)y * dstride;
        for (int x = 0; x < w; ++x) {
            uint32_t acc = 0;
            for (int j = 0; j < 7; ++j) acc += (uint32_t)kGauss7[j] * rows[j][x];
            D[x] = (uint8_t)((acc + 32768u) >> 16);
        }
    }
}

// ---------------------------------------------------------------------------------------------
// copyMakeBorder(BORDER_REFLECT_101): dst is (w+2b) x (h+2b).
static inline void border101(const uint8_t* src, int w, int h, int sstride, uint8_t* dst, int dstride, int b) {
    for (int y = -b; y < h + b; ++y) {
        const uint8_t* S = src + (size_t)reflect101(y, h) * sstride;
        uint8_t* D = dst + (size_t)(y + b) * dstride;
        for (int x = -b; x < w + b; ++x) D[x + b] = S[reflect101(x, w)];
    }
}

// ---------------------------------------------------------------------------------------------
// FAST-9/16.  Ring offsets (dx,dy), k = 0..15, as in cv::makeOffsets(patternSize = 16).
static const int kRingDx[16] = {0, 1, 2, 3, 3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1};
static const int kRingDy[16] = {3, 3, 2, 1, 0, -1, -2, -3, -3, -3, -2, -1, 0, 1, 2, 3};

// Threshold-independent corner score: the largest t for which the pixel is a FAST-9 corner,
// i.e. max over the 16 arcs of 9 contiguous ring pixels of min(|diff|) with one sign, minus 1.
// (cv::cornerScore<16> returns max(t, that) for a pixel that passed the test at t.)  <= 0: never a corner.
static inline int fast_score(const uint8_t* p, const int* ofs) {
    int d[25];
    const int v = p[0];
    for (int k = 0; k < 16; ++k) d[k] = v - p[ofs[k]];
    for (int k = 16; k < 25; ++k) d[k] = d[k - 16];
    int best = 0;  // max over arcs of min(d) (brighter-centre) and of min(-d)
    for (int s = 0; s < 16; ++s) {
        int mn = d[s], mx = d[s];
        for (int k = 1; k < 9; ++k) { mn = std::min(mn, d[s + k]); mx = std::max(mx, d[s + k]); }
        best = std::max(best, std::max(mn, -mx));
    }
    return best - 1;
}

struct FastPt { int x, y, score; };

// cv::FAST(img, kps, threshold, nonmaxSuppression=true) on a w x h view: pixels closer than 3 to
// the view edge are never corners; non-max suppression is strict (>) against the 8 neighbours,
// where a neighbour that is not a corner at this threshold (or lies outside the tested region)
// counts as score 0.  Output order is row-major.  response = score.
static inline void fast9_16(const uint8_t* img, int w, int h, int stride, int threshold, std::vector<FastPt>& out) {
    out.clear();
    if (w < 7 || h < 7) return;
    int ofs[16];
    for (int k = 0; k < 16; ++k) ofs[k] = kRingDy[k] * stride + kRingDx[k];
    std::vector<int> sc((size_t)w * 3, 0);  // three rolling score rows
    auto row = [&](int y) -> int* { return &sc[(size_t)((y % 3 + 3) % 3) * w]; };
    auto compute_row = [&](int y) {
        int* r = row(y);
        std::fill(r, r + w, 0);
        if (y < 3 || y >= h - 3) return;
        const uint8_t* P = img + (size_t)y * stride;
        for (int x = 3; x < w - 3; ++x) {
            const uint8_t* p = P + x;
            const int v = p[0], lo = v - threshold, hi = v + threshold;
            // quick reject: a 9-arc must contain ring pixel 0 or 8, and 4 or 12
            int a = p[ofs[0]], b = p[ofs[8]];
            if (!((a > hi) || (a < lo) || (b > hi) || (b < lo))) continue;
            int s = fast_score(p, ofs);
            if (s >= threshold) r[x] = s;
        }
    };
    compute_row(2); compute_row(3);
    for (int y = 3; y < h - 3; ++y) {
        compute_row(y + 1);
        const int* pp = row(y - 1); const int* cu = row(y); const int* nx = row(y + 1);
        for (int x = 3; x < w - 3; ++x) {
            int s = cu[x];
            if (s == 0) continue;
            if (s > cu[x - 1] && s > cu[x + 1] && s > pp[x - 1] && s > pp[x] && s > pp[x + 1] &&
                s > nx[x - 1] && s > nx[x] && s > nx[x + 1])
                out.push_back({x, y, s});
        }
    }
}

// ---------------------------------------------------------------------------------------------
// cv::fastAtan2(y, x) in degrees, float32, every operation individually rounded (no FMA).
static inline float fast_atan2(float y, float x) {
    const float p1 = 0.9997878412794807f * (float)(180 / M_PI);
    const float p3 = -0.3258083974640975f * (float)(180 / M_PI);
    const float p5 = 0.1555786518463281f * (float)(180 / M_PI);
    const float p7 = -0.04432655554792128f * (float)(180 / M_PI);
    volatile float ax = std::fabs(x), ay = std::fabs(y);
    volatile float a, c, c2, t;
    if (ax >= ay) {
        t = ax + (float)2.2204460492503131e-16;
        c = ay / t;
        c2 = c * c;
        t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1;
        a = t * c;
    } else {
        t = ay + (float)2.2204460492503131e-16;
        c = ax / t;
        c2 = c * c;
        t = p7 * c2; t = t + p5; t = t * c2; t = t + p3; t = t * c2; t = t + p1;
        t = t * c;
        a = 90.f - t;
    }
    if (x < 0) a = 180.f - a;
    if (y < 0) a = 360.f - a;
    return a;
}


// cv::remap(src, dst, map1 = x (CV_32FC1), map2 = y (CV_32FC1), INTER_LINEAR, BORDER_CONSTANT 0) on 8UC1, as the EuRoC
// stereo example rectifies its images (reference Examples/Stereo/stereo_euroc.cc:97-98,136-137).  OpenCV >= 3 fixed point:
// coordinates to 1/32 px (cvRound(x * 32), half to even), integer part saturated to short, bilinear weights
// (32-fy)(32-fx)*32 ... summing to 2^15, result (sum + 2^14) >> 15; pixels outside the source read 0.
static inline void remap_linear_u8(const uint8_t* src, int sw, int sh, int sstride, const float* mapx, const float* mapy, uint8_t* dst, int dw,
                                   int dh, int dstride) {
    for (int y = 0; y < dh; ++y)
        for (int x = 0; x < dw; ++x) {
            const float fxs = mapx[(size_t)y * dw + x] * 32.f, fys = mapy[(size_t)y * dw + x] * 32.f;
            // cvRound saturates through cvtss2si: out-of-range and NaN give INT_MIN
            const int sx = (fxs >= -2147483648.f && fxs < 2147483648.f) ? (int)lrintf(fxs) : INT_MIN;
            const int sy = (fys >= -2147483648.f && fys < 2147483648.f) ? (int)lrintf(fys) : INT_MIN;
            const int fx = sx & 31, fy = sy & 31;
            const int ix = std::max(-32768, std::min(32767, sx >> 5)), iy = std::max(-32768, std::min(32767, sy >> 5));
            const int w00 = (32 - fy) * (32 - fx) * 32, w01 = (32 - fy) * fx * 32, w10 = fy * (32 - fx) * 32, w11 = fy * fx * 32;
            auto px = [&](int yy, int xx) -> int { return (xx >= 0 && xx < sw && yy >= 0 && yy < sh) ? src[(size_t)yy * sstride + xx] : 0; };
            const int v = px(iy, ix) * w00 + px(iy, ix + 1) * w01 + px(iy + 1, ix) * w10 + px(iy + 1, ix + 1) * w11;
            dst[(size_t)y * dstride + x] = (uint8_t)((v + (1 << 14)) >> 15);
        }
}

// cv::cvtColor(src, dst, CV_RGB2GRAY / CV_BGR2GRAY / CV_RGBA2GRAY / CV_BGRA2GRAY) on 8U (reference src/Tracking.cc:174-199):
// OpenCV 4 fixed point, 15 bits: (R*9798 + G*19235 + B*3735 + 2^14) >> 15.  rgb_order != 0: channel 0 is R.
static inline void gray_u8(const uint8_t* src, int w, int h, int sstride, int channels, int rgb_order, uint8_t* dst, int dstride) {
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            const uint8_t* p = src + (size_t)y * sstride + (size_t)x * channels;
            const int r = rgb_order ? p[0] : p[2], g = p[1], b = rgb_order ? p[2] : p[0];
            dst[(size_t)y * dstride + x] = (uint8_t)((r * 9798 + g * 19235 + b * 3735 + (1 << 14)) >> 15);
        }
}
}  // namespace orc
