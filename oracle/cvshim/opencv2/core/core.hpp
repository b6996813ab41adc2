// TEST INFRASTRUCTURE — minimal stand-in for the OpenCV C++ API used by the reference's
// src/ORBextractor.cc, so that the reference source compiles VERBATIM where it lies (OpenCV C++ is
// not installed in the authoring container).  Only the container types are re-created here; the
// arithmetic of resize / GaussianBlur / FAST / fastAtan2 / copyMakeBorder is oracle/prim.hpp, which
// is pinned bit-exactly against the real cv2 4.13.0 (tests/test_oracle_primitives.py).
// Semantics that the reference relies on and that are reproduced deliberately:
//   * Mat is a ref-counted header; operator()(Rect), rowRange, colRange return views;
//   * writing through resize / copyMakeBorder / GaussianBlur / `m = Mat::zeros(...)` into a Mat
//     that already has the right size and type reuses its buffer (cv::Mat::create semantics), which
//     is how the reference fills the pyramid ROI and the descriptor rows in place.
#pragma once
#include <algorithm>
#include <cassert>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <climits>
#include <cfloat>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>
#include "../../../prim.hpp"

typedef unsigned char uchar;
#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_64F 6
#define CV_PI 3.1415926535897932384626433832795

static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum { INTER_NEAREST = 0, INTER_LINEAR = 1 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T _x, T _y) : x(_x), y(_y) {}
    template <typename U> Point_(const Point_<U>& p) : x((T)p.x), y((T)p.y) {}
    Point_& operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_& operator+=(const Point_& o) { x += o.x; y += o.y; return *this; }
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size { int width, height; Size() : width(0), height(0) {} Size(int w, int h) : width(w), height(h) {} };
struct Rect { int x, y, width, height; Rect() : x(0), y(0), width(0), height(0) {} Rect(int _x, int _y, int w, int h) : x(_x), y(_y), width(w), height(h) {} };

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float _size, float _angle = -1, float _response = 0, int _octave = 0, int _class_id = -1)
        : pt(x, y), size(_size), angle(_angle), response(_response), octave(_octave), class_id(_class_id) {}
};

struct ZerosExpr { int rows, cols, type; };
class Mat;
struct MatT;   // lazy transpose (keeps cv::gemm's GEMM_1_T flag: that path accumulates in double)

static inline size_t cvshim_esz(int type) { return type == CV_8U ? 1 : type == CV_32F ? 4 : 8; }

class Mat {
    struct Buf { int refs; };  // header of a malloc'd block: [Buf][pixels]; bypasses operator new on purpose
    Buf* buf_;
    int type_;
    void retain() { if (buf_) ++buf_->refs; }
    void drop() { if (buf_ && --buf_->refs == 0) std::free(buf_); buf_ = nullptr; }
public:
    int rows, cols;
    size_t step;
    uchar* data;
    Mat() : buf_(nullptr), type_(CV_8U), rows(0), cols(0), step(0), data(nullptr) {}
    Mat(int r, int c, int type) : buf_(nullptr), type_(CV_8U), rows(0), cols(0), step(0), data(nullptr) { create(r, c, type); }
    Mat(Size s, int type) : buf_(nullptr), type_(CV_8U), rows(0), cols(0), step(0), data(nullptr) { create(s.height, s.width, type); }
    Mat(const Mat& o) : buf_(o.buf_), type_(o.type_), rows(o.rows), cols(o.cols), step(o.step), data(o.data) { retain(); }
    Mat(const ZerosExpr& z) : buf_(nullptr), type_(CV_8U), rows(0), cols(0), step(0), data(nullptr) { *this = z; }
    inline Mat(const MatT& t);
    // external (caller-owned) pixels, no ownership
    Mat(int r, int c, int type, void* ext, size_t ext_step = 0) : buf_(nullptr), type_(type), rows(r), cols(c), step(ext_step ? ext_step : (size_t)c * cvshim_esz(type)), data((uchar*)ext) {}
    ~Mat() { drop(); }
    Mat& operator=(const Mat& o) {
        if (this != &o) { Buf* nb = o.buf_; if (nb) ++nb->refs; drop(); buf_ = nb; type_ = o.type_; rows = o.rows; cols = o.cols; step = o.step; data = o.data; }
        return *this;
    }
    Mat& operator=(const ZerosExpr& z) {
        create(z.rows, z.cols, z.type);
        for (int y = 0; y < rows; ++y) std::memset(data + (size_t)y * step, 0, (size_t)cols * elemSize());
        return *this;
    }
    void create(int r, int c, int type) {
        if (data && rows == r && cols == c && type_ == type) return;  // cv::Mat::create keeps a matching buffer
        drop();
        type_ = type;
        rows = r; cols = c; step = (size_t)c * cvshim_esz(type);
        buf_ = (Buf*)std::malloc(sizeof(Buf) + 16 + (size_t)r * step + 64);
        buf_->refs = 1;
        data = (uchar*)buf_ + sizeof(Buf) + 16 - sizeof(Buf) % 16;
    }
    void release() { drop(); rows = cols = 0; step = 0; data = nullptr; }
    static ZerosExpr zeros(int r, int c, int type) { ZerosExpr z = {r, c, type}; return z; }
    static Mat ones(int r, int c, int type) { assert(type == CV_32F); Mat m(r, c, type); for (int y = 0; y < r; ++y) for (int x = 0; x < c; ++x) m.at<float>(y, x) = 1.f; return m; }
    static Mat eye(int r, int c, int type) { assert(type == CV_32F); Mat m = zeros(r, c, type); for (int i = 0; i < std::min(r, c); ++i) m.at<float>(i, i) = 1.f; return m; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    int depth() const { return type_; }
    int channels() const { return 1; }
    size_t elemSize() const { return cvshim_esz(type_); }
    size_t step1() const { return step / elemSize(); }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }
    Mat operator()(const Rect& r) const {
        assert(r.x >= 0 && r.y >= 0 && r.x + r.width <= cols && r.y + r.height <= rows);
        Mat m(*this); m.data = data + (size_t)r.y * step + r.x * elemSize(); m.rows = r.height; m.cols = r.width; return m;
    }
    Mat rowRange(int a, int b) const { assert(0 <= a && a <= b && b <= rows); Mat m(*this); m.data = data + (size_t)a * step; m.rows = b - a; return m; }
    Mat colRange(int a, int b) const { assert(0 <= a && a <= b && b <= cols); Mat m(*this); m.data = data + a * elemSize(); m.cols = b - a; return m; }
    Mat row(int y) const { return rowRange(y, y + 1); }
    Mat col(int x) const { return colRange(x, x + 1); }
    Mat clone() const {
        Mat m; if (empty()) return m;
        m.create(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    inline void copyTo(const class _OutputArray& dst) const;
    void convertTo(Mat& dst, int type) const {
        assert(type_ == CV_8U && type == CV_32F);
        Mat out(rows, cols, CV_32F);
        for (int y = 0; y < rows; ++y) for (int x = 0; x < cols; ++x) out.at<float>(y, x) = (float)data[(size_t)y * step + x];
        dst = out;
    }
    Mat reshape(int, int = 0) const { std::abort(); }   // only behind the reference's distortion path (src/Frame.cc:468)
    inline MatT t() const;
    template <typename T> T& at(int y, int x) { return *(T*)(data + (size_t)y * step + x * sizeof(T)); }
    template <typename T> const T& at(int y, int x) const { return *(const T*)(data + (size_t)y * step + x * sizeof(T)); }
    // cv::Mat::at(i0): element i0 of a row or column vector (column vectors may be strided views)
    template <typename T> T& at(int i) { return cols == 1 ? at<T>(i, 0) : rows == 1 ? at<T>(0, i) : at<T>(i / cols, i % cols); }
    template <typename T> const T& at(int i) const { return cols == 1 ? at<T>(i, 0) : rows == 1 ? at<T>(0, i) : at<T>(i / cols, i % cols); }
    uchar* ptr(int y = 0) { return data + (size_t)y * step; }
    const uchar* ptr(int y = 0) const { return data + (size_t)y * step; }
    template <typename T> T* ptr(int y = 0) { return (T*)(data + (size_t)y * step); }
    template <typename T> const T* ptr(int y = 0) const { return (const T*)(data + (size_t)y * step); }
    inline double dot(const Mat& o) const;
};

class _InputArray {
    const Mat* m_;
public:
    _InputArray(const Mat& m) : m_(&m) {}
    bool empty() const { return m_->empty(); }
    Mat getMat() const { return *m_; }
};
class _OutputArray {
    Mat* m_;
public:
    _OutputArray(Mat& m) : m_(&m) {}
    _OutputArray(const Mat& m) : m_(const_cast<Mat*>(&m)) {}   // temporaries (ROI headers) as copyTo targets
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void release() const { m_->release(); }
    Mat getMat() const { return *m_; }
};
typedef const _InputArray& InputArray;
typedef const _OutputArray& OutputArray;

inline void resize(const Mat& src, Mat& dst, Size dsize, double = 0, double = 0, int interpolation = INTER_LINEAR) {
    assert(interpolation == INTER_LINEAR);
    dst.create(dsize.height, dsize.width, CV_8UC1);
    orc::resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

inline void copyMakeBorder(const Mat& src, Mat& dst, int top, int bottom, int left, int right, int borderType) {
    assert((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && top == bottom && left == right && top == left);
    const int b = top;
    std::vector<uchar> tmp((size_t)(src.rows + 2 * b) * (src.cols + 2 * b));
    orc::border101(src.data, src.cols, src.rows, (int)src.step, tmp.data(), src.cols + 2 * b, b);
    dst.create(src.rows + 2 * b, src.cols + 2 * b, CV_8UC1);
    for (int y = 0; y < dst.rows; ++y) std::memcpy(dst.data + (size_t)y * dst.step, &tmp[(size_t)y * dst.cols], (size_t)dst.cols);
}

inline void GaussianBlur(const Mat& src, Mat& dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT) {
    assert(ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101);
    dst.create(src.rows, src.cols, CV_8UC1);
    orc::gauss7_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step);
}

inline void FAST(const Mat& img, std::vector<KeyPoint>& kps, int threshold, bool nonmaxSuppression = true) {
    assert(nonmaxSuppression);
    std::vector<orc::FastPt> v;
    orc::fast9_16(img.data, img.cols, img.rows, (int)img.step, threshold, v);
    kps.clear();
    for (size_t i = 0; i < v.size(); ++i) kps.push_back(KeyPoint((float)v[i].x, (float)v[i].y, 7.f, -1, (float)v[i].score));
}

inline float fastAtan2(float y, float x) { return orc::fast_atan2(y, x); }

struct KeyPointsFilter {
    // used only by the reference's dead ComputeKeyPointsOld path (call commented out at :1110)
    static void retainBest(std::vector<KeyPoint>& k, int n) {
        if (n >= 0 && (int)k.size() > n) {
            if (n == 0) { k.clear(); return; }
            std::nth_element(k.begin(), k.begin() + n - 1, k.end(), [](const KeyPoint& a, const KeyPoint& b) { return a.response > b.response; });
            const float amb = k[n - 1].response;
            std::vector<KeyPoint>::iterator e = std::partition(k.begin() + n, k.end(), [amb](const KeyPoint& p) { return p.response >= amb; });
            k.resize(e - k.begin());
        }
    }
};

}  // namespace cv
#include "matalg.hpp"
