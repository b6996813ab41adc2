// TEST INFRASTRUCTURE — cv::Mat algebra on small CV_32F matrices for the verbatim build of the
// reference's ORBmatcher.cc / Frame.cc / KeyFrame.cc / MapPoint.cc (oracle/_ref).  Included at the
// end of cvshim's core.hpp.  Expressions are evaluated eagerly, but with the arithmetic cv::gemm /
// cv::norm / scaled conversions of OpenCV 4.13 perform (probed against cv2 4.13.0 in the authoring
// container, tests/test_oracle_primitives.py::test_gemm_semantics):
//   * A*B (+C) with no transpose flag and inner length 2..4 equal to the result's width or height:
//     float products summed left to right in float, unfused; "+C" is one more float add;
//   * every other product (A.t()*B, longer inner lengths): products accumulated in double,
//     result = (float)(alpha * sum);
//   * alpha*A, A/alpha, A.t()*alpha: a * (float)alpha in float (cvtScale 32f -> 32f), division = times (float)(1.0 / alpha);
//   * norm (L2) and dot: accumulated in double.
// The numeric kernels live in oracle/cvshim/matalg.cpp, compiled with -ffp-contract=off.
#pragma once
#include <string>

namespace cv {

enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };

struct MatT { Mat m; double alpha; };

namespace shim {
Mat gemm_plain(const Mat& a, const Mat& b, double alpha);          // a*b, small-matrix float path when it applies
Mat gemm_at(const Mat& a, const Mat& b, double alpha);             // a.t()*b, double accumulation
Mat add(const Mat& a, const Mat& b, float sb);                      // a + sb*b, sb = +-1
Mat scale(const Mat& a, double alpha);
Mat transpose_scaled(const Mat& a, double alpha);
double norm_l2(const Mat& a);
double norm_l1_diff(const Mat& a, const Mat& b);
double dot(const Mat& a, const Mat& b);
void copy(const Mat& src, Mat& dst);
}  // namespace shim

inline Mat::Mat(const MatT& t) : buf_(nullptr), type_(CV_8U), rows(0), cols(0), step(0), data(nullptr) { *this = shim::transpose_scaled(t.m, t.alpha); }
inline MatT Mat::t() const { MatT r = {*this, 1.0}; return r; }
inline double Mat::dot(const Mat& o) const { return shim::dot(*this, o); }
inline void Mat::copyTo(const _OutputArray& dst) const {
    Mat d = dst.getMat();
    if (!(d.data && d.rows == rows && d.cols == cols && d.type() == type())) { dst.create(rows, cols, type()); d = dst.getMat(); }
    shim::copy(*this, d);
}

inline Mat operator*(const Mat& a, const Mat& b) { return shim::gemm_plain(a, b, 1.0); }
inline Mat operator*(const MatT& a, const Mat& b) { return shim::gemm_at(a.m, b, a.alpha); }
inline Mat operator*(const Mat& a, const MatT& b) { return shim::gemm_plain(a, Mat(b), 1.0); }
inline MatT operator-(const MatT& a) { MatT r = {a.m, -a.alpha}; return r; }
inline MatT operator*(double s, const MatT& a) { MatT r = {a.m, a.alpha * s}; return r; }
inline Mat operator-(const Mat& a) { return shim::scale(a, -1.0); }
inline Mat operator+(const Mat& a, const Mat& b) { return shim::add(a, b, 1.f); }
inline Mat operator-(const Mat& a, const Mat& b) { return shim::add(a, b, -1.f); }
inline Mat operator*(double s, const Mat& a) { return shim::scale(a, s); }
inline Mat operator*(const Mat& a, double s) { return shim::scale(a, s); }
inline Mat operator/(const Mat& a, double s) { return shim::scale(a, 1.0 / s); }

inline double norm(const Mat& a) { return shim::norm_l2(a); }
inline double norm(const Mat& a, const Mat& b, int type) { assert(type == NORM_L1); (void)type; return shim::norm_l1_diff(a, b); }

// cv::Mat_<float>(r, c) << a, b, c ...
template <typename T> class Mat_;
template <typename T> struct MatCommaInit {
    Mat m; int i;
    MatCommaInit& operator,(double v) { m.at<T>(i / m.cols, i % m.cols) = (T)v; ++i; return *this; }
    operator Mat() const { return m; }
};
template <typename T> class Mat_ : public Mat {
public:
    Mat_(int r, int c) : Mat(r, c, sizeof(T) == 4 ? CV_32F : CV_64F) {}
    MatCommaInit<T> operator<<(double v) { MatCommaInit<T> ci = {*this, 0}; ci, v; return ci; }
};
template <typename T> inline Mat operator*(const Mat& a, const MatCommaInit<T>& b) { return a * Mat(b); }

typedef Point_<double> Point2d;
template <typename T> struct Point3_ { T x, y, z; Point3_() : x(0), y(0), z(0) {} Point3_(T a, T b, T c) : x(a), y(b), z(c) {} };
typedef Point3_<float> Point3f;
typedef Point3_<double> Point3d;

// The reference never reaches these with the undistorted pinhole settings used by the oracle
// (src/Frame.cc:454 returns early when mDistCoef[0] == 0); they only have to link.
inline void undistortPoints(const Mat&, Mat&, const Mat&, const Mat&, const Mat&, const Mat&) { std::abort(); }

// cv::FileStorage as far as DBoW2's TemplatedVocabulary names it (YAML save/load is never called:
// the oracle loads vocabularies with the reference's own loadFromTextFile).
class FileNode {
public:
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
    FileNode operator[](int) const { std::abort(); }
    size_t size() const { return 0; }
    operator int() const { return 0; }
    operator float() const { return 0.f; }
    operator double() const { return 0.0; }
    operator std::string() const { return std::string(); }
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char*, int) {}
    FileStorage(const std::string&, int) {}
    bool isOpened() const { return false; }
    void release() {}
    FileNode operator[](const char*) const { std::abort(); }
    FileNode operator[](const std::string&) const { std::abort(); }
};
template <typename T> inline FileStorage& operator<<(FileStorage& fs, const T&) { return fs; }

}  // namespace cv
