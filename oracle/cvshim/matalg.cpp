// TEST INFRASTRUCTURE — numeric kernels behind oracle/cvshim's cv::Mat algebra (see matalg.hpp for
// the OpenCV 4.13 semantics they reproduce).  Compiled with -ffp-contract=off: every operation is a
// separately rounded IEEE operation.
#include <opencv2/core/core.hpp>

namespace cv { namespace shim {

static inline float F(const Mat& m, int r, int c) { return m.at<float>(r, c); }

Mat gemm_plain(const Mat& a, const Mat& b, double alpha) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.cols == b.rows);
    const int len = a.cols, h = a.rows, w = b.cols;
    Mat d(h, w, CV_32F);
    const bool small = len >= 2 && len <= 4 && (len == w || len == h);   // cv::gemm's unrolled float case
    for (int r = 0; r < h; ++r)
        for (int c = 0; c < w; ++c) {
            if (small) {
                float t = F(a, r, 0) * F(b, 0, c);
                for (int k = 1; k < len; ++k) { const float p = F(a, r, k) * F(b, k, c); t = t + p; }
                d.at<float>(r, c) = (float)((double)t * alpha);
            } else {
                double s = 0;
                for (int k = 0; k < len; ++k) s += (double)F(a, r, k) * (double)F(b, k, c);
                d.at<float>(r, c) = (float)(s * alpha);
            }
        }
    return d;
}

Mat gemm_at(const Mat& a, const Mat& b, double alpha) {   // a.t() * b
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows);
    const int len = a.rows, h = a.cols, w = b.cols;
    Mat d(h, w, CV_32F);
    for (int r = 0; r < h; ++r)
        for (int c = 0; c < w; ++c) {
            double s = 0;
            for (int k = 0; k < len; ++k) s += (double)F(a, k, r) * (double)F(b, k, c);
            d.at<float>(r, c) = (float)(s * alpha);
        }
    return d;
}

Mat add(const Mat& a, const Mat& b, float sb) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    Mat d(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) d.at<float>(r, c) = sb > 0 ? F(a, r, c) + F(b, r, c) : F(a, r, c) - F(b, r, c);
    return d;
}

Mat scale(const Mat& a, double alpha) {
    assert(a.type() == CV_32F);
    Mat d(a.rows, a.cols, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) d.at<float>(r, c) = F(a, r, c) * (float)alpha;   // cvtScale 32f -> 32f works in float
    return d;
}

Mat transpose_scaled(const Mat& a, double alpha) {
    assert(a.type() == CV_32F);
    Mat d(a.cols, a.rows, CV_32F);
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) d.at<float>(c, r) = alpha == 1.0 ? F(a, r, c) : F(a, r, c) * (float)alpha;
    return d;
}

double norm_l2(const Mat& a) {
    assert(a.type() == CV_32F);
    double s = 0;
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) s += (double)F(a, r, c) * (double)F(a, r, c);
    return std::sqrt(s);
}

double norm_l1_diff(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    double s = 0;
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) s += std::fabs((double)F(a, r, c) - (double)F(b, r, c));
    return s;
}

double dot(const Mat& a, const Mat& b) {
    assert(a.type() == CV_32F && b.type() == CV_32F && a.rows == b.rows && a.cols == b.cols);
    double s = 0;
    for (int r = 0; r < a.rows; ++r)
        for (int c = 0; c < a.cols; ++c) s += (double)F(a, r, c) * (double)F(b, r, c);
    return s;
}

void copy(const Mat& src, Mat& dst) {
    assert(src.rows == dst.rows && src.cols == dst.cols && src.type() == dst.type());
    for (int y = 0; y < src.rows; ++y) std::memmove(dst.data + (size_t)y * dst.step, src.data + (size_t)y * src.step, (size_t)src.cols * src.elemSize());
}

}}  // namespace cv::shim
