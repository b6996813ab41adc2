/* Minimal stand-in for the OpenCV types that appear in the reference's ORBextractor / ORBmatcher /
 * Frame interfaces (include/ORBextractor.h:26, include/ORBmatcher.h:28): cv::Mat (8-bit and float
 * matrices with shared ownership), cv::KeyPoint, cv::Point2f, cv::InputArray / cv::OutputArray.
 *
 * OpenCV C++ is not installed in this image.  In the reference tree, compile with
 * -DORB_B200_WITH_OPENCV: this header then just includes <opencv2/core/core.hpp> and the drop-in
 * classes (include/ORBextractor.h, include/ORBmatcher.h) use the real cv types.  The member names
 * used below (rows, cols, data, step, type(), empty(), create(), ptr<T>(), at<T>(), clone(),
 * row()) are the ones the drop-in classes touch, with OpenCV's meaning. */
#ifndef ORB_B200_CV_H
#define ORB_B200_CV_H

#ifdef ORB_B200_WITH_OPENCV
#include <opencv2/core/core.hpp>
#else

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_32S 4
#define CV_32F 5
#define CV_8UC1 CV_8U
#define CV_32FC1 CV_32F

namespace cv {

typedef unsigned char uchar;

template <class T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<int> Point2i;
typedef Point2i Point;
typedef Point_<float> Point2f;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

/* Same field order and 28-byte layout as OpenCV's cv::KeyPoint == orb_keypoint of the C ABI. */
struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

class Mat {
public:
    int rows, cols;
    uchar* data;
    size_t step;  /* bytes per row */

    Mat() : rows(0), cols(0), data(nullptr), step(0), type_(CV_8U) {}
    Mat(int r, int c, int type) : rows(0), cols(0), data(nullptr), step(0), type_(CV_8U) { create(r, c, type); }
    /* wraps caller memory (no ownership), like cv::Mat(rows, cols, type, void*, step) */
    Mat(int r, int c, int type, void* ext, size_t step_ = 0)
        : rows(r), cols(c), data(static_cast<uchar*>(ext)), step(step_ ? step_ : (size_t)c * esz(type)), type_(type) {}

    static size_t esz(int type) { return type == CV_8U ? 1 : 4; }
    size_t elemSize() const { return esz(type_); }
    int type() const { return type_; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    Size size() const { return Size(cols, rows); }
    bool isContinuous() const { return step == (size_t)cols * elemSize(); }

    void create(int r, int c, int type) {
        if (r == rows && c == cols && type == type_ && data && own_) return;
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        own_ = std::make_shared<std::vector<uchar> >((size_t)r * step);
        data = own_->empty() ? nullptr : own_->data();
    }
    static Mat zeros(int r, int c, int type) { Mat m(r, c, type); if (m.data) std::memset(m.data, 0, (size_t)r * m.step); return m; }
    Mat clone() const {
        Mat m(rows, cols, type_);
        for (int y = 0; y < rows; ++y) std::memcpy(m.data + (size_t)y * m.step, data + (size_t)y * step, (size_t)cols * elemSize());
        return m;
    }
    Mat row(int y) const {
        Mat m; m.rows = 1; m.cols = cols; m.type_ = type_; m.step = step; m.data = data + (size_t)y * step; m.own_ = own_;
        return m;
    }
    void release() { own_.reset(); data = nullptr; rows = cols = 0; step = 0; }

    template <class T> T* ptr(int y = 0) { return reinterpret_cast<T*>(data + (size_t)y * step); }
    template <class T> const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + (size_t)y * step); }
    template <class T> T& at(int y, int x) { return ptr<T>(y)[x]; }
    template <class T> const T& at(int y, int x) const { return ptr<T>(y)[x]; }
    template <class T> T& at(int i) { return rows == 1 ? ptr<T>(0)[i] : ptr<T>(i)[0]; }
    template <class T> const T& at(int i) const { return rows == 1 ? ptr<T>(0)[i] : ptr<T>(i)[0]; }

    /* the drop-in classes take cv::InputArray / cv::OutputArray and call getMat() / create() on them */
    Mat getMat() const { return *this; }
    Mat& getMatRef() { return *this; }

private:
    int type_;
    std::shared_ptr<std::vector<uchar> > own_;
};

typedef const Mat& InputArray;
typedef Mat& OutputArray;

}  // namespace cv

#endif /* ORB_B200_WITH_OPENCV */
#endif /* ORB_B200_CV_H */
