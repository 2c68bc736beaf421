/*
 * cv_shim.h -- the few OpenCV types the ORBextractor / ORBmatcher interfaces mention (cv::Mat, cv::KeyPoint,
 * cv::Point2f, cv::InputArray, cv::OutputArray), for builds WITHOUT OpenCV headers (this image has none).
 *
 * With real OpenCV, define ORBB200_USE_OPENCV before including the facade headers: they then include
 * <opencv2/core/core.hpp> and this file is not used.  The shim keeps the memory layout the facade relies on:
 * cv::KeyPoint is the 28-byte POD {pt.x, pt.y, size, angle, response, octave, class_id} (== orbb200_keypoint)
 * and cv::Mat is a reference-counted row-major buffer with rows / cols / step / data.
 */
#ifndef ORBB200_CV_SHIM_H
#define ORBB200_CV_SHIM_H

#include <cstddef>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#define CV_8U 0
#define CV_8UC1 0
#define CV_32F 5
#define CV_32FC1 5
#define CV_64F 6
#define CV_64FC1 6
/* OpenCV's encoding: type = depth + ((channels - 1) << 3) */
#define CV_8UC3 16
#define CV_8UC4 24

namespace cv {

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
};
typedef Point_<float> Point2f;
typedef Point_<int> Point2i;
typedef Point2i Point;

struct Size {
    int width, height;
    Size() : width(0), height(0) {}
    Size(int w, int h) : width(w), height(h) {}
};

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
    size_t step; /* bytes between rows */
    unsigned char *data;

    Mat() : rows(0), cols(0), step(0), data(nullptr), type_(CV_8UC1) {}
    Mat(int r, int c, int type) : rows(0), cols(0), step(0), data(nullptr), type_(type) { create(r, c, type); }
    /* user-owned buffer (no copy, no ownership), like cv::Mat(rows, cols, type, data, step) */
    Mat(int r, int c, int type, void *user, size_t step_bytes = 0)
        : rows(r), cols(c), step(step_bytes ? step_bytes : (size_t)c * esz(type)), data((unsigned char *)user), type_(type) {}

    void create(int r, int c, int type)
    {
        if (r == rows && c == cols && type == type_ && data && store_ && step == (size_t)c * esz(type)) return;
        rows = r; cols = c; type_ = type; step = (size_t)c * esz(type);
        store_ = std::shared_ptr<unsigned char>(new unsigned char[step * (size_t)(r > 0 ? r : 0) + 1], std::default_delete<unsigned char[]>());
        data = store_.get();
    }
    void release() { store_.reset(); data = nullptr; rows = cols = 0; step = 0; }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    int type() const { return type_; }
    size_t elemSize() const { return esz(type_); }
    bool isContinuous() const { return step == (size_t)cols * esz(type_); }
    Size size() const { return Size(cols, rows); }
    int channels() const { return (type_ >> 3) + 1; }
    int depth() const { return type_ & 7; }

    template <typename T> T *ptr(int r = 0) { return reinterpret_cast<T *>(data + (size_t)r * step); }
    template <typename T> const T *ptr(int r = 0) const { return reinterpret_cast<const T *>(data + (size_t)r * step); }
    unsigned char *ptr(int r = 0) { return data + (size_t)r * step; }
    const unsigned char *ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T &at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T> const T &at(int r, int c) const { return ptr<T>(r)[c]; }

    /* header sharing the same storage, like cv::Mat::row / rowRange / colRange */
    Mat row(int r) const { return rowRange(r, r + 1); }
    Mat rowRange(int r0, int r1) const
    {
        Mat m(*this);
        m.data = data + (size_t)r0 * step;
        m.rows = r1 - r0;
        return m;
    }
    Mat colRange(int c0, int c1) const
    {
        Mat m(*this);
        m.data = data + (size_t)c0 * esz(type_);
        m.cols = c1 - c0;
        return m;
    }
    Mat clone() const
    {
        Mat m(rows, cols, type_);
        for (int r = 0; r < rows; r++) std::memcpy(m.ptr(r), ptr(r), (size_t)cols * esz(type_));
        return m;
    }
    static Mat zeros(int r, int c, int type)
    {
        Mat m(r, c, type);
        std::memset(m.data, 0, m.step * (size_t)(r > 0 ? r : 0));
        return m;
    }
    const Mat &getMat() const { return *this; } /* so `_image.getMat()` style code keeps compiling */

private:
    static size_t esz(int type) { return (size_t)((type & 7) == CV_64F ? 8 : (type & 7) == CV_32F ? 4 : 1) * (size_t)((type >> 3) + 1); }
    int type_;
    std::shared_ptr<unsigned char> store_;
};

typedef const Mat &InputArray;
typedef Mat &OutputArray;

} // namespace cv

#endif /* ORBB200_CV_SHIM_H */
