/*
 * Stand-in for <opencv2/core/core.hpp> -- TEST INFRASTRUCTURE.
 *
 * This image has no OpenCV C++ headers or libraries.  oracle/Makefile compiles the reference's OWN, UNMODIFIED sources
 * where they lie (/root/reference/src/{ORBextractor,ORBmatcher,Frame,KeyFrame,MapPoint,Map}.cc and Thirdparty/DBoW2)
 * into oracle/_ref/liborbslam_ref.so and oracle/_ref/libdbow2_ref.so against this header tree.  It provides exactly
 * the OpenCV surface those files use:
 *   - the container types (Mat with ROI headers sharing storage, Mat_, Point_, Size_, Rect_, KeyPoint, the
 *     InputArray / OutputArray proxies, cvRound / cvFloor / cvCeil),
 *   - CV_32F matrix expressions with OpenCV's evaluation rules: MatExpr folding (A*B + C is ONE gemm, -A.t()*b is a
 *     gemm with alpha = -1 and GEMM_1_T, M/s is convertTo with the float factor 1/s), cv::gemm's small-matrix branch
 *     (flags == 0 and 2 <= len <= 4: products accumulate left to right in FLOAT) against its general branch (double
 *     accumulation), Mat::dot / cv::norm accumulating in double.  These rules are pinned against cv2 4.13.0
 *     (cv2.gemm / cv2.norm / cv2.transpose) by tests/test_ref_pin.py through oracle/orbslam_ref.cpp's probes,
 *   - the five image-library calls of ORBextractor.cc (declared in imgproc / features2d) implemented over the
 *     cv2-pinned primitives of oracle/orb_ref.c.
 * Nothing under orb-slam2_b200/ or include/ includes this tree, except that the facade headers see it as "OpenCV"
 * when a TEST driver defines ORBB200_USE_OPENCV (tests/cpp/ref_match.cpp).
 */
#ifndef ORBB200_ORACLE_OPENCV_CORE_STANDIN
#define ORBB200_ORACLE_OPENCV_CORE_STANDIN
#include <float.h>
#include <limits.h> /* the real cvdef.h includes it; Frame.cc / MapPoint.cc use INT_MAX without including it */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <algorithm>
#include <cassert>
#include <cmath>
#include <iostream>
#include <limits>
#include <memory>
#include <sstream> /* the real header drags these in and the reference sources rely on it */
#include <string>
#include <vector>

typedef unsigned char uchar;
typedef unsigned short ushort;

#define CV_8U 0
#define CV_8S 1
#define CV_16U 2
#define CV_16S 3
#define CV_32S 4
#define CV_32F 5
#define CV_64F 6
#define CV_CN_SHIFT 3
#define CV_MAT_DEPTH(t) ((t) & 7)
#define CV_MAT_CN(t) ((((t) >> CV_CN_SHIFT) & 63) + 1)
#define CV_MAKETYPE(d, cn) (CV_MAT_DEPTH(d) + (((cn) - 1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_8UC4 CV_MAKETYPE(CV_8U, 4)
#define CV_32SC1 CV_MAKETYPE(CV_32S, 1)
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC2 CV_MAKETYPE(CV_32F, 2)
#define CV_64FC1 CV_MAKETYPE(CV_64F, 1)
#define CV_PI 3.1415926535897932384626433832795

/* OpenCV rounds to nearest-even (cvtsd2si / lrint under the default rounding mode) */
static inline int cvRound(double v) { return (int)lrint(v); }
static inline int cvRound(float v) { return (int)lrintf(v); }
static inline int cvRound(int v) { return v; }
static inline int cvFloor(double v) { int i = (int)v; return i - (i > v); }
static inline int cvFloor(float v) { int i = (int)v; return i - (i > v); }
static inline int cvCeil(double v) { int i = (int)v; return i + (i < v); }
static inline int cvCeil(float v) { int i = (int)v; return i + (i < v); }

namespace cv {

enum { NORM_INF = 1, NORM_L1 = 2, NORM_L2 = 4 };
enum { GEMM_1_T = 1, GEMM_2_T = 2, GEMM_3_T = 4 };
enum { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_REFLECT = 2, BORDER_WRAP = 3, BORDER_REFLECT_101 = 4,
       BORDER_REFLECT101 = 4, BORDER_DEFAULT = 4, BORDER_ISOLATED = 16 };

template <typename T> struct Point_ {
    T x, y;
    Point_() : x(0), y(0) {}
    Point_(T x_, T y_) : x(x_), y(y_) {}
    template <typename U> Point_(const Point_<U> &p) : x((T)p.x), y((T)p.y) {}
    Point_ &operator*=(float s) { x = (T)(x * s); y = (T)(y * s); return *this; }   /* float pt: x*s in float, as OpenCV */
    Point_ &operator*=(double s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_ &operator*=(int s) { x = (T)(x * s); y = (T)(y * s); return *this; }
    Point_ &operator+=(const Point_ &o) { x += o.x; y += o.y; return *this; }
    Point_ &operator-=(const Point_ &o) { x -= o.x; y -= o.y; return *this; }
};
template <typename T> static inline Point_<T> operator+(const Point_<T> &a, const Point_<T> &b) { return Point_<T>(a.x + b.x, a.y + b.y); }
template <typename T> static inline Point_<T> operator-(const Point_<T> &a, const Point_<T> &b) { return Point_<T>(a.x - b.x, a.y - b.y); }
template <typename T> static inline bool operator==(const Point_<T> &a, const Point_<T> &b) { return a.x == b.x && a.y == b.y; }
typedef Point_<int> Point2i;
typedef Point_<float> Point2f;
typedef Point_<double> Point2d;
typedef Point2i Point;

template <typename T> struct Size_ {
    T width, height;
    Size_() : width(0), height(0) {}
    Size_(T w, T h) : width(w), height(h) {}
    T area() const { return width * height; }
};
typedef Size_<int> Size2i;
typedef Size2i Size;
static inline bool operator==(const Size &a, const Size &b) { return a.width == b.width && a.height == b.height; }
static inline bool operator!=(const Size &a, const Size &b) { return !(a == b); }

template <typename T> struct Rect_ {
    T x, y, width, height;
    Rect_() : x(0), y(0), width(0), height(0) {}
    Rect_(T x_, T y_, T w, T h) : x(x_), y(y_), width(w), height(h) {}
};
typedef Rect_<int> Rect;

template <typename T> struct Scalar_ {
    T val[4];
    Scalar_() { val[0] = val[1] = val[2] = val[3] = 0; }
    Scalar_(T v0, T v1 = 0, T v2 = 0, T v3 = 0) { val[0] = v0; val[1] = v1; val[2] = v2; val[3] = v3; }
    T operator[](int i) const { return val[i]; }
};
typedef Scalar_<double> Scalar;

struct KeyPoint {
    Point2f pt;
    float size, angle, response;
    int octave, class_id;
    KeyPoint() : pt(0, 0), size(0), angle(-1), response(0), octave(0), class_id(-1) {}
    KeyPoint(Point2f p, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(p), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
    KeyPoint(float x, float y, float size_, float angle_ = -1, float response_ = 0, int octave_ = 0, int class_id_ = -1)
        : pt(x, y), size(size_), angle(angle_), response(response_), octave(octave_), class_id(class_id_) {}
};

template <typename T> struct DataType;
template <> struct DataType<uchar> { enum { type = CV_8UC1 }; };
template <> struct DataType<int> { enum { type = CV_32SC1 }; };
template <> struct DataType<float> { enum { type = CV_32FC1 }; };
template <> struct DataType<double> { enum { type = CV_64FC1 }; };

class Mat;
class MatExpr;
class _InputArray;
class _OutputArray;
typedef const _InputArray &InputArray;
typedef const _OutputArray &OutputArray;
typedef const _OutputArray &InputOutputArray;
template <typename T> class Mat_;
template <typename T> class MatCommaInitializer_;

class Mat {
public:
    enum { AUTO_STEP = 0 };
    int flags, dims, rows, cols;
    size_t step; /* bytes between rows */
    uchar *data;

    Mat() : flags(0), dims(2), rows(0), cols(0), step(0), data(0) {}
    Mat(int r, int c, int type) : flags(0), dims(2), rows(0), cols(0), step(0), data(0) { create(r, c, type); }
    Mat(Size sz, int type) : flags(0), dims(2), rows(0), cols(0), step(0), data(0) { create(sz.height, sz.width, type); }
    Mat(int r, int c, int type, const Scalar &s) : flags(0), dims(2), rows(0), cols(0), step(0), data(0) { create(r, c, type); setTo(s.val[0]); }
    /* user-owned buffer: no copy, no ownership */
    Mat(int r, int c, int type, void *user, size_t step_bytes = AUTO_STEP)
        : flags(type), dims(2), rows(r), cols(c), step(step_bytes ? step_bytes : (size_t)c * esz(type)), data((uchar *)user) {}
    Mat(Size sz, int type, void *user, size_t step_bytes = AUTO_STEP)
        : flags(type), dims(2), rows(sz.height), cols(sz.width), step(step_bytes ? step_bytes : (size_t)sz.width * esz(type)), data((uchar *)user) {}
    Mat(const Mat &m, const Rect &roi) { *this = m; data = m.data + (size_t)roi.y * m.step + (size_t)roi.x * m.elemSize(); rows = roi.height; cols = roi.width; }
    Mat(const MatExpr &e);
    template <typename T> Mat(const MatCommaInitializer_<T> &ci);
    Mat &operator=(const MatExpr &e);

    /* cv::Mat::create: nothing happens when shape and type already agree (also for a ROI header: ComputePyramid's
       resize / copyMakeBorder write in place through that rule, ORBextractor.cc:1363-1370) */
    void create(int r, int c, int type)
    {
        type &= 0xFFF;
        if (data && r == rows && c == cols && type == this->type()) return;
        flags = type; rows = r; cols = c; step = (size_t)c * esz(type);
        store_ = std::shared_ptr<uchar>(new uchar[step * (size_t)(r > 0 ? r : 0) + 64], std::default_delete<uchar[]>());
        data = store_.get();
    }
    void create(Size sz, int type) { create(sz.height, sz.width, type); }
    void release() { store_.reset(); data = 0; rows = cols = 0; step = 0; }
    bool empty() const { return data == 0 || rows == 0 || cols == 0; }
    int type() const { return flags & 0xFFF; }
    int depth() const { return CV_MAT_DEPTH(flags); }
    int channels() const { return CV_MAT_CN(flags); }
    size_t elemSize() const { return esz(flags); }
    size_t elemSize1() const { return esz1(flags); }
    size_t step1() const { return step / elemSize1(); }
    size_t total() const { return (size_t)rows * cols; }
    bool isContinuous() const { return rows <= 1 || step == (size_t)cols * elemSize(); }
    Size size() const { return Size(cols, rows); }

    uchar *ptr(int r = 0) { return data + (size_t)r * step; }
    const uchar *ptr(int r = 0) const { return data + (size_t)r * step; }
    template <typename T> T *ptr(int r = 0) { return reinterpret_cast<T *>(data + (size_t)r * step); }
    template <typename T> const T *ptr(int r = 0) const { return reinterpret_cast<const T *>(data + (size_t)r * step); }
    template <typename T> T &at(int r, int c) { return ptr<T>(r)[c]; }
    template <typename T> const T &at(int r, int c) const { return ptr<T>(r)[c]; }
    /* single index: element i of a continuous matrix or of a single row / column (cv::Mat::at(int)) */
    template <typename T> T &at(int i) { return const_cast<T &>(static_cast<const Mat *>(this)->at<T>(i)); }
    template <typename T> const T &at(int i) const
    {
        if (isContinuous() || rows == 1) return reinterpret_cast<const T *>(data)[i];
        if (cols == 1) return *reinterpret_cast<const T *>(data + (size_t)i * step);
        return ptr<T>(i / cols)[i % cols];
    }

    Mat row(int r) const { return rowRange(r, r + 1); }
    Mat col(int c) const { return colRange(c, c + 1); }
    Mat rowRange(int r0, int r1) const { Mat m(*this); m.data = data + (size_t)r0 * step; m.rows = r1 - r0; return m; }
    Mat colRange(int c0, int c1) const { Mat m(*this); m.data = data + (size_t)c0 * elemSize(); m.cols = c1 - c0; return m; }
    Mat operator()(const Rect &roi) const { return Mat(*this, roi); }

    Mat clone() const { Mat m; copyTo(m); return m; }
    void copyTo(OutputArray dst) const;
    void copyTo(Mat &m) const
    {
        if (empty()) { m.release(); return; }
        Mat src(*this); /* keeps the storage alive when m aliases *this */
        m.create(src.rows, src.cols, src.type());
        if (m.data == src.data && m.step == src.step) return;
        const size_t rb = (size_t)src.cols * src.elemSize();
        for (int r = 0; r < src.rows; r++) memmove(m.ptr(r), src.ptr(r), rb);
    }
    /* cv::Mat::convertTo for the depth pairs the reference uses: 8U -> 32F (Frame.cc:675, 692) and 32F -> 32F with a
       scale (MatExpr evaluation).  OpenCV narrows alpha / beta to float for a 32F destination: dst = src*(float)a + (float)b */
    void convertTo(Mat &m, int rtype, double alpha = 1, double beta = 0) const
    {
        Mat src(*this);
        if (rtype < 0) rtype = src.type();
        rtype = CV_MAKETYPE(CV_MAT_DEPTH(rtype), src.channels());
        Mat dst;
        if (m.data && m.rows == src.rows && m.cols == src.cols && m.type() == rtype) dst = m; else dst.create(src.rows, src.cols, rtype);
        const int n = src.cols * src.channels();
        const float a = (float)alpha, b = (float)beta;
        const bool plain = alpha == 1 && beta == 0;
        for (int r = 0; r < src.rows; r++) {
            if (src.depth() == CV_8U && dst.depth() == CV_32F) {
                const uchar *s = src.ptr(r); float *d = dst.ptr<float>(r);
                for (int i = 0; i < n; i++) d[i] = plain ? (float)s[i] : (float)s[i] * a + b;
            } else if (src.depth() == CV_32F && dst.depth() == CV_32F) {
                const float *s = src.ptr<float>(r); float *d = dst.ptr<float>(r);
                for (int i = 0; i < n; i++) d[i] = plain ? s[i] : s[i] * a + b;
            } else if (src.depth() == dst.depth() && plain) {
                memmove(dst.ptr(r), src.ptr(r), (size_t)src.cols * src.elemSize());
            } else {
                std::cerr << "cv shim: convertTo " << src.depth() << " -> " << dst.depth() << " not provided" << std::endl;
                abort();
            }
        }
        m = dst;
    }
    void convertTo(OutputArray dst, int rtype, double alpha = 1, double beta = 0) const;
    /* same data seen with another channel count (Frame.cc:491, 503: N x 2 x 1ch <-> N x 1 x 2ch) */
    Mat reshape(int cn, int new_rows = 0) const
    {
        assert(isContinuous() && new_rows == 0);
        Mat m(*this);
        const int total_w = cols * channels();
        m.cols = total_w / cn;
        m.flags = CV_MAKETYPE(depth(), cn);
        return m;
    }
    Mat &setTo(double v)
    {
        for (int r = 0; r < rows; r++)
            for (int c = 0; c < cols * channels(); c++) {
                switch (depth()) {
                case CV_8U: ptr(r)[c] = (uchar)v; break;
                case CV_32S: ptr<int>(r)[c] = (int)v; break;
                case CV_32F: ptr<float>(r)[c] = (float)v; break;
                case CV_64F: ptr<double>(r)[c] = v; break;
                default: abort();
                }
            }
        return *this;
    }
    Mat &operator=(const Scalar &s) { return setTo(s.val[0]); }

    MatExpr t() const;
    MatExpr inv() const; /* declared for completeness; never evaluated on this path (aborts) */
    double dot(InputArray m) const;

    static MatExpr zeros(int r, int c, int type);
    static MatExpr zeros(Size sz, int type);
    static MatExpr ones(int r, int c, int type);
    static MatExpr eye(int r, int c, int type);

private:
    static size_t esz1(int type) { static const size_t s[8] = {1, 1, 2, 2, 4, 4, 8, 2}; return s[type & 7]; }
    static size_t esz(int type) { return esz1(type) * (size_t)CV_MAT_CN(type); }
    std::shared_ptr<uchar> store_;
};

/* ---- matrix expressions with OpenCV's folding rules (modules/core/src/matop.cpp) --------------------------------- */
void gemm(InputArray a, InputArray b, double alpha, InputArray c, double beta, OutputArray d, int flags = 0);
void transpose(InputArray src, OutputArray dst);

class MatExpr {
public:
    enum Kind { IDENT, T, ADDEX, GEMM };
    Kind kind;
    Mat a, b, c;
    double alpha, beta;
    int gflags;
    MatExpr() : kind(IDENT), alpha(1), beta(0), gflags(0) {}
    MatExpr(const Mat &m) : kind(IDENT), a(m), alpha(1), beta(0), gflags(0) {}
    /* ADDEX: alpha*a + beta*b (b may be empty);  T: alpha * a^T;  GEMM: alpha * op(a) * op(b) + beta * c */
    static MatExpr make(Kind k, const Mat &a, const Mat &b, const Mat &c, double alpha, double beta, int gflags = 0)
    {
        MatExpr e; e.kind = k; e.a = a; e.b = b; e.c = c; e.alpha = alpha; e.beta = beta; e.gflags = gflags; return e;
    }
    bool isScaled() const { return kind == ADDEX && b.empty() && beta == 0; } /* alpha * a */
    operator Mat() const { Mat m; eval(m); return m; }
    void eval(Mat &m) const;
    /* members the reference calls on expressions */
    MatExpr t() const { return Mat(*this).t(); }
    double dot(const Mat &m) const;
    Mat clone() const { return Mat(*this); }
    template <typename U> U &at(int r, int c) { tmp_ = Mat(*this); return tmp_.at<U>(r, c); }
private:
    Mat tmp_;
    friend class Mat;
};

class _InputArray {
public:
    _InputArray() : m_(0) {}
    _InputArray(const Mat &m) : m_(const_cast<Mat *>(&m)) {}
    _InputArray(const MatExpr &e) : own_(new Mat(e)) { m_ = own_.get(); }
    template <typename T> _InputArray(const Mat_<T> &m);
    Mat getMat() const { return m_ ? *m_ : Mat(); }
    bool empty() const { return !m_ || m_->empty(); }
    int type() const { return m_ ? m_->type() : 0; }
    Size size() const { return m_ ? m_->size() : Size(); }
protected:
    Mat *m_;
    std::shared_ptr<Mat> own_;
};
class _OutputArray : public _InputArray {
public:
    _OutputArray() {}
    _OutputArray(Mat &m) : _InputArray(m) {}
    _OutputArray(const Mat &m) : _InputArray(m) {} /* OpenCV has it too: Rwc.copyTo(Twc.rowRange(0,3).colRange(0,3)), KeyFrame.cc:82 */
    template <typename T> _OutputArray(Mat_<T> &m);
    void create(int r, int c, int type) const { m_->create(r, c, type); }
    void create(Size sz, int type) const { m_->create(sz, type); }
    void release() const { m_->release(); }
    Mat &getMatRef() const { return *m_; }
    bool needed() const { return m_ != 0; }
};
static inline const _InputArray &noArray() { static _InputArray n; return n; }

inline void Mat::copyTo(OutputArray dst) const { copyTo(dst.getMatRef()); }
inline void Mat::convertTo(OutputArray dst, int rtype, double alpha, double beta) const { convertTo(dst.getMatRef(), rtype, alpha, beta); }
inline Mat::Mat(const MatExpr &e) : flags(0), dims(2), rows(0), cols(0), step(0), data(0) { e.eval(*this); }
/* Mat = expression: OpenCV evaluates INTO the destination, and Mat::create is a no-op when shape and type already agree, so
   a header that views other storage is overwritten in place -- computeDescriptors relies on it
   (ORBextractor.cc:1252: `descriptors = Mat::zeros(...)` on a rowRange view of the output matrix) */
inline Mat &Mat::operator=(const MatExpr &e)
{
    Mat m;
    e.eval(m);
    if (data && m.data != data && rows == m.rows && cols == m.cols && type() == m.type()) {
        const size_t rb = (size_t)cols * elemSize();
        for (int r = 0; r < rows; r++) memmove(ptr(r), m.ptr(r), rb);
    } else
        *this = m;
    return *this;
}

template <typename T> class Mat_ : public Mat {
public:
    Mat_() {}
    Mat_(int r, int c) : Mat(r, c, DataType<T>::type) {}
    Mat_(const Mat &m) : Mat(m) {}
    T &operator()(int r, int c) { return this->template at<T>(r, c); }
    const T &operator()(int r, int c) const { return this->template at<T>(r, c); }
};
template <typename T> _InputArray::_InputArray(const Mat_<T> &m) : m_(const_cast<Mat_<T> *>(&m)) {}
template <typename T> _OutputArray::_OutputArray(Mat_<T> &m) : _InputArray(static_cast<const Mat &>(m)) {}

template <typename T> class MatCommaInitializer_ {
public:
    MatCommaInitializer_(const Mat_<T> &m) : m_(m), i_(0) {}
    template <typename U> MatCommaInitializer_ &operator,(U v) { put((T)v); return *this; }
    void put(T v) { m_.template at<T>(i_ / m_.cols, i_ % m_.cols) = v; i_++; }
    operator Mat_<T>() const { return m_; }
    Mat_<T> m_;
    int i_;
};
template <typename T, typename U> static inline MatCommaInitializer_<T> operator<<(const Mat_<T> &m, U v)
{
    MatCommaInitializer_<T> ci(m);
    ci.put((T)v);
    return ci;
}
template <typename T> inline Mat::Mat(const MatCommaInitializer_<T> &ci) { *this = static_cast<const Mat &>(ci.m_); }

/* element-wise float add / subtract (one IEEE operation per element, like cv::add / cv::subtract on CV_32F) */
Mat shim_addsub(const Mat &a, const Mat &b, bool sub);

inline double MatExpr::dot(const Mat &m) const { return Mat(*this).dot(m); }
inline MatExpr Mat::t() const { return MatExpr::make(MatExpr::T, *this, Mat(), Mat(), 1, 0); }
inline MatExpr Mat::zeros(int r, int c, int type) { Mat m(r, c, type); m.setTo(0); return MatExpr(m); }
inline MatExpr Mat::zeros(Size sz, int type) { return zeros(sz.height, sz.width, type); }
inline MatExpr Mat::ones(int r, int c, int type) { Mat m(r, c, type); m.setTo(1); return MatExpr(m); }
inline MatExpr Mat::eye(int r, int c, int type)
{
    Mat m(r, c, type);
    m.setTo(0);
    for (int i = 0; i < std::min(r, c); i++) {
        if (m.depth() == CV_32F) m.at<float>(i, i) = 1.f; else if (m.depth() == CV_64F) m.at<double>(i, i) = 1.; else abort();
    }
    return MatExpr(m);
}

/* unary minus: -(alpha*A^T) keeps the transposed form; anything else becomes a scaled matrix */
static inline MatExpr operator-(const MatExpr &e)
{
    if (e.kind == MatExpr::T) return MatExpr::make(MatExpr::T, e.a, Mat(), Mat(), -e.alpha, 0);
    if (e.isScaled()) return MatExpr::make(MatExpr::ADDEX, e.a, Mat(), Mat(), -e.alpha, 0);
    return MatExpr::make(MatExpr::ADDEX, Mat(e), Mat(), Mat(), -1, 0);
}
static inline MatExpr operator-(const Mat &m) { return MatExpr::make(MatExpr::ADDEX, m, Mat(), Mat(), -1, 0); }

/* scale: s*A, A*s, A/s (= A * (1/s), the division happens once in double: matop.cpp operator / (Mat, double)) */
static inline MatExpr shim_scale(const MatExpr &e, double s)
{
    if (e.kind == MatExpr::T) return MatExpr::make(MatExpr::T, e.a, Mat(), Mat(), e.alpha * s, 0);
    if (e.kind == MatExpr::GEMM) return MatExpr::make(MatExpr::GEMM, e.a, e.b, e.c, e.alpha * s, e.beta * s, e.gflags);
    if (e.isScaled()) return MatExpr::make(MatExpr::ADDEX, e.a, Mat(), Mat(), e.alpha * s, 0);
    return MatExpr::make(MatExpr::ADDEX, Mat(e), Mat(), Mat(), s, 0);
}
static inline MatExpr operator*(double s, const Mat &m) { return shim_scale(MatExpr(m), s); }
static inline MatExpr operator*(const Mat &m, double s) { return shim_scale(MatExpr(m), s); }
static inline MatExpr operator*(double s, const MatExpr &e) { return shim_scale(e, s); }
static inline MatExpr operator*(const MatExpr &e, double s) { return shim_scale(e, s); }
static inline MatExpr operator/(const Mat &m, double s) { return shim_scale(MatExpr(m), 1. / s); }
static inline MatExpr operator/(const MatExpr &e, double s) { return shim_scale(e, 1. / s); }

/* product: MatOp::matmul -- a transposed or scaled left / right operand folds into the gemm's flags and alpha */
static inline MatExpr shim_matmul(const MatExpr &e1, const MatExpr &e2)
{
    int flags = 0;
    double scale = 1;
    Mat m1, m2;
    if (e1.kind == MatExpr::T) { flags |= GEMM_1_T; scale *= e1.alpha; m1 = e1.a; }
    else if (e1.isScaled()) { scale *= e1.alpha; m1 = e1.a; }
    else m1 = Mat(e1);
    if (e2.kind == MatExpr::T) { flags |= GEMM_2_T; scale *= e2.alpha; m2 = e2.a; }
    else if (e2.isScaled()) { scale *= e2.alpha; m2 = e2.a; }
    else m2 = Mat(e2);
    return MatExpr::make(MatExpr::GEMM, m1, m2, Mat(), scale, 0, flags);
}
static inline MatExpr operator*(const Mat &a, const Mat &b) { return shim_matmul(MatExpr(a), MatExpr(b)); }
static inline MatExpr operator*(const MatExpr &a, const Mat &b) { return shim_matmul(a, MatExpr(b)); }
static inline MatExpr operator*(const Mat &a, const MatExpr &b) { return shim_matmul(MatExpr(a), b); }
static inline MatExpr operator*(const MatExpr &a, const MatExpr &b) { return shim_matmul(a, b); }

/* sum / difference: a product without a C term absorbs a plain matrix (MatOp_GEMM::add / subtract); everything
   else is evaluated and combined element-wise (MatOp_AddEx with alpha = 1, beta = +-1 -> cv::add / cv::subtract) */
static inline MatExpr shim_add(const MatExpr &e1, const MatExpr &e2, bool sub)
{
    if (e1.kind == MatExpr::GEMM && e1.c.empty() && e2.kind == MatExpr::IDENT)
        return MatExpr::make(MatExpr::GEMM, e1.a, e1.b, e2.a, e1.alpha, sub ? -1 : 1, e1.gflags);
    if (!sub && e2.kind == MatExpr::GEMM && e2.c.empty() && e1.kind == MatExpr::IDENT)
        return MatExpr::make(MatExpr::GEMM, e2.a, e2.b, e1.a, e2.alpha, 1, e2.gflags);
    return MatExpr(shim_addsub(Mat(e1), Mat(e2), sub));
}
static inline MatExpr operator+(const Mat &a, const Mat &b) { return shim_add(MatExpr(a), MatExpr(b), false); }
static inline MatExpr operator+(const MatExpr &a, const Mat &b) { return shim_add(a, MatExpr(b), false); }
static inline MatExpr operator+(const Mat &a, const MatExpr &b) { return shim_add(MatExpr(a), b, false); }
static inline MatExpr operator+(const MatExpr &a, const MatExpr &b) { return shim_add(a, b, false); }
static inline MatExpr operator-(const Mat &a, const Mat &b) { return shim_add(MatExpr(a), MatExpr(b), true); }
static inline MatExpr operator-(const MatExpr &a, const Mat &b) { return shim_add(a, MatExpr(b), true); }
static inline MatExpr operator-(const Mat &a, const MatExpr &b) { return shim_add(MatExpr(a), b, true); }
static inline MatExpr operator-(const MatExpr &a, const MatExpr &b) { return shim_add(a, b, true); }

double norm(InputArray a, int normType = NORM_L2);
double norm(InputArray a, InputArray b, int normType = NORM_L2);
float fastAtan2(float y, float x);

/* yml persistence of DBoW2 (TemplatedVocabulary::save / load): declared so that the template parses; the members
   are aborting stubs in oracle/cv_shim_impl.cpp and never reached */
class FileNode {
public:
    size_t size() const;
    FileNode operator[](const char *) const;
    FileNode operator[](const std::string &) const;
    FileNode operator[](int) const;
    operator int() const;
    operator double() const;
    operator std::string() const;
};
class FileStorage {
public:
    enum { READ = 0, WRITE = 1 };
    FileStorage(const char *, int);
    FileStorage(const std::string &, int);
    bool isOpened() const;
    void release();
    FileNode operator[](const std::string &) const;
    FileNode operator[](const char *) const;
};
template <typename T> FileStorage &operator<<(FileStorage &fs, const T &)
{
    std::cerr << "cv shim: cv::FileStorage is not provided" << std::endl;
    abort();
    return fs;
}
} // namespace cv
#endif
