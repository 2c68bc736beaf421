/*
 * oracle/cv_shim_impl.cpp -- TEST INFRASTRUCTURE: the out-of-line half of oracle/shim/opencv2/ (see core/core.hpp there).
 *
 * The reference's front end delegates to OpenCV (un-vendored, absent from this image).  To compile the reference's own
 * unmodified sources into oracle/_ref/liborbslam_ref.so these entry points stand in for the library:
 *   image calls  -> the cv2-pinned primitives of oracle/orb_ref.c (resize / FAST / GaussianBlur / fastAtan2),
 *   CV_32F algebra -> OpenCV's evaluation rules restated (matmul.cpp gemm small-matrix branch vs general branch,
 *                   matop.cpp expression folding, convertTo's float scale, norm / dot in double),
 *   undistortPoints -> OpenCV's 5-iteration fixed point in double (same statement as oracle/prologue_oracle.py).
 * Each rule is pinned to cv2 4.13.0 by tests/test_ref_pin.py through the probes at the end of oracle/orbslam_ref.cpp.
 * Unsupported argument combinations abort loudly -- nothing here guesses.
 */
#include <opencv2/opencv.hpp>

extern "C" {
typedef struct { int x, y, resp; } ref_cand;
void ref_resize_linear_u8(const uint8_t *src, int sw, int sh, int sstride, uint8_t *dst, int dw, int dh, int dstride);
void ref_gaussian7_u8(const uint8_t *src, int w, int h, int sstride, uint8_t *dst, int dstride);
int ref_fast9_nms(const uint8_t *roi, int stride, int w, int h, int th, ref_cand *out, int cap);
float ref_fast_atan2(float y, float x);
}

#define SHIM_REQUIRE(cond, what)                                                         \
    do {                                                                                 \
        if (!(cond)) {                                                                   \
            std::cerr << "cv shim: unsupported call: " << what << " (" #cond ")" << std::endl; \
            abort();                                                                     \
        }                                                                                \
    } while (0)

namespace cv {

float fastAtan2(float y, float x) { return ref_fast_atan2(y, x); }

/* ---- CV_32F algebra ------------------------------------------------------------------------------------------------ */
void transpose(InputArray src_, OutputArray dst_)
{
    Mat src = src_.getMat().clone();
    SHIM_REQUIRE(src.depth() == CV_32F && src.channels() == 1, "transpose type");
    Mat dst(src.cols, src.rows, src.type());
    for (int r = 0; r < src.rows; r++)
        for (int c = 0; c < src.cols; c++) dst.at<float>(c, r) = src.at<float>(r, c);
    dst_.getMatRef() = dst;
}

/* cv::gemm for CV_32F (modules/core/src/matmul.cpp): D = alpha*op(A)*op(B) + beta*op(C) */
void gemm(InputArray a_, InputArray b_, double alpha_, InputArray c_, double beta_, OutputArray d_, int flags)
{
    /* cv::hal::gemm32f takes float alpha / beta: both are narrowed once, then used in double arithmetic */
    const double alpha = (double)(float)alpha_, beta = (double)(float)beta_;
    const Mat A = a_.getMat(), B = b_.getMat(), C = c_.getMat();
    SHIM_REQUIRE(A.type() == CV_32FC1 && B.type() == CV_32FC1 && (C.empty() || C.type() == CV_32FC1), "gemm type");
    const bool tA = (flags & GEMM_1_T) != 0, tB = (flags & GEMM_2_T) != 0, tC = (flags & GEMM_3_T) != 0;
    const int M = tA ? A.cols : A.rows, len = tA ? A.rows : A.cols;
    const int len2 = tB ? B.cols : B.rows, N = tB ? B.rows : B.cols;
    SHIM_REQUIRE(len == len2, "gemm inner size");
    if (!C.empty()) SHIM_REQUIRE((tC ? C.cols : C.rows) == M && (tC ? C.rows : C.cols) == N, "gemm C size");
    Mat D(M, N, CV_32FC1);
    /* small-matrix branch: "flags == 0 && 2 <= len && len <= 4 && (len == d_size.width || len == d_size.height)":
       the products of one output element accumulate left to right in FLOAT, the scaling by alpha and the C term
       are applied in double and narrowed once */
    const bool small = flags == 0 && 2 <= len && len <= 4 && (len == N || len == M);
    for (int i = 0; i < M; i++)
        for (int j = 0; j < N; j++) {
            const double cv = C.empty() ? 0.0 : (double)(tC ? C.at<float>(j, i) : C.at<float>(i, j));
            if (small) {
                float t = A.at<float>(i, 0) * B.at<float>(0, j);
                for (int k = 1; k < len; k++) t = t + A.at<float>(i, k) * B.at<float>(k, j);
                D.at<float>(i, j) = (float)(t * alpha + cv * beta);
            } else {
                /* GEMMSingleMul<float, double>: double accumulation in k order */
                double s = 0;
                for (int k = 0; k < len; k++)
                    s += (double)(tA ? A.at<float>(k, i) : A.at<float>(i, k)) * (double)(tB ? B.at<float>(j, k) : B.at<float>(k, j));
                D.at<float>(i, j) = C.empty() ? (float)(s * alpha) : (float)(s * alpha + cv * beta);
            }
        }
    d_.getMatRef() = D;
}

Mat shim_addsub(const Mat &a, const Mat &b, bool sub)
{
    SHIM_REQUIRE(a.type() == b.type() && a.rows == b.rows && a.cols == b.cols && a.depth() == CV_32F, "add/subtract operands");
    Mat d(a.rows, a.cols, a.type());
    const int n = a.cols * a.channels();
    for (int r = 0; r < a.rows; r++) {
        const float *pa = a.ptr<float>(r), *pb = b.ptr<float>(r);
        float *pd = d.ptr<float>(r);
        for (int i = 0; i < n; i++) pd[i] = sub ? pa[i] - pb[i] : pa[i] + pb[i];
    }
    return d;
}

void MatExpr::eval(Mat &m) const
{
    switch (kind) {
    case IDENT: m = a; break;
    case T: { /* MatOp_T::assign: transpose, then convertTo when alpha != 1 */
        Mat t;
        transpose(a, t);
        if (alpha != 1) t.convertTo(t, -1, alpha);
        m = t;
        break;
    }
    case ADDEX: { /* only alpha*a reaches evaluation (sums are resolved when they are formed) */
        SHIM_REQUIRE(b.empty() && beta == 0, "MatExpr ADDEX form");
        Mat t;
        a.convertTo(t, -1, alpha);
        if (t.data == a.data) t = t.clone();
        m = t;
        break;
    }
    case GEMM: {
        Mat t;
        gemm(a, b, alpha, c, beta, t, gflags);
        m = t;
        break;
    }
    }
}

MatExpr Mat::inv() const
{
    std::cerr << "cv shim: Mat::inv is not on the front-end path" << std::endl;
    abort();
}

/* Mat::dot on CV_32F: dotProd_32f -> below one SIMD width the tail loop accumulates (double)a*b in order */
double Mat::dot(InputArray m_) const
{
    const Mat m = m_.getMat();
    SHIM_REQUIRE(type() == CV_32FC1 && m.type() == CV_32FC1 && total() == m.total(), "dot operands");
    SHIM_REQUIRE(total() < 4, "dot longer than the scalar tail (SIMD block order not restated)");
    double r = 0;
    const int n = (int)total();
    for (int i = 0; i < n; i++) r += (double)at<float>(i) * (double)m.at<float>(i);
    return r;
}

/* cv::norm L2 / L1 on CV_32F: squares (absolute values) accumulate in double */
double norm(InputArray a_, int normType)
{
    const Mat a = a_.getMat();
    SHIM_REQUIRE(a.depth() == CV_32F && a.channels() == 1, "norm type");
    double s = 0;
    for (int r = 0; r < a.rows; r++)
        for (int c = 0; c < a.cols; c++) {
            const float v = a.at<float>(r, c);
            if (normType == NORM_L2) s += (double)v * v;
            else if (normType == NORM_L1) s += std::fabs((double)v);
            else SHIM_REQUIRE(false, "norm kind");
        }
    return normType == NORM_L2 ? std::sqrt(s) : s;
}
double norm(InputArray a_, InputArray b_, int normType)
{
    const Mat a = a_.getMat(), b = b_.getMat();
    SHIM_REQUIRE(a.type() == CV_32FC1 && b.type() == CV_32FC1 && a.rows == b.rows && a.cols == b.cols, "norm(a, b) operands");
    double s = 0;
    for (int r = 0; r < a.rows; r++)
        for (int c = 0; c < a.cols; c++) {
            const float d = a.at<float>(r, c) - b.at<float>(r, c);
            if (normType == NORM_L1) s += (double)std::fabs(d);
            else if (normType == NORM_L2) s += (double)d * d;
            else SHIM_REQUIRE(false, "norm kind");
        }
    return normType == NORM_L2 ? std::sqrt(s) : s;
}

/* ---- image calls of ORBextractor.cc ---------------------------------------------------------------------------------- */
static inline int reflect101(int p, int n)
{
    if (n == 1) return 0;
    while (p < 0 || p >= n) p = p < 0 ? -p : 2 * (n - 1) - p;
    return p;
}

void resize(InputArray src_, OutputArray dst_, Size dsize, double fx, double fy, int interpolation)
{
    const Mat src = src_.getMat();
    SHIM_REQUIRE(src.type() == CV_8UC1 && interpolation == INTER_LINEAR && fx == 0 && fy == 0 && dsize.width > 0 && dsize.height > 0,
                 "resize: only 8UC1 INTER_LINEAR with an explicit size");
    dst_.create(dsize, src.type()); /* a ROI of the right shape is written in place */
    Mat dst = dst_.getMat();
    ref_resize_linear_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, dst.cols, dst.rows, (int)dst.step);
}

void copyMakeBorder(InputArray src_, OutputArray dst_, int top, int bottom, int left, int right, int borderType, const Scalar &)
{
    const Mat src = src_.getMat();
    SHIM_REQUIRE((borderType & ~BORDER_ISOLATED) == BORDER_REFLECT_101 && src.type() == CV_8UC1, "copyMakeBorder: only 8UC1 REFLECT_101");
    dst_.create(src.rows + top + bottom, src.cols + left + right, src.type());
    Mat dst = dst_.getMat();
    /* in place (src is the interior of dst, ORBextractor.cc:1370): the interior copy is the identity and the border
       only reads interior pixels, so any order is safe */
    for (int r = 0; r < dst.rows; r++) {
        const uchar *s = src.ptr(reflect101(r - top, src.rows));
        uchar *d = dst.ptr(r);
        if (d + left != s) memmove(d + left, s, (size_t)src.cols);
    }
    for (int r = 0; r < dst.rows; r++) {
        uchar *d = dst.ptr(r);
        for (int c = 0; c < left; c++) d[c] = d[left + reflect101(c - left, src.cols)];
        for (int c = 0; c < right; c++) d[left + src.cols + c] = d[left + reflect101(src.cols + c, src.cols)];
    }
}

void GaussianBlur(InputArray src_, OutputArray dst_, Size ksize, double sigmaX, double sigmaY, int borderType)
{
    const Mat src = src_.getMat();
    SHIM_REQUIRE(src.type() == CV_8UC1 && ksize.width == 7 && ksize.height == 7 && sigmaX == 2 && sigmaY == 2 && borderType == BORDER_REFLECT_101,
                 "GaussianBlur: only 8UC1 7x7 sigma 2 REFLECT_101");
    dst_.create(src.size(), src.type());
    Mat dst = dst_.getMat();
    ref_gaussian7_u8(src.data, src.cols, src.rows, (int)src.step, dst.data, (int)dst.step); /* in-place safe: row pass buffered */
}

void FAST(InputArray image_, std::vector<KeyPoint> &keypoints, int threshold, bool nonmaxSuppression)
{
    const Mat img = image_.getMat();
    SHIM_REQUIRE(img.type() == CV_8UC1 && nonmaxSuppression, "FAST: only 8UC1 with non-max suppression");
    keypoints.clear();
    if (img.empty()) return;
    std::vector<ref_cand> out((size_t)img.rows * img.cols);
    const int n = ref_fast9_nms(img.data, (int)img.step, img.cols, img.rows, threshold, out.data(), (int)out.size());
    keypoints.reserve(n);
    for (int i = 0; i < n; i++) keypoints.push_back(KeyPoint((float)out[i].x, (float)out[i].y, 7.f, -1, (float)out[i].resp));
}

/* only reachable from the dead ComputeKeyPointsOld; keeps the npoints strongest (ties with the weakest kept are kept) */
void KeyPointsFilter::retainBest(std::vector<KeyPoint> &kp, int npoints)
{
    if (npoints < 0 || (int)kp.size() <= npoints) return;
    if (npoints == 0) { kp.clear(); return; }
    std::vector<KeyPoint> s(kp);
    std::nth_element(s.begin(), s.begin() + npoints - 1, s.end(), [](const KeyPoint &a, const KeyPoint &b) { return a.response > b.response; });
    const float cut = s[npoints - 1].response;
    std::vector<KeyPoint> keep;
    for (size_t i = 0; i < kp.size(); i++)
        if (kp[i].response >= cut) keep.push_back(kp[i]);
    kp.swap(keep);
}

/* cv::undistortPoints(src, dst, K, distCoeffs, Mat(), P) for CV_32FC2 points (Frame.cc:501, :553): double arithmetic,
   five fixed-point iterations, the icdist < 0 fallback, re-projection by P with an identity R */
void undistortPoints(InputArray src_, OutputArray dst_, InputArray K_, InputArray dist_, InputArray R_, InputArray P_)
{
    const Mat src = src_.getMat(), K = K_.getMat(), dist = dist_.getMat(), P = P_.getMat();
    SHIM_REQUIRE(src.type() == CV_32FC2 && (src.cols == 1 || src.rows == 1) && src.isContinuous(), "undistortPoints: CV_32FC2 vector");
    SHIM_REQUIRE(K.type() == CV_32FC1 && K.rows == 3 && K.cols == 3 && R_.empty(), "undistortPoints: 3x3 CV_32F camera matrix, no R");
    SHIM_REQUIRE(dist.type() == CV_32FC1 && (dist.total() == 4 || dist.total() == 5 || dist.total() == 8), "undistortPoints: distortion vector");
    SHIM_REQUIRE(P.empty() || (P.type() == CV_32FC1 && P.rows == 3 && P.cols >= 3), "undistortPoints: P");
    const int n = (int)src.total();
    Mat out(src.rows, src.cols, CV_32FC2);
    double k[12] = {0};
    for (int i = 0; i < (int)dist.total(); i++) k[i] = (double)dist.at<float>(i);
    const double fx = K.at<float>(0, 0), fy = K.at<float>(1, 1), cx = K.at<float>(0, 2), cy = K.at<float>(1, 2);
    const double ifx = 1. / fx, ify = 1. / fy;
    double PP[3][3] = {{1, 0, 0}, {0, 1, 0}, {0, 0, 1}};
    if (!P.empty())
        for (int r = 0; r < 3; r++)
            for (int c = 0; c < 3; c++) PP[r][c] = (double)P.at<float>(r, c);
    const float *s = src.ptr<float>();
    float *d = out.ptr<float>();
    for (int i = 0; i < n; i++) {
        double x = ((double)s[2 * i] - cx) * ifx, y = ((double)s[2 * i + 1] - cy) * ify;
        const double x0 = x, y0 = y;
        for (int j = 0; j < 5; j++) {
            const double r2 = x * x + y * y;
            const double icdist = (1 + ((k[7] * r2 + k[6]) * r2 + k[5]) * r2) / (1 + ((k[4] * r2 + k[1]) * r2 + k[0]) * r2);
            if (icdist < 0) { x = x0; y = y0; break; }
            const double dx = 2 * k[2] * x * y + k[3] * (r2 + 2 * x * x) + k[8] * r2 + k[9] * r2 * r2;
            const double dy = k[2] * (r2 + 2 * y * y) + 2 * k[3] * x * y + k[10] * r2 + k[11] * r2 * r2;
            x = (x0 - dx) * icdist;
            y = (y0 - dy) * icdist;
        }
        const double xx = PP[0][0] * x + PP[0][1] * y + PP[0][2];
        const double yy = PP[1][0] * x + PP[1][1] * y + PP[1][2];
        const double ww = 1. / (PP[2][0] * x + PP[2][1] * y + PP[2][2]);
        d[2 * i] = (float)(xx * ww);
        d[2 * i + 1] = (float)(yy * ww);
    }
    dst_.getMatRef() = out;
}

/* yml persistence (DBoW2 TemplatedVocabulary::save / load): never on the path */
#define SHIM_NOYML { std::cerr << "cv shim: cv::FileStorage is not provided" << std::endl; abort(); }
size_t FileNode::size() const SHIM_NOYML
FileNode FileNode::operator[](const char *) const SHIM_NOYML
FileNode FileNode::operator[](const std::string &) const SHIM_NOYML
FileNode FileNode::operator[](int) const SHIM_NOYML
FileNode::operator int() const SHIM_NOYML
FileNode::operator double() const SHIM_NOYML
FileNode::operator std::string() const SHIM_NOYML
FileStorage::FileStorage(const char *, int) SHIM_NOYML
FileStorage::FileStorage(const std::string &, int) SHIM_NOYML
bool FileStorage::isOpened() const SHIM_NOYML
void FileStorage::release() SHIM_NOYML
FileNode FileStorage::operator[](const std::string &) const SHIM_NOYML
FileNode FileStorage::operator[](const char *) const SHIM_NOYML
} // namespace cv
