/*
 * ORBextractor.h -- source-compatible replacement of ORB_SLAM2::ORBextractor (reference include/ORBextractor.h:53-135)
 * over the C-ABI of liborbb200.so (include/orbb200.h).  Header-only host C++; link with -lorbb200.
 *
 * Same constructor (nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST), same operator()(image, mask, keypoints,
 * descriptors), same accessors, same public mvImagePyramid -- so Frame::ExtractORB (src/Frame.cc:285-292),
 * Frame::Frame (:73-79) and Tracking::Tracking (src/Tracking.cc:118-130) compile unchanged.
 *
 * Behaviour kept from the reference:
 *   - empty image: return with the outputs untouched (src/ORBextractor.cc:1264);
 *   - the mask is ignored;  image must be CV_8UC1 (:1269, an assert there; an error line + 0 keypoints here);
 *   - keypoints are cleared then filled, descriptors is created N x 32 CV_8U or released when N == 0 (:1288-1298);
 *   - no exceptions, no return code: a CUDA failure prints one line to std::cerr and yields 0 keypoints, so that
 *     Frame::Frame takes its `if(mvKeys.empty()) return;` exit (src/Frame.cc:89).
 * There is no CPU fallback.
 *
 * Extensions (do not exist in the reference; defaults keep its behaviour):
 *   SetDevice(int)               choose the GPU before the first call (default: current device)
 *   SetPyramidMode(mode)         PYRAMID_EAGER (default): mvImagePyramid is downloaded after every call, because
 *                                Frame::ComputeStereoMatches reads it (src/Frame.cc:584,674,686,691);
 *                                PYRAMID_NONE: leave it empty (use orbb200::ComputeStereoMatches on the device instead)
 *   Handle()                     the underlying orbb200_extractor*
 */
#ifndef ORBB200_ORBEXTRACTOR_H
#define ORBB200_ORBEXTRACTOR_H

#include <iostream>
#include <list>
#include <vector>

#ifdef ORBB200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif
#include "../orbb200.h"

/* The classes live in ORB_SLAM2 so that the reference's callers compile unchanged.  A test that links the reference's own
   classes of the same name beside these (tests/cpp/ref_match.cpp) renames the namespace with -DORBB200_NAMESPACE=... */
#ifndef ORBB200_NAMESPACE
#define ORBB200_NAMESPACE ORB_SLAM2
#endif
namespace ORBB200_NAMESPACE {

class ORBextractor;

/* mvImagePyramid, lazily: the levels stay on the device until somebody reads them (Frame::ComputeStereoMatches does,
 * src/Frame.cc:584,674,691; orbb200::ComputeStereoMatches does not).  The first access after an extraction fetches the image's
 * whole pyramid slab with ONE device-to-host copy; the cv::Mat of level l is a header over that host slab (rows `pitch` bytes
 * apart, like the reference's ROI-of-a-bordered-image levels, src/ORBextractor.cc:1358-1360).  Indexing, size(), begin()/end()
 * and conversion to const std::vector<cv::Mat>& cover every use the reference makes of the member. */
class LazyPyramid {
public:
    LazyPyramid() : owner_(nullptr), stale_(false) {}
    size_t size() const { return levels_.size(); }
    bool empty() const { return levels_.empty(); }
    void resize(size_t n) { levels_.resize(n); }
    cv::Mat &operator[](size_t l) { sync(); return levels_[l]; }
    const cv::Mat &operator[](size_t l) const { sync(); return levels_[l]; }
    cv::Mat &at(size_t l) { sync(); return levels_.at(l); }
    const cv::Mat &at(size_t l) const { sync(); return levels_.at(l); }
    std::vector<cv::Mat>::const_iterator begin() const { sync(); return levels_.begin(); }
    std::vector<cv::Mat>::const_iterator end() const { sync(); return levels_.end(); }
    operator const std::vector<cv::Mat> &() const { sync(); return levels_; }
    bool resident_on_host() const { return !stale_; } /* false: the levels of the last extraction have not been downloaded */

private:
    friend class ORBextractor;
    inline void sync() const;
    ORBextractor *owner_;
    mutable bool stale_;
    mutable std::vector<cv::Mat> levels_;
};

class ORBextractor {
public:
    enum { HARRIS_SCORE = 0, FAST_SCORE = 1 };
    enum PyramidMode { PYRAMID_LAZY = 0, PYRAMID_EAGER = 1, PYRAMID_NONE = 2 }; /* LAZY: one D2H on first read; EAGER: at the end of operator() */

    ORBextractor(int nfeatures_, float scaleFactor_, int nlevels_, int iniThFAST_, int minThFAST_)
        : nfeatures(nfeatures_), scaleFactor(scaleFactor_), nlevels(nlevels_), iniThFAST(iniThFAST_), minThFAST(minThFAST_),
          handle_(nullptr), device_(-1), pyramidMode_(PYRAMID_LAZY)
    {
        static_assert(sizeof(cv::KeyPoint) == sizeof(orbb200_keypoint), "cv::KeyPoint must be the 28-byte POD");
        mvScaleFactor.resize(nlevels);
        mvInvScaleFactor.resize(nlevels);
        mvLevelSigma2.resize(nlevels);
        mvInvLevelSigma2.resize(nlevels);
        mnFeaturesPerLevel.resize(nlevels);
        umax.resize(16);
        if (orbb200_compute_level_tables(nfeatures, scaleFactor_, nlevels, mvScaleFactor.data(), mvInvScaleFactor.data(),
                                         mvLevelSigma2.data(), mvInvLevelSigma2.data(), mnFeaturesPerLevel.data(), umax.data()))
            std::cerr << "ORBextractor: " << orbb200_last_error() << std::endl;
        mvImagePyramid.resize(nlevels);
        mvImagePyramid.owner_ = this;
    }

    ~ORBextractor() { orbb200_extractor_destroy(handle_); } /* the reference never deletes its extractors; harmless */

    ORBextractor(const ORBextractor &) = delete;
    ORBextractor &operator=(const ORBextractor &) = delete;

    // Compute the ORB features and descriptors on an image.  Mask is ignored (as in the reference).
    void operator()(cv::InputArray _image, cv::InputArray /*_mask*/, std::vector<cv::KeyPoint> &_keypoints,
                    cv::OutputArray _descriptors)
    {
        const cv::Mat image = _image.getMat();
        if (image.empty()) return; // :1264
        _keypoints.clear();
        if (image.type() != CV_8UC1) { fail("image must be CV_8UC1", _descriptors); return; }
        if (!ensure_handle()) { fail(orbb200_last_error(), _descriptors); return; }
        const int cap = orbb200_max_keypoints_for_shape(handle_, image.cols, image.rows);
        _keypoints.resize(cap);
        descBuf_.resize((size_t)cap * 32);
        int n = 0;
        const int rc = orbb200_extract(handle_, image.data, image.cols, image.rows, (int)image.step,
                                       reinterpret_cast<orbb200_keypoint *>(_keypoints.data()), descBuf_.data(), cap, &n);
        if (rc != ORBB200_OK) { _keypoints.clear(); fail(orbb200_last_error(), _descriptors); return; }
        _keypoints.resize(n);
        if (n == 0) {
            _descriptors.release();
        } else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), descBuf_.data() + (size_t)i * 32, 32);
        }
        pyramid_ready();
    }

    /* Extension: the cv::cvtColor prologue of Tracking::GrabImage* (src/Tracking.cc:177-202, 218-231, 255-266)
     * fused into the extraction.  imColor is CV_8UC3 / CV_8UC4, bRGB = Tracking::mbRGB (channel 0 is red).  imGray receives
     * what cvtColor would have produced (Tracking keeps it as mImGray) -- it is level 0 of the device pyramid. */
    void ExtractColor(const cv::Mat &imColor, bool bRGB, std::vector<cv::KeyPoint> &_keypoints, cv::OutputArray _descriptors,
                      cv::Mat *imGray = nullptr)
    {
        if (imColor.empty()) return;
        _keypoints.clear();
        const int ch = imColor.channels();
        if (imColor.depth() != CV_8U || (ch != 3 && ch != 4)) { fail("colour image must be CV_8UC3 or CV_8UC4", _descriptors); return; }
        if (!ensure_handle()) { fail(orbb200_last_error(), _descriptors); return; }
        const int cap = orbb200_max_keypoints_for_shape(handle_, imColor.cols, imColor.rows);
        _keypoints.resize(cap);
        descBuf_.resize((size_t)cap * 32);
        int n = 0;
        const int rc = orbb200_extract_color(handle_, imColor.data, imColor.cols, imColor.rows, (int)imColor.step, ch, bRGB ? 1 : 0,
                                             reinterpret_cast<orbb200_keypoint *>(_keypoints.data()), descBuf_.data(), cap, &n);
        if (rc != ORBB200_OK) { _keypoints.clear(); fail(orbb200_last_error(), _descriptors); return; }
        _keypoints.resize(n);
        if (n == 0) {
            _descriptors.release();
        } else {
            _descriptors.create(n, 32, CV_8U);
            cv::Mat d = _descriptors.getMat();
            for (int i = 0; i < n; i++) std::memcpy(d.ptr(i), descBuf_.data() + (size_t)i * 32, 32);
        }
        pyramid_ready();
        if (imGray) {
            int w = 0, h = 0;
            imGray->create(imColor.rows, imColor.cols, CV_8UC1);
            orbb200_pyramid_level(handle_, 0, 0, imGray->data, (int)imGray->step, &w, &h);
        }
    }

    int inline GetLevels() { return nlevels; }
    float inline GetScaleFactor() { return (float)scaleFactor; }
    std::vector<float> inline GetScaleFactors() { return mvScaleFactor; }
    std::vector<float> inline GetInverseScaleFactors() { return mvInvScaleFactor; }
    std::vector<float> inline GetScaleSigmaSquares() { return mvLevelSigma2; }
    std::vector<float> inline GetInverseScaleSigmaSquares() { return mvInvLevelSigma2; }

    LazyPyramid mvImagePyramid;

    /* ---- extensions -------------------------------------------------------------------------------- */
    void SetDevice(int device) { device_ = device; }
    void SetPyramidMode(PyramidMode m) { pyramidMode_ = m; }
    orbb200_extractor *Handle() { ensure_handle(); return handle_; }
    const std::vector<int> &GetFeaturesPerLevel() const { return mnFeaturesPerLevel; }

protected:
    bool ensure_handle()
    {
        if (handle_) return true;
        return orbb200_extractor_create(&handle_, nfeatures, (float)scaleFactor, nlevels, iniThFAST, minThFAST, device_) == ORBB200_OK;
    }
    void fail(const char *msg, cv::OutputArray d)
    {
        std::cerr << "ORBextractor (orbb200): " << msg << std::endl;
        d.release();
    }
    friend class LazyPyramid;
    /* after an extraction: mark the host copy stale (LAZY), refresh it now (EAGER), or leave it alone (NONE) */
    void pyramid_ready()
    {
        mvImagePyramid.stale_ = pyramidMode_ != PYRAMID_NONE;
        if (pyramidMode_ == PYRAMID_EAGER) mvImagePyramid.sync();
    }
    /* one device-to-host copy of the last image's pyramid slab; level headers over it */
    void download_pyramid()
    {
        if (!handle_) return;
        int offs[ORBB200_MAX_LEVELS], pit[ORBB200_MAX_LEVELS], ws[ORBB200_MAX_LEVELS], hs[ORBB200_MAX_LEVELS];
        size_t bytes = 0;
        if (orbb200_pyramid_layout(handle_, offs, pit, ws, hs, &bytes) != ORBB200_OK) return;
        const uint8_t *slab = nullptr; /* pinned, owned by the handle: valid until the next download */
        if (orbb200_pyramid_map(handle_, 0, &slab) != ORBB200_OK) {
            std::cerr << "ORBextractor (orbb200): " << orbb200_last_error() << std::endl;
            return;
        }
        for (int l = 0; l < nlevels; l++)
            mvImagePyramid.levels_[l] = cv::Mat(hs[l], ws[l], CV_8UC1, const_cast<uint8_t *>(slab) + offs[l], (size_t)pit[l]);
    }

    int nfeatures;
    double scaleFactor;
    int nlevels;
    int iniThFAST;
    int minThFAST;
    std::vector<int> mnFeaturesPerLevel;
    std::vector<int> umax;
    std::vector<float> mvScaleFactor;
    std::vector<float> mvInvScaleFactor;
    std::vector<float> mvLevelSigma2;
    std::vector<float> mvInvLevelSigma2;

    orbb200_extractor *handle_;
    int device_;
    PyramidMode pyramidMode_;
    std::vector<unsigned char> descBuf_;
};

inline void LazyPyramid::sync() const
{
    if (!stale_ || !owner_) return;
    stale_ = false;
    owner_->download_pyramid();
}

} // namespace ORBB200_NAMESPACE (ORB_SLAM2 by default)

namespace orbb200 {

/* Device replacement of Frame::ComputeStereoMatches (src/Frame.cc:577-751): uses the device-resident pyramids and
 * outputs of the last operator() call of both extractors.  mvuRight / mvDepth get N = #left keypoints entries
 * (-1 where unmatched), exactly what the reference leaves in Frame::mvuRight / mvDepth.  Returns the match count
 * (or -1 on error, after printing it). */
inline int ComputeStereoMatches(ORBB200_NAMESPACE::ORBextractor &left, ORBB200_NAMESPACE::ORBextractor &right, int nLeft, float mbf, float mb,
                                std::vector<float> &mvuRight, std::vector<float> &mvDepth)
{
    const int cap = orbb200_last_capacity(left.Handle());
    std::vector<float> u(cap > 0 ? cap : 1, -1.f), d(cap > 0 ? cap : 1, -1.f);
    int nvalid = 0;
    mvuRight.assign(nLeft, -1.0f);
    mvDepth.assign(nLeft, -1.0f);
    if (nLeft == 0) return 0;
    if (orbb200_stereo_match(left.Handle(), right.Handle(), mbf, mb, u.data(), d.data(), &nvalid) != ORBB200_OK) {
        std::cerr << "ComputeStereoMatches (orbb200): " << orbb200_last_error() << std::endl;
        return -1;
    }
    for (int i = 0; i < nLeft && i < cap; i++) { mvuRight[i] = u[i]; mvDepth[i] = d[i]; }
    return nvalid;
}

/* Device replacement of Frame::UndistortKeyPoints (src/Frame.cc:470-515).  mK is the 3x3 CV_32F camera matrix, mDistCoef the
 * 4x1 / 5x1 CV_32F coefficients (k1, k2, p1, p2[, k3]) exactly as Frame holds them; k1 == 0 copies (:473-477). */
inline void UndistortKeyPoints(const std::vector<cv::KeyPoint> &mvKeys, const cv::Mat &mK, const cv::Mat &mDistCoef,
                               std::vector<cv::KeyPoint> &mvKeysUn, int device = -1)
{
    const int nCoef = mDistCoef.rows * mDistCoef.cols;
    float dist[5] = {0, 0, 0, 0, 0};
    for (int i = 0; i < nCoef && i < 5; i++) dist[i] = mDistCoef.ptr<float>()[i];
    if (dist[0] == 0.0f || mvKeys.empty()) { mvKeysUn = mvKeys; return; }
    const float cam[4] = {mK.at<float>(0, 0), mK.at<float>(1, 1), mK.at<float>(0, 2), mK.at<float>(1, 2)};
    mvKeysUn.resize(mvKeys.size());
    if (orbb200_undistort_keypoints(reinterpret_cast<const orbb200_keypoint *>(mvKeys.data()), (int)mvKeys.size(), cam, dist,
                                    reinterpret_cast<orbb200_keypoint *>(mvKeysUn.data()), device) != ORBB200_OK) {
        std::cerr << "UndistortKeyPoints (orbb200): " << orbb200_last_error() << std::endl;
        mvKeysUn = mvKeys;
    }
}

} // namespace orbb200

#endif /* ORBB200_ORBEXTRACTOR_H */
