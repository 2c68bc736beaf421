/*
 * Rectify.h -- the stereo rectification of Examples/Stereo/stereo_euroc.cc (:60-97 the maps, :131-132 the per-frame cv::remap) on
 * the device.  Header-only over the C-ABI (orbb200_remap_*, orbb200_init_undistort_rectify_map).
 *
 *   orbb200::StereoRectifier rect(K_l, D_l, R_l, P_l, K_r, D_r, R_r, P_r, cols, rows);   // cv::Mat arguments as the example reads them
 *   rect(imLeft, imRight, imLeftRect, imRightRect);                                       // == the two cv::remap calls
 *
 * Results are bit-exact with OpenCV (tests/test_gpu_remap.py against cv2 4.13).  Matrices may be CV_64F or CV_32F; P may be 3x4
 * (its left 3x3 block is used, like P.rowRange(0,3).colRange(0,3)).
 */
#ifndef ORBB200_RECTIFY_H
#define ORBB200_RECTIFY_H

#include <iostream>
#include <vector>

#include "../orbb200.h"
#ifdef ORBB200_USE_OPENCV
#include <opencv2/core/core.hpp>
#else
#include "cv_shim.h"
#endif

namespace orbb200 {

class StereoRectifier {
public:
    StereoRectifier(const cv::Mat &K_l, const cv::Mat &D_l, const cv::Mat &R_l, const cv::Mat &P_l, const cv::Mat &K_r, const cv::Mat &D_r,
                    const cv::Mat &R_r, const cv::Mat &P_r, int cols, int rows, int device = -1)
        : left_(nullptr), right_(nullptr), cols_(cols), rows_(rows)
    {
        std::vector<float> mx((size_t)cols * rows), my((size_t)cols * rows);
        if (!maps(K_l, D_l, R_l, P_l, mx, my) || orbb200_remap_create(&left_, mx.data(), my.data(), cols, rows, device) != ORBB200_OK) fail();
        if (!maps(K_r, D_r, R_r, P_r, mx, my) || orbb200_remap_create(&right_, mx.data(), my.data(), cols, rows, device) != ORBB200_OK) fail();
    }
    ~StereoRectifier()
    {
        orbb200_remap_destroy(left_);
        orbb200_remap_destroy(right_);
    }
    StereoRectifier(const StereoRectifier &) = delete;
    StereoRectifier &operator=(const StereoRectifier &) = delete;

    bool ok() const { return left_ && right_; }
    orbb200_remap *Left() const { return left_; }
    orbb200_remap *Right() const { return right_; }

    /* cv::remap(imLeft, imLeftRect, M1l, M2l, INTER_LINEAR); cv::remap(imRight, imRightRect, M1r, M2r, INTER_LINEAR) */
    bool operator()(const cv::Mat &imLeft, const cv::Mat &imRight, cv::Mat &imLeftRect, cv::Mat &imRightRect) const
    {
        return one(left_, imLeft, imLeftRect) && one(right_, imRight, imRightRect);
    }

private:
    static double at(const cv::Mat &m, int i)
    {
        const int r = i / m.cols, c = i % m.cols;
        return m.type() == CV_64F ? m.at<double>(r, c) : (double)m.at<float>(r, c);
    }
    bool maps(const cv::Mat &K, const cv::Mat &D, const cv::Mat &R, const cv::Mat &P, std::vector<float> &mx, std::vector<float> &my) const
    {
        double k[9], r[9], p[9], d[5] = {0, 0, 0, 0, 0};
        if (K.rows != 3 || K.cols != 3 || R.rows != 3 || R.cols != 3 || P.rows < 3 || P.cols < 3) return false;
        for (int i = 0; i < 9; i++) { k[i] = at(K, i); r[i] = at(R, i); }
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) p[3 * i + j] = at(P, i * P.cols + j);
        const int nd = D.rows * D.cols < 5 ? D.rows * D.cols : 5;
        for (int i = 0; i < nd; i++) d[i] = at(D, i);
        return orbb200_init_undistort_rectify_map(k, d, nd, r, p, cols_, rows_, mx.data(), my.data()) == ORBB200_OK;
    }
    bool one(orbb200_remap *h, const cv::Mat &src, cv::Mat &dst) const
    {
        if (!h || src.empty()) return false;
        dst.create(rows_, cols_, CV_8UC1);
        const uint8_t *in[1] = {src.ptr<uint8_t>()};
        uint8_t *out[1] = {dst.ptr<uint8_t>()};
        if (orbb200_remap_batch(h, in, 1, src.cols, src.rows, (int)src.step, out, (int)dst.step) != ORBB200_OK) { fail(); return false; }
        return true;
    }
    void fail() const { std::cerr << "StereoRectifier (orbb200): " << orbb200_last_error() << std::endl; }

    orbb200_remap *left_, *right_;
    int cols_, rows_;
};

} // namespace orbb200

#endif /* ORBB200_RECTIFY_H */
