/*
 * Stand-in for the reference's include/Converter.h -- TEST INFRASTRUCTURE (oracle/Makefile puts this directory first
 * on the include path).  The reference header drags in <Eigen/Dense> and g2o's Eigen-based types
 * (include/Converter.h:26-28); Eigen is absent from this image.  src/Frame.cc:463 and src/KeyFrame.cc:63 use ONE
 * function of it, Converter::toDescriptorVector (src/Converter.cc:27-36: the rows of the descriptor matrix as a
 * vector of 1x32 headers), which oracle/orbslam_ref.cpp defines with the same meaning.
 */
#ifndef ORBB200_ORACLE_CONVERTER_STANDIN
#define ORBB200_ORACLE_CONVERTER_STANDIN
#include <opencv2/core/core.hpp>
#include <vector>
namespace ORB_SLAM2 {
class Converter {
public:
    static std::vector<cv::Mat> toDescriptorVector(const cv::Mat &Descriptors);
};
} // namespace ORB_SLAM2
#endif
