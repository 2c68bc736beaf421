/* Stand-in for <opencv2/imgproc/imgproc.hpp> -- TEST INFRASTRUCTURE (see opencv2/core/core.hpp in this tree).
 * The image-library calls of the reference's front end, implemented in oracle/cv_shim_impl.cpp over the cv2-pinned
 * primitives of oracle/orb_ref.c: cv::resize INTER_LINEAR 8UC1 (ORBextractor.cc:1366), cv::copyMakeBorder
 * REFLECT_101 (:1370, :1382), cv::GaussianBlur 7x7 sigma 2 REFLECT_101 8UC1 (:1314), cv::undistortPoints
 * (Frame.cc:501, :553).  Any other argument combination aborts instead of guessing. */
#ifndef ORBB200_ORACLE_OPENCV_IMGPROC_STANDIN
#define ORBB200_ORACLE_OPENCV_IMGPROC_STANDIN
#include "../core/core.hpp"
namespace cv {
enum { INTER_NEAREST = 0, INTER_LINEAR = 1, INTER_CUBIC = 2, INTER_AREA = 3 };
void resize(InputArray src, OutputArray dst, Size dsize, double fx = 0, double fy = 0, int interpolation = INTER_LINEAR);
void copyMakeBorder(InputArray src, OutputArray dst, int top, int bottom, int left, int right, int borderType,
                    const Scalar &value = Scalar());
void GaussianBlur(InputArray src, OutputArray dst, Size ksize, double sigmaX, double sigmaY = 0, int borderType = BORDER_DEFAULT);
void undistortPoints(InputArray src, OutputArray dst, InputArray cameraMatrix, InputArray distCoeffs,
                     InputArray R = noArray(), InputArray P = noArray());
} // namespace cv
#endif
