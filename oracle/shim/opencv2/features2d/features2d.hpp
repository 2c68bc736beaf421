/* Stand-in for <opencv2/features2d/features2d.hpp> -- TEST INFRASTRUCTURE (see opencv2/core/core.hpp in this tree).
 * cv::FAST (TYPE_9_16, ORBextractor.cc:988, :993) over oracle/orb_ref.c's cv2-pinned ref_fast9_nms;
 * KeyPointsFilter::retainBest is only called from the dead ComputeKeyPointsOld (:1205, :1223) and merely has to link. */
#ifndef ORBB200_ORACLE_OPENCV_FEATURES2D_STANDIN
#define ORBB200_ORACLE_OPENCV_FEATURES2D_STANDIN
#include "../core/core.hpp"
namespace cv {
void FAST(InputArray image, std::vector<KeyPoint> &keypoints, int threshold, bool nonmaxSuppression = true);
class KeyPointsFilter {
public:
    static void retainBest(std::vector<KeyPoint> &keypoints, int npoints);
};
} // namespace cv
#endif
