/* Stand-in for <opencv/cv.h> (include/ORBextractor.h:26) -- TEST INFRASTRUCTURE (see opencv2/core/core.hpp in this tree). */
#include "../opencv2/opencv.hpp"
