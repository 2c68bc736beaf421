/* Stand-in for <opencv2/opencv.hpp> -- TEST INFRASTRUCTURE (see opencv2/core/core.hpp in this tree). */
#include "core/core.hpp"
#include "features2d/features2d.hpp"
#include "imgproc/imgproc.hpp"
