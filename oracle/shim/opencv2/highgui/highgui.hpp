/* Stand-in for <opencv2/highgui/highgui.hpp> -- TEST INFRASTRUCTURE: nothing on the path uses highgui. */
#include "../core/core.hpp"
