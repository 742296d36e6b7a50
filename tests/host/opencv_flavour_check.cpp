// Compile-only check of the -DORBX_WITH_OPENCV overloads of orbx.hpp against a cv namespace
// (here the test-only stand-in oracle/ref_shim/cvshim.h; a maintainer includes the real OpenCV).
#include "cvshim.h"
#define ORBX_WITH_OPENCV
#include "orbx.hpp"
void use(orbx::ORBextractor& ex, const cv::Mat& im, std::vector<cv::KeyPoint>& k, cv::Mat& d) { ex(im, cv::Mat(), k, d); }
