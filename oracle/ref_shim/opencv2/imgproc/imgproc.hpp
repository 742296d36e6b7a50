/* stand-in for <opencv2/imgproc/imgproc.hpp>: see cvshim.h (test infrastructure only) */
#include "../../cvshim.h"
