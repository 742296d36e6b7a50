/* stand-in for <opencv2/features2d/features2d.hpp>: see cvshim.h (test infrastructure only) */
#include "../../cvshim.h"
