/* stand-in for <opencv/cv.h>: see cvshim.h (test infrastructure only) */
#include "../cvshim.h"
