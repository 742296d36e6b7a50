/* stand-in for <opencv2/highgui/highgui.hpp>: see cvshim.h (test infrastructure only) */
#include "../../cvshim.h"
