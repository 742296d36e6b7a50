/* stand-in for <opencv2/core/core.hpp>: see cvshim.h (test infrastructure only) */
#include "../../cvshim.h"
