#include "cvshim2.h"
