#ifndef ORBX_SHIM_G2O_SE3
#define ORBX_SHIM_G2O_SE3
namespace g2o { struct SE3Quat {}; }
#endif
