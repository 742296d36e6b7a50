#ifndef ORBX_SHIM_G2O_SIM3
#define ORBX_SHIM_G2O_SIM3
namespace g2o { struct Sim3 {}; }
#endif
