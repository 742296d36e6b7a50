/* stand-in for the PCL types that include/KeyFrame.h:126 and src/KeyFrame.cc:703-728 name (dense mapping: not on the pinned paths) */
#ifndef ORBX_SHIM_PCL_POINT_TYPES
#define ORBX_SHIM_PCL_POINT_TYPES
#include <cstdint>
#include <vector>
#include "../boost/make_shared.hpp"
namespace pcl {
struct PointXYZRGBA { float x, y, z; uint8_t b, g, r, a; };
template <typename P> struct PointCloud {
    typedef boost::shared_ptr<PointCloud<P>> Ptr;
    std::vector<P> points;
    unsigned width = 0, height = 0;
    bool is_dense = true;
    void push_back(const P& p) { points.push_back(p); }
    size_t size() const { return points.size(); }
};
}
#endif
