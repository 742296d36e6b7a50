/* stand-in: the reference only needs boost::shared_ptr / boost::make_shared for its PCL cloud member (src/KeyFrame.cc:93) */
#ifndef ORBX_SHIM_BOOST_MAKE_SHARED
#define ORBX_SHIM_BOOST_MAKE_SHARED
#include <memory>
namespace boost {
template <typename T> using shared_ptr = std::shared_ptr<T>;
template <typename T, typename... A> shared_ptr<T> make_shared(A&&... a) { return std::make_shared<T>(std::forward<A>(a)...); }
}
#endif
