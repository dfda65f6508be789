// Shim (test infrastructure only): boost::shared_ptr as named by the reference's robot headers. Boost is absent here.
#ifndef GVX_SHIM_BOOST_SHARED_PTR_HPP
#define GVX_SHIM_BOOST_SHARED_PTR_HPP
#include <memory>
namespace boost { template <class T> using shared_ptr = std::shared_ptr<T>; }
#endif
