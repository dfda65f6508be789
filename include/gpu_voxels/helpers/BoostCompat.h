// gpu_voxels/helpers/BoostCompat.h -- the reference's public API hands out boost::shared_ptr (GV/GpuVoxelsMap.h:46,
// GV/GpuVoxels.h:62-66) and its examples name boost:: directly (boost::shared_ptr<ProbVoxelMap>,
// boost::dynamic_pointer_cast<...>). This build does not depend on Boost: the same roles are played by the std::
// types. Where Boost is NOT installed, the handful of names caller code uses are made available in namespace boost as
// aliases of the std:: ones, so that such code compiles unchanged; where Boost IS installed nothing is declared here
// (convert with std:: in that case). Define GPU_VOXELS_B200_NO_BOOST_COMPAT to switch the aliases off.
#ifndef GPU_VOXELS_B200_HELPERS_BOOSTCOMPAT_H
#define GPU_VOXELS_B200_HELPERS_BOOSTCOMPAT_H

#include <memory>

#if !defined(GPU_VOXELS_B200_NO_BOOST_COMPAT)
#if defined(__has_include)
#if !__has_include(<boost/shared_ptr.hpp>)
namespace boost {
using std::shared_ptr;
using std::weak_ptr;
using std::make_shared;
using std::dynamic_pointer_cast;
using std::static_pointer_cast;
using std::const_pointer_cast;
}  // namespace boost
#endif
#endif
#endif

#endif
