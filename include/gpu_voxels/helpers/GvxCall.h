// gpu_voxels/helpers/GvxCall.h -- glue between the C++ facade and the C ABI (include/gvx.h): lazy gvx_init and the
// reference's error convention (log, carry on) for status codes.
#ifndef GPU_VOXELS_B200_HELPERS_GVXCALL_H
#define GPU_VOXELS_B200_HELPERS_GVXCALL_H

#include <cstdlib>
#include <iostream>

#include <gvx.h>

namespace gpu_voxels {

// The reference exits with 123 when no GPU can be initialised (GV/GpuVoxels.cpp:33-42); the C ABI never exits, the
// facade keeps the reference's behaviour for its callers.
inline void gvxEnsureInit()
{
  static const int rc = gvx_init(-1);
  if (rc != GVX_OK)
  {
    std::cerr << "<Error GpuVoxels> CUDA device initialisation failed: " << gvx_last_error() << std::endl;
    std::exit(123);
  }
}

// true on success; failures are logged like the reference's HANDLE_CUDA_ERROR (execution continues)
inline bool gvxCheck(int rc, const char* what)
{
  if (rc == GVX_OK) return true;
  std::cerr << "<Error GpuVoxels> " << what << ": " << gvx_last_error() << std::endl;
  return false;
}

}  // namespace gpu_voxels
#endif
