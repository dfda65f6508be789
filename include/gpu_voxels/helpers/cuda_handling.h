// gpu_voxels/helpers/cuda_handling.h -- the device queries of the reference's helper library
// (GV/helpers/cuda_handling.h:42-61, .cu:66-149) over gvx_device_info. The process drives ONE device (the one gvx_init
// selected, one process per GPU), so the reports cover that device; the kernel-side helpers of the original header
// (cuPrintDeviceArray ..., HANDLE_CUDA_ERROR) have no counterpart: callers of the facade do not see the CUDA runtime.
#ifndef GPU_VOXELS_B200_HELPERS_CUDA_HANDLING_H
#define GPU_VOXELS_B200_HELPERS_CUDA_HANDLING_H

#include <sstream>
#include <string>

#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/common_defines.h>

namespace gpu_voxels {

// false (and no exit) when no usable device is there; GpuVoxels' constructor exits with 123 in that case, as the
// reference does (GV/GpuVoxels.cpp:33-42)
inline bool cuTestAndInitDevice() { return gvx_init(-1) == GVX_OK; }

inline bool cuGetNrOfDevices(int* nr_of_devices)
{
  *nr_of_devices = cuTestAndInitDevice() ? 1 : 0;  // the device of this process
  return *nr_of_devices > 0;
}

inline std::string getDeviceInfos()
{
  std::stringstream out;
  int sm_count = 0;
  uint64_t free_bytes = 0, total_bytes = 0;
  char name[64] = {0};
  if (!cuTestAndInitDevice() || gvx_device_info(&sm_count, &free_bytes, &total_bytes, name) != GVX_OK)
  {
    out << "No CUDA devices found." << std::endl;
    return out.str();
  }
  out << "Device Information of GPU 0" << std::endl
      << "Model: " << name << std::endl
      << "Multi Processor Count: " << sm_count << std::endl
      << "Global Memory: " << cBYTE2MBYTE * float(total_bytes) << " MB" << std::endl << std::endl;
  return out.str();
}

inline std::string getDeviceMemoryInfo()
{
  std::stringstream out;
  uint64_t free_bytes = 0, total_bytes = 0;
  gvxEnsureInit();
  gvxCheck(gvx_sync(), "getDeviceMemoryInfo");
  gvxCheck(gvx_device_info(NULL, &free_bytes, &total_bytes, NULL), "getDeviceMemoryInfo");
  out << "Device memory status:" << std::endl;
  out << "-----------------------------------" << std::endl;
  out << "total memory (MB)  : " << float(total_bytes) * cBYTE2MBYTE << std::endl;
  out << "free  memory (MB)  : " << float(free_bytes) * cBYTE2MBYTE << std::endl;
  out << "used  memory (MB)  : " << float(total_bytes - free_bytes) * cBYTE2MBYTE << std::endl;
  out << "-----------------------------------" << std::endl;
  return out.str();
}

}  // namespace gpu_voxels
#endif
