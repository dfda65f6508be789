// gpu_voxels/helpers/MathHelpers.h -- host helpers of the reference's API (GV/helpers/MathHelpers.h:34-76,
// GV/helpers/MathHelpers.cpp:29-160): the linear launch shape the reference's kernels are started with, and the linear
// interpolation of joint states the planners use between configurations.
// The kernels of libgvx.so size their own grids (multiples of the SM count); computeLinearLoad is
// kept for callers that launch kernels of their own on wrapped map memory with the reference's launch shape.
#ifndef GPU_VOXELS_B200_HELPERS_MATHHELPERS_H
#define GPU_VOXELS_B200_HELPERS_MATHHELPERS_H

#include <cassert>
#include <cstdint>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/common_defines.h>

namespace gpu_voxels {

#ifndef GPU_VOXELS_B200_LAUNCH_LIMITS
#define GPU_VOXELS_B200_LAUNCH_LIMITS
// GV/helpers/common_defines.h:38-40
const uint32_t cMAX_NR_OF_BLOCKS = 65535;
const uint32_t cMAX_THREADS_PER_BLOCK = 1024;
#endif

// ceil(n / 1024) blocks of 1024 threads, at most 65535 blocks (kernels then stride over the rest); n == 0 gives a
// 1 x 1 launch and an error message, because an empty grid is a CUDA launch error (MathHelpers.cpp:29-77)
inline void computeLinearLoad(const uint32_t nr_of_items, uint32_t* blocks, uint32_t* threads_per_block)
{
  if (nr_of_items == 0)
  {
    std::cerr << "<Error Gpu_voxels_helpers> Number of Items is 0. Blocks and Threads per Block is set to 1. "
                 "Size 0 would lead to a Cuda ERROR" << std::endl;
    *blocks = 1;
    *threads_per_block = 1;
    return;
  }
  *threads_per_block = cMAX_THREADS_PER_BLOCK;
  const uint32_t at_once = cMAX_NR_OF_BLOCKS * cMAX_THREADS_PER_BLOCK;
  *blocks = nr_of_items <= at_once ? (nr_of_items + cMAX_THREADS_PER_BLOCK - 1) / cMAX_THREADS_PER_BLOCK
                                   : cMAX_NR_OF_BLOCKS;
}

// value1 * (1 - ratio) + value2 * ratio; ratios outside [0, 1] extrapolate. The float overload evaluates
// (1.0 - ratio) and the sum in double and value2 * ratio in float, like the reference's expression does
// (MathHelpers.cpp:84-87), so the rounded result is the same.
inline float interpolateLinear(float value1, float value2, float ratio)
{
  const double weight1 = 1.0 - ratio;
  const float part2 = value2 * ratio;
  return float(value1 * weight1 + part2);
}
inline double interpolateLinear(double value1, double value2, double ratio)
{
  return value1 * (1.0 - ratio) + value2 * ratio;
}
inline std::vector<float> interpolateLinear(const std::vector<float>& joint_state1,
                                            const std::vector<float>& joint_state2, float ratio)
{
  assert(joint_state1.size() == joint_state2.size());
  std::vector<float> result(joint_state1.size());
  for (std::size_t i = 0; i < result.size(); ++i) result[i] = interpolateLinear(joint_state1[i], joint_state2[i], ratio);
  return result;
}
inline std::vector<double> interpolateLinear(const std::vector<double>& joint_state1,
                                             const std::vector<double>& joint_state2, double ratio)
{
  assert(joint_state1.size() == joint_state2.size());
  std::vector<double> result(joint_state1.size());
  for (std::size_t i = 0; i < result.size(); ++i) result[i] = interpolateLinear(joint_state1[i], joint_state2[i], ratio);
  return result;
}
// joint maps: every joint of joint_state1 is looked up in joint_state2 (MathHelpers.cpp:141-160)
inline std::map<std::string, float> interpolateLinear(const std::map<std::string, float>& joint_state1,
                                                      const std::map<std::string, float>& joint_state2, float ratio)
{
  assert(joint_state1.size() == joint_state2.size());
  std::map<std::string, float> result(joint_state1);
  for (std::map<std::string, float>::iterator joint = result.begin(); joint != result.end(); ++joint)
    joint->second = interpolateLinear(joint->second, joint_state2.at(joint->first), ratio);  // throws if missing
  return result;
}

}  // namespace gpu_voxels
#endif
