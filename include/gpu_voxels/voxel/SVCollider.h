// gpu_voxels/voxel/SVCollider.h -- the swept-volume collider's parameters (GV/voxel/SVCollider.h:33-88, .hpp:34-53):
// a collision threshold for probabilistic voxels and a window of bits around each swept-volume ID. The collision
// arithmetic itself lives in libgvx.so (gvx_collide_map_map_bitcheck / gvx_collide_map_map_types).
#ifndef GPU_VOXELS_B200_VOXEL_SVCOLLIDER_H
#define GPU_VOXELS_B200_VOXEL_SVCOLLIDER_H

#include <cstddef>

namespace gpu_voxels {

class SVCollider
{
public:
  SVCollider() : m_threshold(100.0f), m_type_range(10) {}  // GV/voxel/SVCollider.hpp:34-39
  SVCollider(const float coll_threshold, const std::size_t window_size = 0)
    : m_threshold(coll_threshold), m_type_range(window_size) {}
  float getThreshold() const { return m_threshold; }
  std::size_t getWindowSize() const { return m_type_range; }

private:
  float m_threshold;
  std::size_t m_type_range;
};

}  // namespace gpu_voxels
#endif
