// gpu_voxels/voxel/SVCollider.h -- the swept-volume collider (GV/voxel/SVCollider.h:33-88, .hpp:33-137): int8
// thresholds for probabilistic voxels and a window of +-N bits around each swept-volume ID. The map-level collision
// arithmetic runs in libgvx.so (gvx_collide_map_map_bitcheck / gvx_collide_map_map_types); the collide() members here
// are the same predicates on host value types, for code that works on downloaded voxels.
#ifndef GPU_VOXELS_B200_VOXEL_SVCOLLIDER_H
#define GPU_VOXELS_B200_VOXEL_SVCOLLIDER_H

#include <cstddef>
#include <stdint.h>

#include <gpu_voxels/helpers/common_defines.h>
#include <gpu_voxels/voxel/BitVoxel.h>
#include <gpu_voxels/voxel/ProbabilisticVoxel.h>

namespace gpu_voxels {

class SVCollider
{
public:
  SVCollider() : m_threshold1(100), m_threshold2(100), m_type_range(10), m_threshold(100.0f) {}  // .hpp:33-38
  SVCollider(const float coll_threshold, const std::size_t window_size = 0)
    : m_threshold1(floatToProbability(coll_threshold)), m_threshold2(floatToProbability(coll_threshold)),
      m_type_range(window_size), m_threshold(coll_threshold) {}
  SVCollider(const Probability threshold1, const Probability threshold2, const std::size_t window_size = 0)
    : m_threshold1(threshold1), m_threshold2(threshold2), m_type_range(window_size),
      m_threshold((float(threshold1) - float(MIN_PROBABILITY)) / (float(MAX_PROBABILITY) - float(MIN_PROBABILITY))) {}

  // what the facade hands to the C ABI
  float getThreshold() const { return m_threshold; }
  std::size_t getWindowSize() const { return m_type_range; }

  // note the strict comparison on the second voxel: the reference's Prob x Prob swept-volume predicate (.hpp:54-57)
  bool collide(const ProbabilisticVoxel& v1, const ProbabilisticVoxel& v2) const
  {
    return v1.getOccupancy() >= m_threshold1 && v2.getOccupancy() > m_threshold2;
  }
  bool collide(const ProbabilisticVoxel& v1) const { return v1.getOccupancy() >= m_threshold1; }
  template <std::size_t length> bool collide(const ProbabilisticVoxel& v1, const BitVoxel<length>& v2) const
  {
    return v1.getOccupancy() >= m_threshold1 && !v2.bitVector().noneButEmpty();
  }
  template <std::size_t length> bool collide(const BitVoxel<length>& v1, const ProbabilisticVoxel& v2) const
  {
    return collide(v2, v1);  // i.e. against threshold 1, as there
  }
  // swept volume x swept volume: windowed bit test, `collisions` receives the colliding IDs (.hpp:84-90)
  template <std::size_t length>
  bool collide(const BitVoxel<length>& v1, const BitVoxel<length>& v2, BitVector<length>* collisions,
               const uint32_t sv_offset = 0) const
  {
    return bitMarginCollisionCheck<length>(v1.bitVector(), v2.bitVector(), collisions, uint8_t(m_type_range), sv_offset);
  }
  // swept volume x probabilistic map: every meaning of the colliding voxel is reported (.hpp:93-116)
  template <std::size_t length>
  bool collide(const BitVoxel<length>& v1, const ProbabilisticVoxel& v2, BitVector<length>* collisions,
               const uint32_t = 0) const
  {
    if (v2.getOccupancy() < m_threshold2 || v1.bitVector().noneButEmpty()) return false;
    *collisions |= v1.bitVector();
    return true;
  }
  template <std::size_t length>
  bool collide(const ProbabilisticVoxel& v1, const BitVoxel<length>& v2, BitVector<length>* collisions,
               const uint32_t = 0) const
  {
    if (v1.getOccupancy() < m_threshold1 || v2.bitVector().noneButEmpty()) return false;
    *collisions |= v2.bitVector();
    return true;
  }

  static Probability floatToProbability(const float val)
  {
    const float tmp = (val * (float(MAX_PROBABILITY) - float(MIN_PROBABILITY))) + MIN_PROBABILITY;
    return Probability(tmp);
  }

protected:
  Probability m_threshold1, m_threshold2;
  std::size_t m_type_range;
  float m_threshold;
};

}  // namespace gpu_voxels
#endif
