// gpu_voxels/voxel/DefaultCollider.h -- the reference's standard collision predicate as a host value type
// (GV/voxel/DefaultCollider.h:36-80, .hpp:34-98): thresholds in int8 log-odds, computed from a float in [0, 1] with
// C truncation (1.0 -> 127, 0.5 -> 0, 0.1 -> -101). The device kernels of libgvx.so evaluate the same predicates;
// this class is what callers pass to collisionCheckWithCounter(...) and what host code uses on downloaded voxels.
#ifndef GPU_VOXELS_B200_VOXEL_DEFAULTCOLLIDER_H
#define GPU_VOXELS_B200_VOXEL_DEFAULTCOLLIDER_H

#include <cmath>
#include <limits>

#include <gpu_voxels/helpers/common_defines.h>
#include <gpu_voxels/voxel/BitVoxel.h>
#include <gpu_voxels/voxel/ProbabilisticVoxel.h>

namespace gpu_voxels {

class DefaultCollider
{
public:
  DefaultCollider() : m_threshold1(100), m_threshold2(100), m_float(std::numeric_limits<float>::quiet_NaN()) {}
  DefaultCollider(const float coll_threshold)
    : m_threshold1(floatToProbability(coll_threshold)), m_threshold2(floatToProbability(coll_threshold)),
      m_float(coll_threshold) {}
  DefaultCollider(const Probability threshold1, const Probability threshold2)
    : m_threshold1(threshold1), m_threshold2(threshold2), m_float(std::numeric_limits<float>::quiet_NaN()) {}

  bool collide(const ProbabilisticVoxel& v1, const ProbabilisticVoxel& v2) const
  {
    return v1.getOccupancy() >= m_threshold1 && v2.getOccupancy() >= m_threshold2;
  }
  bool collide(const ProbabilisticVoxel& v1) const { return v1.getOccupancy() >= m_threshold1; }
  template <std::size_t length> bool collide(const ProbabilisticVoxel& v1, const BitVoxel<length>& v2) const
  {
    return v1.getOccupancy() >= m_threshold1 && !v2.bitVector().noneButEmpty();
  }
  template <std::size_t length> bool collide(const BitVoxel<length>& v1, const ProbabilisticVoxel& v2) const
  {
    return v2.getOccupancy() >= m_threshold2 && !v1.bitVector().noneButEmpty();
  }
  template <std::size_t length> bool collide(const BitVoxel<length>& v1, const BitVoxel<length>& v2) const
  {
    return !v1.bitVector().noneButEmpty() && !v2.bitVector().noneButEmpty();
  }

  // val * 254 - 127 in fp32, truncated towards zero (GV/voxel/DefaultCollider.hpp:94-98)
  static Probability floatToProbability(const float val)
  {
    const float tmp = (val * (float(MAX_PROBABILITY) - float(MIN_PROBABILITY))) + MIN_PROBABILITY;
    return Probability(tmp);
  }

  Probability getThreshold1() const { return m_threshold1; }
  Probability getThreshold2() const { return m_threshold2; }
  // A float that floatToProbability maps onto this collider's thresholds, for the C ABI (which takes the float the
  // reference's public collideWith takes). false if the two thresholds differ: the kernels apply ONE threshold.
  bool asFloatThreshold(float* thr) const
  {
    if (m_threshold1 != m_threshold2) return false;
    if (m_float == m_float) { *thr = m_float; return true; }
    float guess = (float(m_threshold1) - float(MIN_PROBABILITY)) / (float(MAX_PROBABILITY) - float(MIN_PROBABILITY));
    for (int step = 0; step < 64 && floatToProbability(guess) != m_threshold1; ++step)
      guess = std::nextafter(guess, floatToProbability(guess) < m_threshold1 ? 2.0f : -1.0f);
    if (floatToProbability(guess) != m_threshold1) return false;
    *thr = guess;
    return true;
  }

protected:
  Probability m_threshold1, m_threshold2;
  float m_float;
};

}  // namespace gpu_voxels
#endif
