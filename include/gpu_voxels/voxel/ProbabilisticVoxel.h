// gpu_voxels/voxel/ProbabilisticVoxel.h -- the 1-byte probabilistic voxel of the reference's API
// (GV/voxel/ProbabilisticVoxel.h:34-116, .hpp:38-108): int8 log-odds occupancy, -128 = unknown, clamped to [-127, 127].
// Host-side value type for code that reads maps back (ProbVoxelMap::getDeviceDataPtr, downloads, file dumps); the
// device kernels of libgvx.so work on the same byte layout.
#ifndef GPU_VOXELS_B200_VOXEL_PROBABILISTICVOXEL_H
#define GPU_VOXELS_B200_VOXEL_PROBABILISTICVOXEL_H

#include <cstdint>
#include <istream>

#include <gpu_voxels/helpers/common_defines.h>

namespace gpu_voxels {

class ProbabilisticVoxel
{
public:
  ProbabilisticVoxel() : m_occupancy(UNKNOWN_PROBABILITY) {}

  // saturating log-odds update in 32-bit arithmetic; an unknown voxel (-128) counts as -128 (.hpp:51-57)
  Probability updateOccupancy(const Probability occupancy)
  {
    int32_t sum = int32_t(m_occupancy) + int32_t(occupancy);
    if (sum < int32_t(MIN_PROBABILITY)) sum = int32_t(MIN_PROBABILITY);
    if (sum > int32_t(MAX_PROBABILITY)) sum = int32_t(MAX_PROBABILITY);
    m_occupancy = Probability(sum);
    return m_occupancy;
  }
  Probability& occupancy() { return m_occupancy; }
  const Probability& occupancy() const { return m_occupancy; }
  Probability getOccupancy() const { return m_occupancy; }

  // FREE -> -127, OCCUPIED / COLLISION -> 127, everything else -> unknown (.hpp:77-92)
  void insert(const BitVoxelMeaning voxel_meaning)
  {
    if (voxel_meaning == eBVM_FREE) m_occupancy = MIN_PROBABILITY;
    else if (voxel_meaning == eBVM_OCCUPIED || voxel_meaning == eBVM_COLLISION) m_occupancy = MAX_PROBABILITY;
    else m_occupancy = UNKNOWN_PROBABILITY;
  }
  // the RAW threshold: occupancy (as an integer) >= col_threshold (a float), not a probability (.hpp:94-98)
  bool isOccupied(float col_threshold) const { return m_occupancy >= col_threshold; }

  static ProbabilisticVoxel reduce(const ProbabilisticVoxel voxel, const ProbabilisticVoxel other_voxel)
  {
    ProbabilisticVoxel res = voxel;
    res.updateOccupancy(other_voxel.getOccupancy());
    return res;
  }
  struct reduce_op
  {
    ProbabilisticVoxel operator()(const ProbabilisticVoxel& a, const ProbabilisticVoxel& b) const { return reduce(a, b); }
  };

  template <typename T> friend T& operator<<(T& os, const ProbabilisticVoxel& dt)
  {
    os << dt.getOccupancy();
    return os;
  }
  friend std::istream& operator>>(std::istream& in, ProbabilisticVoxel& dt)
  {
    Probability tmp;
    in >> tmp;
    dt.occupancy() = tmp;
    return in;
  }

private:
  Probability m_occupancy;
};

static_assert(sizeof(ProbabilisticVoxel) == 1, "ProbabilisticVoxel is one byte, as in the reference");

}  // namespace gpu_voxels
#endif
