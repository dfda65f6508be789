// gpu_voxels/voxel/BitVoxel.h -- BitVector<256> / BitVectorVoxel as seen by callers of the public API
// (GV/helpers/BitVector.h:41-352, GV/voxel/BitVoxel.h:38-108): 32 bytes, bit k = byte k>>3, position k&7.
#ifndef GPU_VOXELS_B200_VOXEL_BITVOXEL_H
#define GPU_VOXELS_B200_VOXEL_BITVOXEL_H

#include <cstring>
#include <istream>
#include <ostream>
#include <string>
#include <stdint.h>

#include <gpu_voxels/helpers/common_defines.h>

namespace gpu_voxels {

template <std::size_t num_bits>
class BitVector
{
public:
  BitVector() { clear(); }
  void clear() { std::memset(m_bytes, 0, sizeof(m_bytes)); }
  bool isZero() const { for (std::size_t i = 0; i < sizeof(m_bytes); ++i) if (m_bytes[i]) return false; return true; }
  // any bit except bit 0 (eBVM_FREE)? (GV/helpers/BitVector.h:183-198)
  bool noneButEmpty() const
  {
    if (m_bytes[0] & 0xFE) return false;
    for (std::size_t i = 1; i < sizeof(m_bytes); ++i) if (m_bytes[i]) return false;
    return true;
  }
  bool getBit(uint32_t index) const { return (m_bytes[index >> 3] >> (index & 7)) & 1; }
  void setBit(uint32_t index) { m_bytes[index >> 3] |= uint8_t(1u << (index & 7)); }
  void clearBit(uint32_t index) { m_bytes[index >> 3] &= uint8_t(~(1u << (index & 7))); }
  uint8_t getByte(uint32_t index) const { return m_bytes[index >> 3]; }
  void setByte(uint32_t index, uint8_t data) { m_bytes[index >> 3] = data; }
  BitVector operator|(const BitVector& o) const { BitVector r; for (std::size_t i = 0; i < sizeof(m_bytes); ++i) r.m_bytes[i] = m_bytes[i] | o.m_bytes[i]; return r; }
  BitVector operator&(const BitVector& o) const { BitVector r; for (std::size_t i = 0; i < sizeof(m_bytes); ++i) r.m_bytes[i] = m_bytes[i] & o.m_bytes[i]; return r; }
  BitVector operator~() const { BitVector r; for (std::size_t i = 0; i < sizeof(m_bytes); ++i) r.m_bytes[i] = uint8_t(~m_bytes[i]); return r; }
  void operator|=(const BitVector& o) { for (std::size_t i = 0; i < sizeof(m_bytes); ++i) m_bytes[i] |= o.m_bytes[i]; }
  bool operator==(const BitVector& o) const { return std::memcmp(m_bytes, o.m_bytes, sizeof(m_bytes)) == 0; }
  uint8_t* bytes() { return m_bytes; }
  const uint8_t* bytes() const { return m_bytes; }
  // bit 0 first, bit num_bits-1 last, as the reference prints it (GV/helpers/BitVector.h:270-288;
  // GV/test/testing_voxelmap.cu:379)
  friend std::ostream& operator<<(std::ostream& os, const BitVector& v)
  {
    for (uint32_t i = 0; i < num_bits; ++i) os << (v.getBit(i) ? '1' : '0');
    return os;
  }
  // complete bit string, bit 0 first (GV/helpers/BitVector.h:290-312)
  friend std::istream& operator>>(std::istream& in, BitVector& v)
  {
    std::string s;
    in >> s;
    v.clear();
    for (uint32_t i = 0; i < num_bits && i < s.size(); ++i) if (s[i] == '1') v.setBit(i);
    return in;
  }
private:
  uint8_t m_bytes[(num_bits + 7) / 8];
};

template <std::size_t length>
class BitVoxel
{
public:
  BitVoxel() {}
  BitVector<length>& bitVector() { return m_bit_vector; }
  const BitVector<length>& bitVector() const { return m_bit_vector; }
  void insert(const BitVoxelMeaning voxel_meaning) { m_bit_vector.setBit(uint32_t(voxel_meaning)); }
  bool isOccupied(float) const { return !m_bit_vector.noneButEmpty(); }
  friend std::ostream& operator<<(std::ostream& os, const BitVoxel& v) { return os << v.m_bit_vector; }
  friend std::istream& operator>>(std::istream& in, BitVoxel& v) { return in >> v.m_bit_vector; }
private:
  BitVector<length> m_bit_vector;
};

typedef BitVoxel<BIT_VECTOR_LENGTH> BitVectorVoxel;  // GV/helpers/common_defines.h:99
static_assert(sizeof(BitVectorVoxel) == 32, "BitVectorVoxel must be 32 bytes");

}  // namespace gpu_voxels
#endif
