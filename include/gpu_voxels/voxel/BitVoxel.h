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

// Swept-volume IDs move `shift_size` (<= 56) bits towards bit 0; the four non-swept-volume bits of byte 0 end up
// cleared (performLeftShift, GV/helpers/BitVector.h:361-402). Host restatement of what gvx_map_shift_sv_ids /
// gvx_list_shift_sv_ids do per voxel on the device.
template <std::size_t num_bits>
inline void performLeftShift(BitVector<num_bits>& bit_vector, const uint8_t shift_size)
{
  const uint32_t n_bytes = uint32_t((num_bits + 7) / 8);
  uint8_t* v = bit_vector.bytes();
  uint64_t window = 0;
  for (uint32_t k = 0; k < 8 && k < n_bytes; ++k) window |= uint64_t(v[k]) << (8 * k);
  for (uint32_t k = 0; k < n_bytes; ++k)
  {
    uint8_t shifted = uint8_t(window >> shift_size);
    if (k == 0) shifted &= 0xF0;
    window >>= 8;
    if (k + 8 < n_bytes) window |= uint64_t(v[k + 8]) << 56;
    v[k] = shifted;
  }
}

// Do v1 and v2 share a swept-volume bit (4..253) within +-`bit_margin` positions, v1 displaced by `sv_offset`?
// (bitMarginCollisionCheck, GV/helpers/BitVector.h:416-471.) A 64-bit window slides over v2 -- its byte 0 without the
// four non-swept-volume bits -- and each visited byte of `collisions` (may be NULL) is OVERWRITTEN with the hits
// found for that byte of v1; sv_offset % 8 moves the first visited bit, sv_offset / 8 shifts the probe. Host
// restatement of the per-voxel test of gvx_collide_map_map_bitcheck; margins above 31 are undefined there.
template <std::size_t num_bits>
inline bool bitMarginCollisionCheck(const BitVector<num_bits>& v1, const BitVector<num_bits>& v2,
                                    BitVector<num_bits>* collisions, const uint8_t bit_margin, const uint32_t sv_offset)
{
  const uint8_t* a = v1.bytes();
  const uint8_t* b = v2.bytes();
  uint64_t window = uint64_t(b[0] & 0xF0) << 32;
  for (uint32_t k = 1; k < 4; ++k) window |= uint64_t(b[k]) << (32 + 8 * k);
  const uint32_t first_bit = uint8_t(sv_offset % 8), probe_shift = uint8_t(sv_offset / 8);
  bool any = false;
  for (uint32_t bit = uint32_t(eBVM_SWEPT_VOLUME_START) + first_bit; bit < uint32_t(eBVM_SWEPT_VOLUME_END); bit += 8)
  {
    uint64_t probe = uint64_t(a[bit >> 3]) << (32 - bit_margin + probe_shift);
    uint8_t hits = 0;
    for (uint32_t step = 0; step <= 2u * bit_margin; ++step, probe <<= 1)
      hits |= uint8_t((probe & window) >> (32 - bit_margin + step));
    if (collisions) collisions->setByte(bit, hits);
    any = any || hits != 0;
    window >>= 8;
    if (bit < num_bits - 32) window |= uint64_t(b[(bit + 32) >> 3]) << 56;
  }
  if (collisions) return !collisions->isZero();  // the reference looks at the whole output vector, old bytes included
  return any;
}

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
