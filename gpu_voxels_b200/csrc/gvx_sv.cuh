// gvx_sv.cuh -- the swept-volume bit-window check on one voxel pair, shared by the map x map and list x list kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gvx {

// The reference's window arithmetic on one voxel pair, byte for byte: a 64-bit window holds 4 bytes of `other` above
// bit 32 (its first byte without the four non-swept-volume bits) and what has been shifted down below; `low` moves the
// first visited bit (and can skip byte 0), `high` shifts the probe byte of `this` -- the reference names the two the
// other way round. a8/b8: the two voxels as bytes; out: collision bits, zero on entry; bytes not visited stay zero.
__device__ __forceinline__ bool margin_check(const uint8_t* a8, const uint8_t* b8, uint8_t* out, uint32_t margin,
                                             uint32_t sv_offset) {
  unsigned long long window = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) window += (unsigned long long)(k == 0 ? (b8[0] & 0xF0) : b8[k]) << (32 + 8 * k);
  const uint32_t low = sv_offset & 7u, high = (sv_offset >> 3) & 0xFFu;
  uint32_t any = 0;
  for (uint32_t i = 4 + low; i < 254; i += 8) {
    uint32_t o = 0;
    unsigned long long probe = (unsigned long long)a8[i >> 3] << (32 - margin + high);
    for (uint32_t s = 0; s <= 2 * margin; ++s) {
      o |= (uint32_t)((probe & window) >> (32 - margin + s));
      probe <<= 1;
    }
    o &= 0xFFu;
    out[i >> 3] = (uint8_t)o;
    any |= o;
    window >>= 8;
    if (i < 224) window += (unsigned long long)b8[(i + 32) >> 3] << 56;
  }
  return any != 0;
}

union VoxelBytes {
  uint4 q[2];
  uint32_t w[8];
  uint8_t b[32];
};

}  // namespace gvx
