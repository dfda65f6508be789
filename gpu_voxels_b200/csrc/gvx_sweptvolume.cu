// gvx_sweptvolume.cu -- swept-volume collision check with a bit window, and the swept-volume ID shift.
//
//   k_collide_bitcheck  BitVoxelMap::collisionCheckBitvector(BitVoxelMap*, SVCollider(thr, window), types, sv_offset)
//                       (GV/voxelmap/BitVoxelMap.hpp:91-135, kernelCollideVoxelMapsBitvector
//                       GV/voxelmap/kernels/VoxelMapOperations.hpp:187-239, bitMarginCollisionCheck
//                       GV/helpers/BitVector.h:416-471); collideWithTypes(BitVectorVoxelMap*) is window 0, offset 0
//   k_shift_sv_ids      BitVoxelMap::shiftLeftSweptVolumeIDs (GV/voxelmap/BitVoxelMap.hpp:234-246, kernelShiftBitVector
//                       VoxelMapOperations.hpp:615-624, performLeftShift GV/helpers/BitVector.h:361-402)
//
// Both are HBM streaming kernels in the reference (64 B resp. 32 B + 32 B per voxel). Here the check reads the
// occupancy summary of `other` first (1 bit per voxel): a voxel of `other` without any bit above eBVM_FREE cannot
// contribute a swept-volume bit, so only candidate voxels are fetched at all.
#include "gvx_internal.h"
#include "gvx_reduce.cuh"
#include "gvx_sv.cuh"

namespace gvx {

template <bool MARK>
__device__ __forceinline__ void bitcheck_voxel(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, long long i,
                                               const VoxelBytes& vb, uint32_t margin, uint32_t sv_offset,
                                               uint32_t* occ_a_mark, uint32_t& cnt, uint32_t* acc) {
  VoxelBytes va;
  const uint4* pa = reinterpret_cast<const uint4*>(a + i * kBitVoxelBytes);
  va.q[0] = __ldcg(pa); va.q[1] = __ldcg(pa + 1);
  if ((va.w[0] | va.w[1] | va.w[2] | va.w[3] | va.w[4] | va.w[5] | va.w[6] | va.w[7]) == 0u) return;
  VoxelBytes out;
  out.q[0] = make_uint4(0, 0, 0, 0); out.q[1] = make_uint4(0, 0, 0, 0);
  if (!margin_check(va.b, vb.b, out.b, margin, sv_offset)) return;
  ++cnt;
#pragma unroll
  for (int k = 0; k < 8; ++k) acc[k] |= out.w[k];
  if (MARK && !(va.w[0] & (1u << 2))) {  // eBVM_COLLISION
    atomicOr(reinterpret_cast<uint32_t*>(a + i * kBitVoxelBytes), 1u << 2);
    if (occ_a_mark) summary_set(occ_a_mark, (uint64_t)i);
  }
}

// occ_b != nullptr: one summary word (32 voxels) per thread and step, candidates only; else one voxel per thread and
// step, `other` read first and `this` only where `other` has a swept-volume bit.
template <bool MARK>
__global__ void __launch_bounds__(kThreads)
k_collide_bitcheck(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, const uint32_t* __restrict__ occ_a,
                   const uint32_t* __restrict__ occ_b, uint32_t* occ_a_mark, long long n, uint32_t margin,
                   uint32_t sv_offset, ReduceScratch* scratch, ResultBlock* out_d, ResultBlock* out_h,
                   unsigned long long* count_out_d, PeerReduce pr) {
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long stride = (long long)gridDim.x * kThreads;
  if (occ_b) {
    const long long n_groups = (n + 31) >> 5;
    for (long long g = (long long)blockIdx.x * kThreads + threadIdx.x; g < n_groups; g += stride) {
      uint32_t cand = __ldg(occ_b + g);
      if (occ_a) cand &= __ldg(occ_a + g);
      if (g == n_groups - 1 && (n & 31)) cand &= (1u << (n & 31)) - 1u;
      while (cand) {
        const int bit = __ffs(cand) - 1;
        cand &= cand - 1;
        const long long i = (g << 5) + bit;
        VoxelBytes vb;
        const uint4* pb = reinterpret_cast<const uint4*>(b + i * kBitVoxelBytes);
        vb.q[0] = __ldg(pb); vb.q[1] = __ldg(pb + 1);
        if (((vb.w[0] & ~0xFu) | vb.w[1] | vb.w[2] | vb.w[3] | vb.w[4] | vb.w[5] | vb.w[6] | vb.w[7]) == 0u) continue;
        bitcheck_voxel<MARK>(a, b, i, vb, margin, sv_offset, occ_a_mark, cnt, acc);
      }
    }
  } else {
    for (long long i = (long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
      VoxelBytes vb;
      const BitVoxelWords w = ld_voxel_stream(b + i * kBitVoxelBytes);
#pragma unroll
      for (int k = 0; k < 8; ++k) vb.w[k] = w.w[k];
      if (((vb.w[0] & ~0xFu) | vb.w[1] | vb.w[2] | vb.w[3] | vb.w[4] | vb.w[5] | vb.w[6] | vb.w[7]) == 0u) continue;
      bitcheck_voxel<MARK>(a, b, i, vb, margin, sv_offset, occ_a_mark, cnt, acc);
    }
  }
  finish_block<true>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

cudaError_t launch_collide_bitcheck(const LaunchCtx& c, void* a, const void* b, int64_t n, uint32_t margin,
                                    uint32_t sv_offset, bool mark, const uint32_t* occ_a, const uint32_t* occ_b,
                                    uint32_t* occ_a_mark, unsigned long long* count_out_d) {
  if (n < 0) n = 0;
  const long long units = occ_b ? ((n + 31) >> 5) : n;
  long long grid = (units + kThreads - 1) / kThreads;
  const long long cap = (long long)c.sm_count * 8;
  if (grid > cap) grid = cap;
  if (grid > kMaxGrid) grid = kMaxGrid;
  if (grid < 1) grid = 1;
  ++*c.launches;
  if (mark)
    k_collide_bitcheck<true><<<(int)grid, kThreads, 0, c.stream>>>((uint8_t*)a, (const uint8_t*)b, occ_a, occ_b, occ_a_mark,
                                                                   n, margin, sv_offset, c.scratch_d, c.result_d,
                                                                   c.result_hmap_d, count_out_d, c.peer);
  else
    k_collide_bitcheck<false><<<(int)grid, kThreads, 0, c.stream>>>((uint8_t*)a, (const uint8_t*)b, occ_a, occ_b, occ_a_mark,
                                                                    n, margin, sv_offset, c.scratch_d, c.result_d,
                                                                    c.result_hmap_d, count_out_d, c.peer);
  return cudaGetLastError();
}

// out = (in >> shift) as a 256-bit little-endian number, bits 0..3 cleared; shift <= 56. One voxel per thread and
// step; untouched (all-zero) voxels are not written back. The summary word of every 32 voxels is re-derived.
__global__ void __launch_bounds__(kThreads)
k_shift_sv_ids(uint8_t* __restrict__ data, unsigned long long n, uint32_t shift, uint32_t* __restrict__ occ) {
  const unsigned long long stride = (unsigned long long)gridDim.x * kThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  const uint32_t q = shift >> 5, r = shift & 31u;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kThreads + threadIdx.x; i < n_round; i += stride) {
    bool occupied = false;
    if (i < n) {
      const BitVoxelWords v = ld_voxel(data + i * kBitVoxelBytes);
      uint32_t in[10];
#pragma unroll
      for (int k = 0; k < 8; ++k) in[k] = v.w[k];
      in[8] = in[9] = 0u;
      uint32_t any_in = 0, any_out = 0, o[8];
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        const uint32_t lo = q ? in[k + 1] : in[k], hi = q ? in[k + 2] : in[k + 1];
        o[k] = __funnelshift_r(lo, hi, r);
        any_in |= in[k];
      }
      o[0] &= ~0xFu;
#pragma unroll
      for (int k = 0; k < 8; ++k) any_out |= o[k];
      if (any_in) {
        uint4* p = reinterpret_cast<uint4*>(data + i * kBitVoxelBytes);
        p[0] = make_uint4(o[0], o[1], o[2], o[3]);
        p[1] = make_uint4(o[4], o[5], o[6], o[7]);
      }
      occupied = any_out != 0u;  // bits 0..3 are clear, so "any bit" == "any meaning except eBVM_FREE"
    }
    if (occ) {
      const uint32_t w = __ballot_sync(0xffffffffu, occupied);
      if ((threadIdx.x & 31) == 0) occ[i >> 5] = w;
    }
  }
}

cudaError_t launch_shift_sv_ids(const LaunchCtx& c, void* data, uint64_t n_voxels, uint32_t shift, uint32_t* occ) {
  unsigned long long grid = (n_voxels + kThreads - 1) / kThreads;
  const unsigned long long cap = (unsigned long long)c.sm_count * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  ++*c.launches;
  k_shift_sv_ids<<<(int)grid, kThreads, 0, c.stream>>>((uint8_t*)data, n_voxels, shift, occ);
  return cudaGetLastError();
}

}  // namespace gvx
