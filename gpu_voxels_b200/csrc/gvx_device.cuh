// gvx_device.cuh -- device-side building blocks shared by the sm_100a kernels.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace gvx {

// programmatic dependent launch (gvx_internal.h pdl_launch): no-ops in a kernel that was launched plainly
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }


constexpr int kBitVoxelBytes = 32;
constexpr int kMapBit = 0;   // MT_BITVECTOR_VOXELMAP
constexpr int kMapProb = 4;  // MT_PROBAB_VOXELMAP

// One BitVectorVoxel (256 bits) as eight little-endian 32-bit words: bit k of the reference's
// uint8_t[32] (byte k>>3, position k&7; GV/helpers/BitVector.h:223-228) is bit k&31 of word k>>5.
struct __align__(32) BitVoxelWords {
  uint32_t w[8];
};

// 256-bit streaming load (sm_100a: LDG.E.256), read-only path, no L1 allocation.
__device__ __forceinline__ BitVoxelWords ld_voxel_stream(const void* p) {
  BitVoxelWords r;
  asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]),
                 "=r"(r.w[7])
               : "l"(p));
  return r;
}
// 256-bit coherent load (for data this kernel may also write)
__device__ __forceinline__ BitVoxelWords ld_voxel(const void* p) {
  BitVoxelWords r;
  asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]),
                 "=r"(r.w[7])
               : "l"(p));
  return r;
}
__device__ __forceinline__ uint4 ld_stream_u4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ int8_t ld_stream_i8(const void* p) {
  int r;
  asm volatile("ld.global.nc.L1::no_allocate.s8 %0, [%1];" : "=r"(r) : "l"(p));
  return (int8_t)r;
}

// !BitVector::noneButEmpty(): any bit other than bit 0 (eBVM_FREE) set (GV/helpers/BitVector.h:183-198)
__device__ __forceinline__ bool bit_occupied(const BitVoxelWords& v) {
  return (((v.w[0] & 0xFFFFFFFEu) | v.w[1] | v.w[2] | v.w[3]) | (v.w[4] | v.w[5] | v.w[6] | v.w[7])) != 0u;
}

// float -> voxel coordinate, GV/voxelmap/kernels/VoxelMapOperations.h:123-133:
// static_cast<uint32_t>(floor(c / side)). IEEE division (this file is never built with -use_fast_math), floor,
// saturating convert (NaN and negatives -> 0); compiles to F2I.U32.FLOOR.NTZ like the reference kernel.
__device__ __forceinline__ uint32_t coord_to_voxel(float c, float side) {
  return static_cast<uint32_t>(floorf(__fdiv_rn(c, side)));
}

// Matrix4f * Vector3f (GV/helpers/cuda_matrices.h:740-746) with the exact contraction nvcc 12.9 emits for the
// reference's kernelTransformCloud: fadd(a14, fma(a13, z, fma(a11, x, fmul(a12, y)))). The _rn intrinsics are
// never re-associated or re-contracted by the compiler.
__device__ __forceinline__ float affine_row(const float* __restrict__ a, float x, float y, float z) {
  float t = __fmul_rn(a[1], y);
  t = __fmaf_rn(a[0], x, t);
  t = __fmaf_rn(a[2], z, t);
  return __fadd_rn(a[3], t);
}

// Lower-bound search in the exclusive prefix of cloud sizes: cloud c with prefix[c] <= i < prefix[c+1].
// (The reference walks cloud_sizes[] linearly from 0 in every thread, VoxelMapOperations.hpp:396-406.)
__device__ __forceinline__ uint32_t find_cloud(const uint32_t* prefix, uint32_t n_clouds, uint32_t i) {
  uint32_t lo = 0, hi = n_clouds;
  while (hi - lo > 1) {
    const uint32_t mid = (lo + hi) >> 1;
    if (prefix[mid] <= i) lo = mid; else hi = mid;
  }
  return lo;
}

// set voxel idx's bit in a map summary (skips the atomic when the bit is already there)
__device__ __forceinline__ void summary_set(uint32_t* occ, uint64_t idx) {
  uint32_t* w = occ + (idx >> 5);
  const uint32_t b = 1u << (idx & 31);
  if (!(__ldcg(w) & b)) atomicOr(w, b);
}

__device__ __forceinline__ uint64_t warp_sum_u32(uint32_t v) { return (uint64_t)__reduce_add_sync(0xffffffffu, v); }

// Scratch shared by every reducing kernel of one context: accumulators the CTAs add into with one atomic each, and a
// self-resetting ticket (atomicInc wraps to 0 when the last CTA arrives). The last CTA takes the accumulators with
// atomicExch(.., 0), so no memset is needed between launches.
struct ResultBlock {
  unsigned long long count;
  uint32_t mask[8];
  uint32_t flag;
  uint32_t pad;
  unsigned long long per_meaning[256];
};
// Fused all-reduce of a collision count over NVLink peer memory (one process per GPU): every rank owns a mailbox in
// its HBM with 2 (epoch parity) x kMaxPeers slots of {value, epoch}; peers[r] is rank r's mailbox mapped into this
// process (CUDA IPC). The last CTA of a counting kernel stores its rank's count into its slot of EVERY mailbox
// (st.global on mapped peer pointers), then waits until all slots of its own mailbox carry the current epoch and
// sums them -- the collective is part of the kernel, no NCCL launch follows. world <= 1 disables it.
constexpr int kMaxPeers = 16;
struct PeerSlot { unsigned long long value; unsigned long long epoch; };
// The mailbox allocation of a rank: the scalar slots [2 (epoch parity)][kMaxPeers], then a vector area
// [2][kMaxPeers (source rank)][kPeerVecWords] of packed words (epoch16 << 48 | value48) for the 257-word all-reduce of
// list x map per-meaning counts (total + 256 counts).
constexpr int kPeerVecWords = 257;
constexpr size_t kPeerVecOffset = 2 * kMaxPeers * sizeof(PeerSlot);
constexpr size_t kPeerMailboxBytes = kPeerVecOffset + 2ull * kMaxPeers * kPeerVecWords * sizeof(unsigned long long);
struct PeerReduce {
  PeerSlot* peers[kMaxPeers];
  int rank;
  int world;
  unsigned long long epoch;  // > 0, identical on all ranks for one collective, strictly increasing
  // the same for the collectives that use the vector area (a sequence of its own: a slot of the area is rewritten every
  // second VECTOR collective, so its 16-bit tag can never meet a stale word of 65 536 collectives ago, however many
  // scalar-only collectives lie in between); 0: no vector part
  unsigned long long vec_epoch;
};
// Sensor-frame dispatch over NVLink peer memory (gvx_update_logodds_dispatch, gvx_dispatch.cu; one process per GPU,
// z-slab maps). Every rank owns ONE allocation, mapped into all peers through CUDA IPC: a DispatchHead followed by
// the inboxes, uint32 entries[2 (epoch parity)][world (source rank)][capacity]. An entry is the slab-local linear
// index of a voxel that receives one log-odds hit.
//   ready[p][s]  written by source s into the DESTINATION's head: epoch16 << 48 | number of entries s stored into
//                inbox[p][s] for this epoch (a release after the entries);
//   ack[d]       written by destination d into the SOURCE's head: the last epoch whose entries d has consumed; the
//                source must see ack[d] + 2 >= epoch before it overwrites inbox[epoch & 1] at d (credit);
//   counters, ticket: local bookkeeping of the owner's kernel, self-resetting.
struct DispatchHead {
  unsigned long long ready[2][kMaxPeers];
  unsigned long long ack[kMaxPeers];
  unsigned int counters[kMaxPeers];
  unsigned int ticket_sent, pad_[3];
};
constexpr size_t kDispatchHeadBytes = 1024;
static_assert(sizeof(DispatchHead) <= kDispatchHeadBytes, "dispatch head grew beyond its slot");
struct PeerDispatch {
  unsigned char* base[kMaxPeers];  // rank r's allocation as mapped into this process
  uint32_t z_bounds[kMaxPeers + 1];  // rank r owns global z in [z_bounds[r], z_bounds[r + 1])
  uint32_t capacity;               // entries per (parity, source) inbox
  int rank;
  int world;
  unsigned long long epoch;        // > 0, identical on all ranks for one collective call, strictly increasing
};
constexpr int kMaxGrid = 148 * 16;  // upper bound on CTAs of a reducing kernel
struct ReduceScratch {
  unsigned long long total;  // low 48 bits: sum of the CTAs' counts, above: arrivals (finish_block); zeroed by the last CTA
  uint32_t mask[8];          // OR of the CTAs' meaning masks; same protocol
  unsigned int ticket;       // k_insert's flag protocol: atomicInc wraps to 0 when the last CTA arrives
  unsigned int flag_accum;
  unsigned int any_flag;     // early exit of the boolean collisionCheck: raised by the first CTA that finds a collision,
                             // lowered by the last CTA of that launch (finish_block)
  unsigned int pad_[3];
  // per-meaning histogram of list x map collides (k_collide_list_map<MODE 3>): every CTA adds its 256 bins, the last CTA
  // reads, publishes and ZEROES them -- like `total`, so no memset precedes a launch
  unsigned long long pm_acc[256];
};

}  // namespace gvx
