// gvx_reduce.cuh -- CTA-level finish shared by every counting kernel (map x map, list x map).
#pragma once
#include "gvx_device.cuh"

namespace gvx {

constexpr int kThreads = 256;  // block size of every reducing kernel

// One WARP: lane r publishes `mine` to rank r's mailbox and waits for rank r's value of this epoch in its own mailbox;
// the lanes' values are then summed. Value (48 bits) and epoch (16 bits) travel in ONE 64-bit word, so a single relaxed
// system-scope store per peer is enough -- nothing else written by this kernel is read through the mailbox -- and the
// world - 1 NVLink stores and polls overlap instead of queueing behind one thread's release fences.
// The wait is bounded (~4 s of SM clocks): a rank that never arrives yields ~0ull instead of a hung GPU.
__device__ __forceinline__ unsigned long long peer_allreduce_sum(unsigned long long mine, const PeerReduce& pr) {
  const int lane = threadIdx.x & 31;
  const int par = (int)(pr.epoch & 1ull);
  const unsigned long long tag = (pr.epoch & 0xFFFFull) << 48;
  unsigned long long v = 0;
  bool ok = true;
  if (lane < pr.world) {
    PeerSlot* dst = pr.peers[lane] + par * kMaxPeers + pr.rank;
    const unsigned long long word = tag | (mine & 0xFFFFFFFFFFFFull);
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(&dst->value), "l"(word) : "memory");
    const PeerSlot* src = pr.peers[pr.rank] + par * kMaxPeers + lane;
    const long long t0 = clock64();
    for (;;) {
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(&src->value) : "memory");
      if ((v >> 48) == (pr.epoch & 0xFFFFull)) break;
      if (clock64() - t0 > 8000000000ll) { ok = false; break; }
    }
    v &= 0xFFFFFFFFFFFFull;
  }
  const bool all_ok = __all_sync(0xffffffffu, ok);
  unsigned int lo = (unsigned int)v, hi = (unsigned int)(v >> 32);
  unsigned long long sum = 0;
  for (int r = 0; r < pr.world; ++r)  // world <= 16: a handful of shuffles, no 64-bit warp reduce needed
    sum += ((unsigned long long)__shfl_sync(0xffffffffu, hi, r) << 32) | __shfl_sync(0xffffffffu, lo, r);
  return all_ok ? sum : ~0ull;
}

// The same for a vector of up to kPeerVecWords values, by ALL kThreads threads of one CTA: thread t owns word t (thread 0
// also word 256). Every word is one packed 64-bit value (epoch16 << 48 | value48) stored into every peer's vector area
// with a relaxed system-scope store and polled in the own area until it carries the current epoch: no flag, no fence,
// the n_words x world stores and polls of a CTA overlap. v (in registers) is replaced by the sum over the ranks;
// a rank that never arrives yields ~0ull (bounded spin).
__device__ __forceinline__ unsigned long long peer_allreduce_word(unsigned long long mine, int word, const PeerReduce& pr) {
  const int par = (int)(pr.vec_epoch & 1ull);
  const unsigned long long tag = (pr.vec_epoch & 0xFFFFull) << 48;
  const unsigned long long packed = tag | (mine & 0xFFFFFFFFFFFFull);
  for (int r = 0; r < pr.world; ++r) {
    unsigned long long* dst = reinterpret_cast<unsigned long long*>(reinterpret_cast<unsigned char*>(pr.peers[r]) + kPeerVecOffset) +
                              ((size_t)par * kMaxPeers + pr.rank) * kPeerVecWords + word;
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(dst), "l"(packed) : "memory");
  }
  unsigned long long sum = 0;
  bool ok = true;
  const long long t0 = clock64();
  for (int r = 0; r < pr.world && ok; ++r) {
    const unsigned long long* src = reinterpret_cast<const unsigned long long*>(reinterpret_cast<const unsigned char*>(pr.peers[pr.rank]) + kPeerVecOffset) +
                                    ((size_t)par * kMaxPeers + r) * kPeerVecWords + word;
    unsigned long long v;
    for (;;) {
      asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(src) : "memory");
      if ((v >> 48) == (pr.vec_epoch & 0xFFFFull)) break;
      if (clock64() - t0 > 8000000000ll) { ok = false; break; }
    }
    sum += v & 0xFFFFFFFFFFFFull;
  }
  return ok ? sum : ~0ull;
}

// Block-level finish shared by all counting kernels: reduce (count, mask) over the CTA, add it into the context's
// accumulator together with an arrival ticket, one atomic per CTA (8 more for a mask); the last CTA to arrive has the total,
// optionally all-reduces the count over the peers, and publishes the result to device memory and to the mapped pinned
// host block. Returns true in the last CTA.
constexpr int kTicketShift = 48;
constexpr unsigned long long kTicketOne = 1ull << kTicketShift;

template <bool TYPES>
__device__ __forceinline__ bool finish_block(uint32_t my_count, const uint32_t (&acc)[8], ReduceScratch* s,
                                             ResultBlock* out_d, ResultBlock* out_h,
                                             unsigned long long* count_out_d, const PeerReduce& pr) {
  __shared__ unsigned long long sh_cnt[kThreads / 32];
  __shared__ uint32_t sh_mask[8];
  __shared__ bool sh_last;
  __shared__ unsigned long long sh_tot;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (TYPES && threadIdx.x < 8) sh_mask[threadIdx.x] = 0u;
  const unsigned long long wsum = warp_sum_u32(my_count);
  if (lane == 0) sh_cnt[warp] = wsum;
  __syncthreads();
  if (TYPES) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      const uint32_t m = __reduce_or_sync(0xffffffffu, acc[k]);
      if (lane == 0 && m) atomicOr(&sh_mask[k], m);
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    unsigned long long c = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) c += sh_cnt[w];
    if (TYPES) {
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (sh_mask[k]) atomicOr(&s->mask[k], sh_mask[k]);
      __threadfence();  // the mask bits are in L2 before this CTA's arrival is
    }
    // count and arrival ticket travel in ONE 64-bit atomic (count in the low 48 bits, arrivals above; grids are
    // <= kMaxGrid CTAs, counts <= the voxels of a 180 GB map): the tail of a reducing kernel is one L2 round trip per
    // CTA, and the last arrival already holds the total
    const unsigned long long old = atomicAdd(&s->total, c + kTicketOne);
    sh_last = (unsigned)(old >> kTicketShift) + 1u == gridDim.x;
    if (sh_last) {
      sh_tot = (old & (kTicketOne - 1ull)) + c;
      s->total = 0ull;  // nobody else arrives in this launch; the next launch starts behind a kernel boundary
      s->any_flag = 0u; // every CTA has left its loop: nobody reads or raises the early-exit flag any more
      if (TYPES) __threadfence();
    }
  }
  __syncthreads();
  if (sh_last && threadIdx.x < 32) {  // block-uniform branch; the first warp publishes
    unsigned long long tot = sh_tot;
    if (pr.world > 1) tot = peer_allreduce_sum(tot, pr);
    if (threadIdx.x == 0) {
      // out_h == nullptr: an asynchronous call whose result stays on the device -- no store into mapped host memory
      // and no system-scope fence (a PCIe round trip) at the end of the kernel
      out_d->count = tot;
      out_d->flag = tot != 0;
      if (out_h) { out_h->count = tot; out_h->flag = tot != 0; }
      if (count_out_d) *count_out_d = tot;
      if (TYPES) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint32_t m = atomicExch(&s->mask[k], 0u);
          out_d->mask[k] = m;
          if (out_h) out_h->mask[k] = m;
        }
      }
      if (out_h) __threadfence_system();
    }
  }
  __syncthreads();
  return sh_last;
}

}  // namespace gvx
