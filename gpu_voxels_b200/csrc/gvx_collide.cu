// gvx_collide.cu -- collideWith map x map for sm_100a.
//
// Replaces kernelCollideVoxelMapsDebug / kernelCollideVoxelMaps / kernelCollideVoxelMapsBitvector
// (GV/voxelmap/kernels/VoxelMapOperations.hpp:78-239) and the host tail of
// collisionCheckWithCounterRelativeTransform (GV/voxelmap/TemplateVoxelMap.hpp:486-519: 128 KiB D2H of
// uint16 block counters + a 65k-iteration host sum).
//
// Shape: a pure HBM stream over two grids. One BitVectorVoxel = one 256-bit load (LDG.E.256, sm_100a), so a
// warp reads 1 KiB contiguous per instruction; probabilistic voxels are byte loads of one 32-B sector per warp
// (Bit x Prob) or 128-bit loads of 16 voxels per thread (Prob x Prob). Tests and popcounts stay in registers,
// counts go warp-REDUX -> shared -> one partial per CTA; the last CTA to arrive (self-resetting ticket) sums
// the partials and writes the 64-bit result to device memory and to a mapped pinned host word, so one launch
// and one stream sync deliver the answer. Algorithmic traffic: 33 / 33 / 2 / 64 B per voxel
// (Bit x Prob, Prob x Bit, Prob x Prob, Bit x Bit).
#include "gvx_internal.h"
#include "gvx_reduce.cuh"

namespace gvx {

constexpr int kUnroll = 4;
constexpr int kCtasPerSm = 4;

template <bool COHERENT>
__device__ __forceinline__ BitVoxelWords load_bit(const uint8_t* p) {
  if (COHERENT) {
    BitVoxelWords r;
    asm volatile("ld.global.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(r.w[0]), "=r"(r.w[1]), "=r"(r.w[2]), "=r"(r.w[3]), "=r"(r.w[4]), "=r"(r.w[5]), "=r"(r.w[6]),
                   "=r"(r.w[7])
                 : "l"(p));
    return r;
  }
  return ld_voxel_stream(p);
}

// Early exit of the boolean collisionCheck (kernelCollideVoxelMaps, GV/voxelmap/kernels/VoxelMapOperations.hpp:78-120:
// any-collision flag, no write-back): the first thread that counts a collision raises ReduceScratch::any_flag, every
// thread looks at it once per loop iteration and leaves. The count that reaches finish_block is then a lower bound
// (> 0 iff there is a collision), which is all gvx_collide_map_map_any reports. The last CTA lowers the flag again.
__device__ __forceinline__ bool any_raised(const ReduceScratch* s) {
  unsigned int v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(&s->any_flag) : "memory");
  return v != 0u;
}
__device__ __forceinline__ void any_raise(ReduceScratch* s) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(&s->any_flag), "r"(1u) : "memory");
}

// `thr` of every collide kernel carries DefaultCollider's TWO int8 thresholds (GV/voxel/DefaultCollider.hpp:39-49:
// DefaultCollider(Probability t1, Probability t2); the float form sets both to the same value): t1, applied to a
// ProbabilisticVoxel on side a, in bits 0..7; t2 for side b in bits 8..15.
__device__ __forceinline__ int thr_of_a(int thr) { return (int)(int8_t)(thr & 0xFF); }
__device__ __forceinline__ int thr_of_b(int thr) { return (int)(int8_t)((thr >> 8) & 0xFF); }
__device__ __forceinline__ uint32_t thr4(int t) { return (uint32_t)(uint8_t)t * 0x01010101u; }

// ---- generic kernel: one voxel per thread per step, at least one side is a BitVectorVoxel map (or an
// unaligned Prob x Prob fallback). AT/BT are MapType values. DefaultCollider::collide semantics
// (GV/voxel/DefaultCollider.hpp:51-80); TYPES adds SVCollider's mask |= a (GV/voxel/SVCollider.hpp:96-106).
template <int AT, int BT, bool MARK, bool TYPES, bool ANY = false>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
k_collide_voxelwise(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, long long i_begin, long long i_end,
                    long long off, int thr, ReduceScratch* scratch, ResultBlock* out_d, ResultBlock* out_h,
                    unsigned long long* count_out_d, PeerReduce pr) {
  constexpr int U = (AT == kMapBit && BT == kMapBit) ? 2 : kUnroll;  // 128 B in flight per thread either way
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const long long stride = (long long)gridDim.x * kThreads * U;
  for (long long base = i_begin + (long long)blockIdx.x * kThreads * U; base < i_end; base += stride) {
    if (ANY) {
      if (cnt) { any_raise(scratch); break; }
      if (any_raised(scratch)) break;
    }
    BitVoxelWords va[U], vb[U];
    int pa[U], pb[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + u * kThreads + threadIdx.x;
      ok[u] = i < i_end;
      if (ok[u]) {
        if (AT == kMapBit) va[u] = load_bit<MARK>(a + (size_t)(i + off) * kBitVoxelBytes);
        else pa[u] = MARK ? (int)((const volatile int8_t*)a)[i + off] : (int)ld_stream_i8(a + (i + off));
        if (BT == kMapBit) vb[u] = ld_voxel_stream(b + (size_t)i * kBitVoxelBytes);
        else pb[u] = (int)ld_stream_i8(b + i);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!ok[u]) continue;
      const long long i = base + u * kThreads + threadIdx.x;
      const bool occ_a = (AT == kMapBit) ? bit_occupied(va[u]) : (pa[u] >= thr_of_a(thr));
      const bool occ_b = (BT == kMapBit) ? bit_occupied(vb[u]) : (pb[u] >= thr_of_b(thr));
      if (occ_a && occ_b) {
        ++cnt;
        if (TYPES && AT == kMapBit) {
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] |= va[u].w[k];
        }
        if (MARK) {  // a[i].insert(eBVM_COLLISION): bit 2 / occupancy 127; write only when it changes something
          if (AT == kMapBit) {
            if (!(va[u].w[0] & 4u)) *(uint32_t*)(a + (size_t)(i + off) * kBitVoxelBytes) = va[u].w[0] | 4u;
          } else {
            if (pa[u] != 127) ((int8_t*)a)[i + off] = 127;
          }
        }
      }
    }
  }
  if (ANY && cnt) any_raise(scratch);
  finish_block<TYPES>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

// ---- Prob x Prob, 16 voxels per 128-bit load. Requires (a + off) and b 16-byte aligned for i % 16 == 0.
// Byte-wise signed >= via the SIMD video instructions; popcount of the 0xFF lanes / 8.
template <bool MARK, bool ANY = false>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
k_collide_prob_prob(int8_t* __restrict__ a, const int8_t* __restrict__ b, long long i_begin, long long i_end,
                    long long off, int thr, ReduceScratch* scratch, ResultBlock* out_d, ResultBlock* out_h,
                    unsigned long long* count_out_d, PeerReduce pr) {
  uint32_t cnt = 0;
  const uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const uint32_t t4a = thr4(thr_of_a(thr)), t4b = thr4(thr_of_b(thr));
  const long long vb = (i_begin + 15) & ~15ll, ve = i_end & ~15ll;
  if (blockIdx.x == 0) {  // unaligned head and tail, at most 30 voxels
    const long long head = (vb < i_end ? vb : i_end) - i_begin;
    const long long tail = ve >= vb ? i_end - ve : 0;  // ve == vb: an empty aligned middle still has a tail
    for (long long k = threadIdx.x; k < head + tail; k += kThreads) {
      const long long i = k < head ? i_begin + k : ve + (k - head);
      const int pa = a[i + off], pb = b[i];
      if (pa >= thr_of_a(thr) && pb >= thr_of_b(thr)) {
        ++cnt;
        if (MARK && pa != 127) a[i + off] = 127;
      }
    }
  }
  const long long nvec = ve > vb ? (ve - vb) >> 4 : 0;
  const uint4* av = reinterpret_cast<const uint4*>(a + vb + off);
  const uint4* bv = reinterpret_cast<const uint4*>(b + vb);
  const long long stride = (long long)gridDim.x * kThreads * kUnroll;
  for (long long base = (long long)blockIdx.x * kThreads * kUnroll; base < nvec; base += stride) {
    if (ANY) {
      if (cnt) { any_raise(scratch); break; }
      if (any_raised(scratch)) break;
    }
    uint4 xa[kUnroll], xb[kUnroll];
    bool ok[kUnroll];
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      const long long v = base + u * kThreads + threadIdx.x;
      ok[u] = v < nvec;
      if (ok[u]) {
        xa[u] = MARK ? __ldcg(av + v) : ld_stream_u4(av + v);
        xb[u] = ld_stream_u4(bv + v);
      }
    }
#pragma unroll
    for (int u = 0; u < kUnroll; ++u) {
      if (!ok[u]) continue;
      const long long v = base + u * kThreads + threadIdx.x;
      uint32_t wa[4] = {xa[u].x, xa[u].y, xa[u].z, xa[u].w};
      const uint32_t wb[4] = {xb[u].x, xb[u].y, xb[u].z, xb[u].w};
      bool changed = false;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const uint32_t hit = __vcmpges4(wa[k], t4a) & __vcmpges4(wb[k], t4b);  // 0xFF per colliding byte
        cnt += __popc(hit) >> 3;
        if (MARK) {
          const uint32_t nw = (wa[k] & ~hit) | (0x7F7F7F7Fu & hit);
          changed |= nw != wa[k];
          wa[k] = nw;
        }
      }
      if (MARK && changed)
        *reinterpret_cast<uint4*>(a + vb + off + (v << 4)) = make_uint4(wa[0], wa[1], wa[2], wa[3]);
    }
  }
  if (ANY && cnt) any_raise(scratch);
  finish_block<false>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

// ---- summary kernel: occupancy of a BitVectorVoxel side comes from its 1-bit-per-voxel summary (Map::occ) instead
// of its 32-byte voxels; a ProbabilisticVoxel side is still streamed (1 B per voxel, 32 voxels per two 128-bit loads).
// One thread handles a group of 32 consecutive voxels per step: mask_a & mask_b, popcount. The voxels themselves are
// only touched where a collision needs a write-back (MARK) or its meaning bits (TYPES) -- sparse.
// Traffic per voxel: Prob x Bit 1.125 B (vs 33), Bit x Bit 0.25 B (vs 64). Requires off == 0.
__device__ __forceinline__ uint32_t prob_group_mask(const uint4& lo, const uint4& hi, uint32_t t4) {
  const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
  uint32_t m = 0;
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const uint32_t c = __vcmpges4(w[k], t4) & 0x01010101u;  // 1 per byte with occupancy >= threshold
    m |= (((c * 0x01020408u) >> 24) & 0xFu) << (4 * k);    // gather the 4 flags into a nibble
  }
  return m;
}

// Candidate bits of one 32-voxel group -> exact collisions: a ProbabilisticVoxel side that was represented by its
// "known" summary gets its threshold applied to the group's 32 bytes; then count, and touch the voxels themselves
// only for the write-back (MARK) or the meaning bits (TYPES) of colliding voxels.
template <int AT, int BT, bool MARK, bool TYPES, bool ASUM, bool BSUM>
__device__ __forceinline__ void resolve_group(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, long long g,
                                              uint32_t hit, uint32_t t4a, uint32_t t4b, uint32_t& cnt, uint32_t (&acc)[8]) {
  if (AT == kMapProb && ASUM) {
    const uint4* p = reinterpret_cast<const uint4*>(a + (g << 5));
    const uint4 lo = __ldcg(p), hi = __ldcg(p + 1);
    hit &= prob_group_mask(lo, hi, t4a);
  }
  if (BT == kMapProb && BSUM && hit) {
    const uint4* p = reinterpret_cast<const uint4*>(b + (g << 5));
    const uint4 lo = __ldg(p), hi = __ldg(p + 1);
    hit &= prob_group_mask(lo, hi, t4b);
  }
  cnt += __popc(hit);
  if (MARK || TYPES) {
    while (hit) {  // colliding voxels only
      const int j = __ffs(hit) - 1;
      hit &= hit - 1;
      const long long i = (g << 5) + j;
      if (AT == kMapBit) {
        uint32_t* vp = reinterpret_cast<uint32_t*>(a + (size_t)i * kBitVoxelBytes);
        if (TYPES) {
          const BitVoxelWords v = ld_voxel(vp);
#pragma unroll
          for (int k = 0; k < 8; ++k) acc[k] |= v.w[k];
          if (MARK && !(v.w[0] & 4u)) vp[0] = v.w[0] | 4u;
        } else if (MARK) {
          const uint32_t w0 = __ldcg(vp);
          if (!(w0 & 4u)) vp[0] = w0 | 4u;  // insert(eBVM_COLLISION); occupancy (and the summary) unchanged
        }
      } else if (MARK) {
        if (((const volatile int8_t*)a)[i] != 127) ((int8_t*)a)[i] = 127;  // stays "known"
      }
    }
  }
}

// voxel-wise handling of the last (n & 31) voxels by the first warp of CTA 0
template <int AT, int BT, bool MARK, bool TYPES>
__device__ __forceinline__ void summary_tail(uint8_t* __restrict__ a, const uint8_t* __restrict__ b,
                                             const uint32_t* __restrict__ occ_a, const uint32_t* __restrict__ occ_b,
                                             long long n, int thr, uint32_t& cnt, uint32_t (&acc)[8]) {
  if (blockIdx.x != 0 || threadIdx.x >= 32) return;
  const long long i = ((n >> 5) << 5) + threadIdx.x;
  if (i >= n) return;
  bool oa, ob;
  int pa = 0;
  if (AT == kMapBit) oa = (__ldg(occ_a + (i >> 5)) >> (i & 31)) & 1u;
  else { pa = ((const volatile int8_t*)a)[i]; oa = pa >= thr_of_a(thr); }
  if (BT == kMapBit) ob = (__ldg(occ_b + (i >> 5)) >> (i & 31)) & 1u;
  else ob = (int)ld_stream_i8(b + i) >= thr_of_b(thr);
  if (!(oa && ob)) return;
  ++cnt;
  if (AT == kMapBit && (MARK || TYPES)) {
    const BitVoxelWords va = ld_voxel(a + (size_t)i * kBitVoxelBytes);
    if (TYPES) {
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] |= va.w[k];
    }
    if (MARK && !(va.w[0] & 4u)) *(uint32_t*)(a + (size_t)i * kBitVoxelBytes) = va.w[0] | 4u;
  } else if (MARK && AT != kMapBit) {
    if (pa != 127) ((int8_t*)a)[i] = 127;
  }
}

// One queued candidate group: both sides' Prob bytes are requested together (no load depends on the other side's
// threshold mask), then the colliding voxels are counted / marked like in resolve_group.
template <int AT, int BT, bool MARK, bool TYPES>
__device__ __forceinline__ void resolve_entry(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, long long g,
                                              uint32_t hit, uint32_t t4a, uint32_t t4b, uint32_t& cnt, uint32_t (&acc)[8]) {
  uint4 alo, ahi, blo, bhi;
  if (AT == kMapProb) {
    const uint4* p = reinterpret_cast<const uint4*>(a + (g << 5));
    alo = __ldcg(p); ahi = __ldcg(p + 1);
  }
  if (BT == kMapProb) {
    const uint4* p = reinterpret_cast<const uint4*>(b + (g << 5));
    blo = __ldg(p); bhi = __ldg(p + 1);
  }
  if (AT == kMapProb) hit &= prob_group_mask(alo, ahi, t4a);
  if (BT == kMapProb) hit &= prob_group_mask(blo, bhi, t4b);
  if (hit) resolve_group<AT, BT, MARK, TYPES, false, false>(a, b, g, hit, t4a, t4b, cnt, acc);  // count + write-back only
}

// Both sides represented by summaries: 128-bit loads, 128 voxels per load pair and thread. Candidate groups (both
// summaries have a common bit) are NOT resolved where they are found -- that would put two dependent ~1 us voxel
// loads into the streaming loop of one lane while its 31 neighbours wait. Each warp appends them to a small queue in
// shared memory and resolves 32 at a time, one per lane, all loads in flight together.
constexpr int kQueuePerWarp = 64;
template <int AT, int BT, bool MARK, bool TYPES, bool ANY = false>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
k_collide_summary2(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, const uint32_t* __restrict__ occ_a,
                   const uint32_t* __restrict__ occ_b, long long n, int thr, ReduceScratch* scratch, ResultBlock* out_d,
                   ResultBlock* out_h, unsigned long long* count_out_d, PeerReduce pr,
                   const uint32_t* __restrict__ coarse_a, const uint32_t* __restrict__ coarse_b) {
  constexpr int U = 4;
  pdl_wait();
  pdl_launch_dependents();  // at most kCtasPerSm CTAs per SM, all resident: the next kernel's CTAs queue up behind
  __shared__ long long sh_g[kThreads / 32][kQueuePerWarp];
  __shared__ uint32_t sh_h[kThreads / 32][kQueuePerWarp];
  long long* qg = sh_g[threadIdx.x >> 5];
  uint32_t* qh = sh_h[threadIdx.x >> 5];
  const unsigned lane = threadIdx.x & 31;
  const unsigned lt = (1u << lane) - 1u;
  int qn = 0;  // warp-uniform number of queued groups
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const uint32_t t4a = thr4(thr_of_a(thr)), t4b = thr4(thr_of_b(thr));
  const long long n_full = n >> 5;     // complete 32-voxel groups
  const long long n_quad = n_full >> 2;  // complete 4-group quads (summaries are 16-byte padded and aligned)
  const uint4* qa = reinterpret_cast<const uint4*>(occ_a);
  const uint4* qb = reinterpret_cast<const uint4*>(occ_b);
  auto drain = [&](int m) {  // the last m (<= 32) queued groups, one per lane
    __syncwarp();
    if ((int)lane < m) resolve_entry<AT, BT, MARK, TYPES>(a, b, qg[qn - m + lane], qh[qn - m + lane], t4a, t4b, cnt, acc);
    qn -= m;
    __syncwarp();
  };
  const long long stride = (long long)gridDim.x * kThreads * U;
  for (long long base = (long long)blockIdx.x * kThreads * U; base < n_quad; base += stride) {  // CTA-uniform trip count
    if (ANY) {  // warp-uniform exit: the loop body holds warp collectives
      if (__any_sync(0xffffffffu, cnt != 0u)) { if (lane == 0) any_raise(scratch); break; }
      if (__any_sync(0xffffffffu, any_raised(scratch))) break;
    }
    uint4 ma[U], mb[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = base + u * kThreads + threadIdx.x;
      ok[u] = q < n_quad;
      // coarse masks (1 bit per 32 quads = per warp and step here, so the test is warp-uniform): a chunk that is empty in
      // either map cannot hold a collision -- neither summary is read for it
      if (ok[u] && coarse_a) { const long long cw = q >> 5; ok[u] = (__ldg(coarse_a + (cw >> 5)) >> (cw & 31)) & 1u; }
      if (ok[u] && coarse_b) { const long long cw = q >> 5; ok[u] = (__ldg(coarse_b + (cw >> 5)) >> (cw & 31)) & 1u; }
      if (ok[u]) { ma[u] = ld_stream_u4(qa + q); mb[u] = ld_stream_u4(qb + q); }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long q = base + u * kThreads + threadIdx.x;
      uint32_t h[4] = {0u, 0u, 0u, 0u};
      if (ok[u]) { h[0] = ma[u].x & mb[u].x; h[1] = ma[u].y & mb[u].y; h[2] = ma[u].z & mb[u].z; h[3] = ma[u].w & mb[u].w; }
      if (!__any_sync(0xffffffffu, (h[0] | h[1] | h[2] | h[3]) != 0u)) continue;
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const bool has = h[k] != 0u;
        const unsigned bal = __ballot_sync(0xffffffffu, has);
        if (!bal) continue;
        if (qn > kQueuePerWarp - 32) drain(32);
        if (has) {
          const int pos = qn + __popc(bal & lt);
          qg[pos] = (q << 2) + k;
          qh[pos] = h[k];
        }
        qn += __popc(bal);
      }
    }
    while (qn >= 32) drain(32);
    if (ANY) while (qn > 0) drain(qn < 32 ? qn : 32);  // find out now whether a candidate is a collision
  }
  while (qn > 0) drain(qn < 32 ? qn : 32);
  // up to 3 complete groups after the last quad
  if (blockIdx.x == 0 && threadIdx.x < (n_full & 3)) {
    const long long g = (n_quad << 2) + threadIdx.x;
    const uint32_t hit = __ldg(occ_a + g) & __ldg(occ_b + g);
    if (hit) resolve_group<AT, BT, MARK, TYPES, true, true>(a, b, g, hit, t4a, t4b, cnt, acc);
  }
  summary_tail<AT, BT, MARK, TYPES>(a, b, occ_a, occ_b, n, thr, cnt, acc);
  if (ANY && cnt) any_raise(scratch);
  finish_block<TYPES>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

// ASUM / BSUM: that side's summary is available. A BitVectorVoxel side always needs it (bit = occupied, exact); for a
// ProbabilisticVoxel side the summary bit means "known" (occupancy != UNKNOWN), a superset of "occupancy >= threshold":
// groups with candidate bits fetch their 32 Prob bytes and apply the threshold, groups without candidates (almost
// all of a sparse map) cost 1 bit per voxel. Without a summary the Prob side's bytes are streamed for every group.
template <int AT, int BT, bool MARK, bool TYPES, bool ASUM, bool BSUM, bool ANY = false>
__global__ void __launch_bounds__(kThreads, kCtasPerSm)
k_collide_summary(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, const uint32_t* __restrict__ occ_a,
                  const uint32_t* __restrict__ occ_b, long long n, int thr, ReduceScratch* scratch, ResultBlock* out_d,
                  ResultBlock* out_h, unsigned long long* count_out_d, PeerReduce pr) {
  static_assert(AT == kMapProb || ASUM, "a BitVectorVoxel side needs its occupancy summary");
  static_assert(BT == kMapProb || BSUM, "a BitVectorVoxel side needs its occupancy summary");
  constexpr int U = 4;
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const uint32_t t4a = thr4(thr_of_a(thr)), t4b = thr4(thr_of_b(thr));
  const long long n_full = n >> 5;  // complete 32-voxel groups
  const long long stride = (long long)gridDim.x * kThreads * U;
  for (long long base = (long long)blockIdx.x * kThreads * U; base < n_full; base += stride) {
    if (ANY) {
      if (cnt) { any_raise(scratch); break; }
      if (any_raised(scratch)) break;
    }
    uint4 alo[U], ahi[U], blo[U], bhi[U];
    uint32_t ma[U], mb[U];
    bool ok[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long g = base + u * kThreads + threadIdx.x;
      ok[u] = g < n_full;
      if (ok[u]) {
        if (ASUM) ma[u] = __ldg(occ_a + g);
        else {
          const uint4* p = reinterpret_cast<const uint4*>(a + (g << 5));
          if (MARK) { alo[u] = __ldcg(p); ahi[u] = __ldcg(p + 1); } else { alo[u] = ld_stream_u4(p); ahi[u] = ld_stream_u4(p + 1); }
        }
        if (BSUM) mb[u] = __ldg(occ_b + g);
        else {
          const uint4* p = reinterpret_cast<const uint4*>(b + (g << 5));
          blo[u] = ld_stream_u4(p); bhi[u] = ld_stream_u4(p + 1);
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!ok[u]) continue;
      const long long g = base + u * kThreads + threadIdx.x;
      if (!ASUM) ma[u] = prob_group_mask(alo[u], ahi[u], t4a);
      if (!BSUM) mb[u] = prob_group_mask(blo[u], bhi[u], t4b);
      const uint32_t hit = ma[u] & mb[u];
      if (hit) resolve_group<AT, BT, MARK, TYPES, ASUM, BSUM>(a, b, g, hit, t4a, t4b, cnt, acc);
    }
  }
  summary_tail<AT, BT, MARK, TYPES>(a, b, occ_a, occ_b, n, thr, cnt, acc);
  if (ANY && cnt) any_raise(scratch);
  finish_block<TYPES>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

// one warp per chunk of 32 summary quads (4096 voxels): any bit set -> the chunk's coarse bit
__global__ void __launch_bounds__(kThreads)
k_coarse_build(const uint4* __restrict__ occ, long long n_quads, uint32_t* __restrict__ coarse) {
  const long long n_chunks = (n_quads + 31) >> 5;
  const int lane = threadIdx.x & 31;
  for (long long c = ((long long)blockIdx.x * kThreads + threadIdx.x) >> 5; c < n_chunks;
       c += ((long long)gridDim.x * kThreads) >> 5) {
    const long long q = (c << 5) + lane;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (q < n_quads) v = ld_stream_u4(occ + q);
    const bool any = __any_sync(0xffffffffu, (v.x | v.y | v.z | v.w) != 0u);
    if (any && lane == 0) atomicOr(coarse + (c >> 5), 1u << (c & 31));
  }
}

cudaError_t launch_coarse_build(const LaunchCtx& c, const uint32_t* occ, long long n_quads, uint32_t* coarse) {
  const long long n_chunks = (n_quads + 31) >> 5;
  long long grid = (n_chunks * 32 + kThreads - 1) / kThreads;
  const long long cap = (long long)c.sm_count * 8;
  if (grid > cap) grid = cap;
  if (grid < 1) grid = 1;
  ++*c.launches;
  k_coarse_build<<<(int)grid, kThreads, 0, c.stream>>>(reinterpret_cast<const uint4*>(occ), n_quads, coarse);
  return cudaGetLastError();
}

template <int AT, int BT, bool ASUM, bool BSUM>
static cudaError_t launch_summary(const LaunchCtx& c, int grid, uint8_t* a, const uint8_t* b, const uint32_t* occ_a,
                                  const uint32_t* occ_b, long long n, int thr, bool mark, bool types,
                                  unsigned long long* out, const uint32_t* coarse_a, const uint32_t* coarse_b) {
  cudaError_t err = cudaSuccess;
  if (c.any) {  // boolean collisionCheck: no write-back, no meaning mask, early exit
    if (ASUM && BSUM)
      err = pdl_launch(k_collide_summary2<AT, BT, false, false, true>, grid, kThreads, c.stream, c.pdl, a, b, occ_a, occ_b, n,
                       thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer, coarse_a, coarse_b);
    else
      k_collide_summary<AT, BT, false, false, ASUM, BSUM, true><<<grid, kThreads, 0, c.stream>>>(
          a, b, occ_a, occ_b, n, thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);
    return err != cudaSuccess ? err : cudaGetLastError();
  }
#define GVX_GO(M, T)                                                                                       \
  do {                                                                                                     \
    if (ASUM && BSUM)                                                                                      \
      err = pdl_launch(k_collide_summary2<AT, BT, M, T>, grid, kThreads, c.stream, c.pdl, a, b, occ_a, occ_b, n, thr, \
                       c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer, coarse_a, coarse_b);         \
    else                                                                                                   \
      k_collide_summary<AT, BT, M, T, ASUM, BSUM><<<grid, kThreads, 0, c.stream>>>(                        \
          a, b, occ_a, occ_b, n, thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);              \
  } while (0)
  if (types && AT == kMapBit) { if (mark) GVX_GO(true, (AT == kMapBit)); else GVX_GO(false, (AT == kMapBit)); }
  else                        { if (mark) GVX_GO(true, false); else GVX_GO(false, false); }
#undef GVX_GO
  return err != cudaSuccess ? err : cudaGetLastError();
}

// CTAs for `work` CTA-iterations when at most max_grid CTAs are resident: every CTA gets the same number of
// iterations (no half-empty last wave)
static int balanced_grid(long long work, int max_grid) {
  if (work < 1) return 1;
  if (work <= max_grid) return (int)work;
  const long long iters = (work + max_grid - 1) / max_grid;
  return (int)((work + iters - 1) / iters);
}

template <int AT, int BT>
static cudaError_t launch_voxelwise(const LaunchCtx& c, int grid, uint8_t* a, const uint8_t* b, long long i0,
                                    long long i1, long long off, int thr, bool mark, bool types,
                                    unsigned long long* out) {
  if (c.any) {
    k_collide_voxelwise<AT, BT, false, false, true><<<grid, kThreads, 0, c.stream>>>(
        a, b, i0, i1, off, thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);
    return cudaGetLastError();
  }
#define GVX_GO(M, T) k_collide_voxelwise<AT, BT, M, T><<<grid, kThreads, 0, c.stream>>>( \
      a, b, i0, i1, off, thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer)
  if (types) { if (mark) GVX_GO(true, true); else GVX_GO(false, true); }
  else       { if (mark) GVX_GO(true, false); else GVX_GO(false, false); }
#undef GVX_GO
  return cudaGetLastError();
}

cudaError_t launch_collide_map_map(const LaunchCtx& c, void* a_, int a_type, const void* b_, int b_type,
                                   int64_t i_begin, int64_t i_end, int64_t off, int8_t thr, bool mark, bool types,
                                   unsigned long long* count_out_d, const uint32_t* occ_a, const uint32_t* occ_b,
                                   const uint32_t* coarse_a, const uint32_t* coarse_b, int thr_b_or_same) {
  const int8_t thr_a = thr, thr_b = thr_b_or_same == kSameThreshold ? thr : (int8_t)thr_b_or_same;
  const int thr_pack = (int)(uint8_t)thr_a | ((int)(uint8_t)thr_b << 8);  // decoded per side in the kernels
  uint8_t* a = (uint8_t*)a_;
  const uint8_t* b = (const uint8_t*)b_;
  if (i_end < i_begin) i_end = i_begin;
  const long long n = i_end - i_begin;
  const int max_grid = c.sm_count * kCtasPerSm < kMaxGrid ? c.sm_count * kCtasPerSm : kMaxGrid;
  ++*c.launches;
  // summary path (no offset, full range): every BitVectorVoxel side has a valid occupancy summary; a ProbabilisticVoxel
  // side uses its "known" summary when it has one and is streamed byte-wise otherwise (Prob x Prob needs both).
  const bool a_bit = a_type == kMapBit, b_bit = b_type == kMapBit;
  // A ProbabilisticVoxel summary marks "known" voxels, a superset of "occupancy >= thr" only for thr >= -127. A float
  // threshold that converts to -128 (DefaultCollider::floatToProbability truncates: thr in (-0.0079, -0.0039], or a
  // wrapped value) makes UNKNOWN voxels collide in the reference (GV/voxel/DefaultCollider.hpp:51-80): stream those.
  if (thr_a == -128 && !a_bit) occ_a = coarse_a = nullptr;
  if (thr_b == -128 && !b_bit) occ_b = coarse_b = nullptr;
  const bool a_sum = occ_a != nullptr, b_sum = occ_b != nullptr;
  if (off == 0 && i_begin == 0 && (!a_bit || a_sum) && (!b_bit || b_sum) && (a_bit || b_bit || (a_sum && b_sum)) &&
      (((uintptr_t)a | (uintptr_t)b) & 15) == 0) {
    const long long units = (a_sum && b_sum) ? (n >> 7) : (n >> 5);  // quads of groups when both sides are summaries
    const int grid = balanced_grid((units + kThreads * 4 - 1) / (kThreads * 4), max_grid);
#define GVX_SUM(AT, BT, AS, BS) return launch_summary<AT, BT, AS, BS>(c, grid, a, b, occ_a, occ_b, n, thr_pack, mark, types, count_out_d, coarse_a, coarse_b)
    if (a_bit && b_bit) GVX_SUM(kMapBit, kMapBit, true, true);
    if (a_bit) { if (b_sum) GVX_SUM(kMapBit, kMapProb, true, true); GVX_SUM(kMapBit, kMapProb, true, false); }
    if (b_bit) { if (a_sum) GVX_SUM(kMapProb, kMapBit, true, true); GVX_SUM(kMapProb, kMapBit, false, true); }
    GVX_SUM(kMapProb, kMapProb, true, true);
#undef GVX_SUM
  }
  // TMA bulk tiles "where they beat plain loads" (north_star; benchmarks/tma_probe.py): measured at 512^3,
  // Bit x Prob gains 5 % (7.15 vs 6.79 TB/s, tiles of 1024 voxels x 3 stages), Prob x Bit and Bit x Bit do not
  // (7.08 vs 7.15, 7.20 vs 7.23 TB/s) -- so auto (-1) feeds only Bit x Prob through TMA; k > 0 forces shape k everywhere.
  if (!c.any && c.tma != 0 && collide_tma_applicable(a, a_type, b, b_type, i_begin, off, types) &&
      (c.tma > 0 || (a_bit && !b_bit && n >= (1ll << 22)))) {
    LaunchCtx ct = c;
    if (ct.tma < 0) ct.tma = 2;
    return launch_collide_tma(ct, a, a_type, b, b_type, n, thr_pack, mark, count_out_d);
  }
  if (a_type == kMapProb && b_type == kMapProb) {
    const bool aligned = ((off & 15) == 0) && (((uintptr_t)a & 15) == 0) && (((uintptr_t)b & 15) == 0);
    if (aligned) {
      const int grid = balanced_grid((n / 16 + kThreads * kUnroll - 1) / (kThreads * kUnroll), max_grid);
      if (c.any) k_collide_prob_prob<false, true><<<grid, kThreads, 0, c.stream>>>((int8_t*)a, (const int8_t*)b, i_begin, i_end, off, thr_pack, c.scratch_d, c.result_d, c.result_hmap_d, count_out_d, c.peer);
      else if (mark) k_collide_prob_prob<true><<<grid, kThreads, 0, c.stream>>>((int8_t*)a, (const int8_t*)b, i_begin, i_end, off, thr_pack, c.scratch_d, c.result_d, c.result_hmap_d, count_out_d, c.peer);
      else k_collide_prob_prob<false><<<grid, kThreads, 0, c.stream>>>((int8_t*)a, (const int8_t*)b, i_begin, i_end, off, thr_pack, c.scratch_d, c.result_d, c.result_hmap_d, count_out_d, c.peer);
      return cudaGetLastError();
    }
  }
  // Bit x Bit advances 2 voxels per thread and step, the other pairs 4
  const int per_step = kThreads * ((a_type == kMapBit && b_type == kMapBit) ? 2 : kUnroll);
  const int grid = balanced_grid((n + per_step - 1) / per_step, max_grid);
  if (a_type == kMapBit && b_type == kMapProb) return launch_voxelwise<kMapBit, kMapProb>(c, grid, a, b, i_begin, i_end, off, thr_pack, mark, types, count_out_d);
  if (a_type == kMapProb && b_type == kMapBit) return launch_voxelwise<kMapProb, kMapBit>(c, grid, a, b, i_begin, i_end, off, thr_pack, mark, false, count_out_d);
  if (a_type == kMapBit && b_type == kMapBit) return launch_voxelwise<kMapBit, kMapBit>(c, grid, a, b, i_begin, i_end, off, thr_pack, mark, false, count_out_d);
  return launch_voxelwise<kMapProb, kMapProb>(c, grid, a, b, i_begin, i_end, off, thr_pack, mark, false, count_out_d);
}

}  // namespace gvx
