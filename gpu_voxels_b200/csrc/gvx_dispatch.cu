// gvx_dispatch.cu -- insertSensorData (log-odds hits) on a z-slab map shared by several GPUs, one sensor frame per rank:
// the all-to-all that routes every hit to the rank owning its voxel is part of the kernel, over NVLink peer memory.
//
// The reference integrates a frame with kernelTransformSensorData + kernelInsertSensorData into ONE map
// (GV/voxelmap/ProbVoxelMap.hpp:52-102, GV/voxelmap/kernels/VoxelMapOperations.cu:27-71, VoxelMapOperations.hpp:559-613).
// With the map cut into z-slabs (one per GPU) and one sensor per rank, gvx_update_logodds_multi lets every rank read
// ALL frames and keep its slab's points: work per rank grows with the number of ranks, and the frames have to be
// all-gathered first. Here every rank transforms and voxelises only ITS frame:
//   phase 1  hits inside the own slab are applied at once; a hit in rank d's slab is appended, as a slab-local
//            uint32 voxel index, to this rank's inbox in d's memory (plain stores through the IPC mapping; slots from
//            a local counter, one atomic per warp and destination). The last CTA publishes the entry counts to the
//            destinations (one packed word each, after a system-scope fence).
//   phase 2  every CTA waits for the counts of this epoch from all sources and applies its share of the inboxes.
//   credits  the NEXT dispatch kernel of a rank starts by telling every source that its inbox of the previous epoch
//            has been consumed; a source checks that credit before it overwrites an inbox (two epochs later).
// updateOccupancy(+delta) with one delta commutes under the saturating clamp (GV/voxel/ProbabilisticVoxel.hpp:51-57),
// so the order in which sources arrive does not matter: the slab ends up byte-identical to the slice of the map the
// reference builds from all frames. Waits are bounded (~4 s of SM clocks); a rank that never arrives sets `error`.
//
// Every CTA of the grid waits (phase 2) for the peers' counts, and a peer publishes its counts only when ALL its CTAs
// have finished phase 1: with CTAs still queued behind the waiting ones two ranks would wait for each other forever.
// The grid is therefore never larger than what is co-resident (occupancy query) and is launched cooperatively.
#include "gvx_internal.h"
#include "gvx_sensor.cuh"

namespace gvx {

__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

constexpr long long kDispatchSpinClocks = 8000000000ll;
constexpr int kDispatchStash = 2;  // chunks of a thread's points whose local hits wait until the counts are published

// one hit per entry: fold the warp's equal voxels into one saturating update (exact for same-sign deltas)
__device__ __forceinline__ void apply_hit(const MapView& m, bool live, uint64_t idx, int delta) {
  const unsigned long long key = live ? idx : (~0ull - (threadIdx.x & 31));
  const unsigned peers = __match_any_sync(0xffffffffu, key);
  if (live && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
    const bool first = logodds_add((int8_t*)m.data, idx, delta * __popc(peers));
    if (m.occ && first) summary_set(m.occ, idx);  // a voxel that was known has its bit already
  }
}

// sensor point i -> (local, idx) or (dest, idx): the arithmetic of k_update_logodds
__device__ __forceinline__ void classify(const MapView& m, const PeerDispatch& pd, const float* __restrict__ xyz,
                                         const float* P, bool sensor_inside, uint64_t layer, unsigned long long i,
                                         unsigned long long n, bool& local, int& dest, uint32_t& idx) {
  local = false; dest = -1; idx = 0;
  if (i >= n) return;
  const float rx = __ldg(xyz + 3 * i), ry = __ldg(xyz + 3 * i + 1), rz = __ldg(xyz + 3 * i + 2);
  float wx, wy, wz;
  sensor_to_world(P, rx, ry, rz, wx, wy, wz);
  if (isnan(wx) || isnan(wy) || isnan(wz)) return;
  const uint32_t cx = coord_to_voxel(wx, m.side), cy = coord_to_voxel(wy, m.side), cz = coord_to_voxel(wz, m.side);
  if (!(cx < m.dx && cy < m.dy && cz < m.dz && sensor_inside)) return;
  if (cz >= m.z0 && cz < m.z1) {
    local = true;
    idx = (uint32_t)((uint64_t)(cz - m.z0) * layer + (uint64_t)cy * m.dx + cx);  // every slab < 2^32 voxels (host check)
  } else if (pd.world > 1) {
    int d = 0;
    while (d + 1 < pd.world && cz >= pd.z_bounds[d + 1]) ++d;
    dest = d;
    idx = (uint32_t)((uint64_t)(cz - pd.z_bounds[d]) * layer + (uint64_t)cy * m.dx + cx);
  }
}

// append the warp's remote hits to the destinations' inboxes: one slot range per (warp, destination)
__device__ __forceinline__ bool send_hits(const PeerDispatch& pd, DispatchHead* me, int par, int dest, uint32_t idx) {
  const int lane = threadIdx.x & 31;
  const unsigned grp = __match_any_sync(0xffffffffu, dest);
  if (dest < 0) return false;
  const int leader = __ffs(grp) - 1;
  uint32_t base = 0;
  if (lane == leader) base = atomicAdd(&me->counters[dest], (unsigned)__popc(grp));
  base = __shfl_sync(grp, base, leader);
  const uint32_t slot = base + (uint32_t)__popc(grp & ((1u << lane) - 1u));
  if (slot < pd.capacity) {
    uint32_t* inbox = reinterpret_cast<uint32_t*>(pd.base[dest] + kDispatchHeadBytes) +
                      ((size_t)par * pd.world + pd.rank) * pd.capacity;
    inbox[slot] = idx;
  }
  return true;
}

__global__ void __launch_bounds__(kInsThreads)
k_update_logodds_dispatch(MapView m, const float* __restrict__ xyz, unsigned long long n, const Affine12 pose, int delta,
                          const PeerDispatch pd, uint32_t* __restrict__ error_h) {
  __shared__ uint32_t sh_cnt[kMaxPeers];
  __shared__ int sh_last, sh_bad;
  const float* P = pose.a;
  const int lane = threadIdx.x & 31;
  const int par = (int)(pd.epoch & 1ull);
  DispatchHead* me = reinterpret_cast<DispatchHead*>(pd.base[pd.rank]);
  const uint64_t layer = (uint64_t)m.dx * m.dy;
  if (threadIdx.x == 0) sh_bad = 0;
  __syncthreads();

  if (pd.world > 1 && threadIdx.x < 32) {
    // credit back: the previous dispatch kernel of this rank has finished (stream order), so the entries the peers sent
    // for epoch - 1 are consumed -- the first CTA tells every source that its inbox of that parity may be overwritten
    if (blockIdx.x == 0 && lane < pd.world && lane != pd.rank && pd.epoch > 1) {
      DispatchHead* src = reinterpret_cast<DispatchHead*>(pd.base[lane]);
      asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(&src->ack[pd.rank]), "l"(pd.epoch - 1ull) : "memory");
    }
    // credit check: destination d has consumed what this rank stored into inbox[par] two epochs ago
    bool ok = true;
    if (lane < pd.world && lane != pd.rank) {
      const long long t0 = clock64();
      while (ld_relaxed_sys(&me->ack[lane]) + 2ull < pd.epoch)
        if (clock64() - t0 > kDispatchSpinClocks) { ok = false; break; }
    }
    if (!__all_sync(0xffffffffu, ok) && lane == 0) sh_bad = 1;
  }

  // ---- phase 1: this rank's frame ----------------------------------------------------------------------------
  const uint32_t sx = coord_to_voxel(P[3], m.side), sy = coord_to_voxel(P[7], m.side), sz = coord_to_voxel(P[11], m.side);
  const bool sensor_inside = sx < m.dx && sy < m.dy && sz < m.dz;
  const unsigned long long nvox = layer * m.dz;
  if (n > nvox) n = nvox;  // the reference's loop bound, VoxelMapOperations.hpp:569
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  const unsigned long long first = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x;
  const unsigned long long chunks = (n_round + stride - 1) / stride;  // grid-uniform
  // the last kDispatchStash chunks: classify now, send the remote hits, keep the local ones in registers until the
  // counts are on their way -- the local atomics then overlap the NVLink latency of the exchange
  const unsigned long long stash0 = chunks > (unsigned long long)kDispatchStash ? chunks - kDispatchStash : 0;
  bool st_local[kDispatchStash];
  int st_dest[kDispatchStash];
  uint32_t st_idx[kDispatchStash];
#pragma unroll
  for (int u = 0; u < kDispatchStash; ++u)
    classify(m, pd, xyz, P, sensor_inside, layer, first + (stash0 + u) * stride, n, st_local[u], st_dest[u], st_idx[u]);
  __syncthreads();  // sh_bad is final: the credits are in (the wait above ran while the other warps classified)
  const bool can_send = !sh_bad;
  bool sent = false;
  for (unsigned long long c = 0; c < stash0; ++c) {  // long frames: the early chunks are handled on the spot
    bool local; int dest; uint32_t idx;
    classify(m, pd, xyz, P, sensor_inside, layer, first + c * stride, n, local, dest, idx);
    apply_hit(m, local, idx, delta);
    if (pd.world > 1) sent |= send_hits(pd, me, par, can_send ? dest : -1, idx);
  }
  if (pd.world > 1) {
#pragma unroll
    for (int u = 0; u < kDispatchStash; ++u) sent |= send_hits(pd, me, par, can_send ? st_dest[u] : -1, st_idx[u]);
    // all entries of this CTA are out; the last CTA of the grid publishes the counts
    const int any_sent = __syncthreads_or(sent);
    if (threadIdx.x == 0) {
      if (any_sent) __threadfence_system();
      const unsigned int t = atomicInc(&me->ticket_sent, gridDim.x - 1);
      sh_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (sh_last && threadIdx.x < 32) {
      __threadfence_system();
      if (lane < pd.world && lane != pd.rank) {
        uint32_t c = atomicExch(&me->counters[lane], 0u);
        if (c > pd.capacity) { c = pd.capacity; *error_h = 2u; }
        DispatchHead* dst = reinterpret_cast<DispatchHead*>(pd.base[lane]);
        st_release_sys(&dst->ready[par][pd.rank], ((pd.epoch & 0xFFFFull) << 48) | c);
      }
    }
  }
#pragma unroll
  for (int u = 0; u < kDispatchStash; ++u) apply_hit(m, st_local[u], st_idx[u], delta);
  if (pd.world <= 1) return;

  // ---- phase 2: what the other ranks sent ----------------------------------------------------------------------
  if (threadIdx.x < 32) {
    uint32_t c = 0;
    bool ok = true;
    if (lane < pd.world && lane != pd.rank) {
      const long long t0 = clock64();
      unsigned long long v;
      for (;;) {
        v = ld_acquire_sys(&me->ready[par][lane]);
        if ((v >> 48) == (pd.epoch & 0xFFFFull)) break;
        if (clock64() - t0 > kDispatchSpinClocks) { ok = false; break; }
      }
      c = ok ? (uint32_t)v : 0u;
      if (c > pd.capacity) c = pd.capacity;
    }
    if (lane < kMaxPeers) sh_cnt[lane] = c;
    if (!__all_sync(0xffffffffu, ok) && lane == 0) sh_bad = 1;
  }
  __syncthreads();
  if (sh_bad && threadIdx.x == 0) { *error_h = 1u; __threadfence_system(); }  // a rank did not arrive in time
  const uint64_t n_slab = layer * (m.z1 - m.z0);
  const uint32_t* inbox0 = reinterpret_cast<const uint32_t*>(pd.base[pd.rank] + kDispatchHeadBytes) +
                           (size_t)par * pd.world * pd.capacity;
  for (int s = 0; s < pd.world; ++s) {
    const uint32_t cnt = sh_cnt[s];
    if (cnt == 0) continue;
    const uint32_t* in = inbox0 + (size_t)s * pd.capacity;
    const uint32_t cnt_round = (cnt + 31u) & ~31u;
    for (unsigned long long j = first; j < cnt_round; j += stride) {
      uint64_t idx = 0;
      bool live = false;
      if (j < cnt) {
        idx = __ldcg(in + j);  // written over NVLink into this GPU's L2: never through L1
        live = idx < n_slab;
      }
      apply_hit(m, live, idx, delta);
    }
  }
}

cudaError_t launch_update_logodds_dispatch(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                           const Affine12& pose, int delta, const PeerDispatch& pd, uint32_t* error_hmap_d) {
  static int per_sm = 0;
  if (!per_sm) {
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, k_update_logodds_dispatch, kInsThreads, 0);
    if (e != cudaSuccess) { per_sm = 0; return e; }
    if (per_sm < 1) per_sm = 1;
  }
  unsigned long long g = (n + kInsThreads - 1) / kInsThreads;
  const unsigned long long cap = (unsigned long long)c.sm_count * per_sm;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  ++*c.launches;
  if (pd.world <= 1) {
    k_update_logodds_dispatch<<<(int)g, kInsThreads, 0, c.stream>>>(m, xyz_d, n, pose, delta, pd, error_hmap_d);
    return cudaGetLastError();
  }
  MapView mv = m;
  unsigned long long nn = n;
  Affine12 P = pose;
  PeerDispatch d = pd;
  void* args[] = {&mv, &xyz_d, &nn, &P, &delta, &d, &error_hmap_d};
  return cudaLaunchCooperativeKernel((const void*)k_update_logodds_dispatch, dim3((unsigned)g), dim3(kInsThreads), args, 0,
                                     c.stream);
}

}  // namespace gvx
