// gvx_list.cu -- BitVectorVoxelList for sm_100a: build (voxelise + sort + merge + unique) and list x map collisions.
//
// Replaces kernelInsertGlobalPointCloud / kernelInsertMetaPointCloud of the list
// (GV/voxellist/kernels/VoxelListOperations.hpp:41-120), TemplateVoxelList::make_unique
// (GV/voxellist/TemplateVoxelList.hpp:141-208: thrust::sort_by_key over a zip of 48-byte payload tuples, a reverse
// inclusive scan with the OR-Merge functor :41-65, unique_by_key) and kernelCollideWithVoxelMap /
// kernelCollideWithVoxelMapBitMask / the types kernel (VoxelListOperations.hpp:126-314) with their D2H + host-sum tails.
//
// Build: only (id, source index) pairs are sorted -- 8 bytes per entry instead of the reference's 48 -- with the
// toolkit's CUB radix sort (LSD radix sort is stable, like the reference's sort_by_key on integer keys); run heads are
// found by comparing neighbours, an inclusive scan of the head flags gives every entry its output slot, and a
// scatter kernel ORs each entry's meaning bits into its slot (one 32-bit atomicOr per new point).
// Collide: one thread per list entry; id load (coalesced), map voxel gather (one 32-B sector), and the 32-B list
// voxel is only fetched when the map voxel is occupied, which is the precondition of every collision predicate.
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include "gvx_internal.h"
#include "gvx_reduce.cuh"

namespace gvx {

// ---- build --------------------------------------------------------------------------------------------------------
// keys/vals for the sort: entries [0,s) are the existing (sorted, unique) list, entries [s,s+n) the new points.
// Points are NOT bounds-checked; the id is getVoxelIndexUnsigned in uint32 arithmetic and wraps for outside points
// (GV/voxelmap/kernels/VoxelMapOperations.h:54-58), exactly like the reference's list insert.
__global__ void __launch_bounds__(256)
k_list_keys(const uint32_t* __restrict__ old_ids, unsigned long long s, const float* __restrict__ xyz,
            unsigned long long n, uint32_t dx, uint32_t dy, float side, MetaDesc d, int single_meaning,
            uint32_t* __restrict__ keys, uint32_t* __restrict__ vals, uint8_t* __restrict__ new_meaning,
            const uint32_t* __restrict__ add_ids, uint32_t add_id_off) {
  const unsigned long long stride = (unsigned long long)gridDim.x * 256;
  for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < s + n; i += stride) {
    uint32_t id;
    if (i < s) {
      id = old_ids[i];
    } else if (add_ids) {  // merge: entry of the other list, id + signed linear offset in uint32 arithmetic
      id = add_ids[i - s] + add_id_off;
    } else {
      const unsigned long long p = i - s;
      const float x = __ldg(xyz + 3 * p), y = __ldg(xyz + 3 * p + 1), z = __ldg(xyz + 3 * p + 2);
      const uint32_t cx = coord_to_voxel(x, side), cy = coord_to_voxel(y, side), cz = coord_to_voxel(z, side);
      id = cz * dx * dy + cy * dx + cx;
      int mg = single_meaning;
      if (d.prefix_d) mg = d.meanings_d[d.n_meanings == 1 ? 0 : find_cloud(d.prefix_d, d.n_clouds, (uint32_t)p)];
      new_meaning[p] = (uint8_t)mg;
    }
    keys[i] = id;
    vals[i] = (uint32_t)i;
  }
}

__global__ void __launch_bounds__(256)
k_list_heads(const uint32_t* __restrict__ keys, unsigned long long n, uint32_t* __restrict__ flags) {
  const unsigned long long stride = (unsigned long long)gridDim.x * 256;
  for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < n; i += stride)
    flags[i] = (i == 0 || keys[i] != keys[i - 1]) ? 1u : 0u;
}

// Scatter every sorted entry into its output slot pos[i]-1. The run head (first entry of equal ids after the stable
// sort) provides id and coordinates, as thrust::unique_by_key keeps the first; every entry ORs its bits in.
__global__ void __launch_bounds__(256)
k_list_merge(const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals, const uint32_t* __restrict__ pos,
             unsigned long long total, unsigned long long s, const uint32_t* __restrict__ old_coords,
             const uint8_t* __restrict__ old_voxels, const float* __restrict__ xyz,
             const uint8_t* __restrict__ new_meaning, float side, uint32_t* __restrict__ out_ids,
             uint32_t* __restrict__ out_coords, uint32_t* __restrict__ out_voxels, ResultBlock* out_d,
             ResultBlock* out_h, const uint32_t* __restrict__ add_coords, const uint8_t* __restrict__ add_voxels,
             int3 add_coord_off, int add_meaning) {
  const unsigned long long stride = (unsigned long long)gridDim.x * 256;
  for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < total; i += stride) {
    const uint32_t key = keys[i], src = vals[i], slot = pos[i] - 1u;
    const bool head = (i == 0) || keys[i - 1] != key;
    uint32_t* ov = out_voxels + (size_t)slot * 8;
    if (src < s) {
      const BitVoxelWords v = ld_voxel_stream(old_voxels + (size_t)src * kBitVoxelBytes);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (v.w[k]) atomicOr(ov + k, v.w[k]);
      if (head) {
        out_ids[slot] = key;
        out_coords[3 * (size_t)slot] = old_coords[3 * (size_t)src];
        out_coords[3 * (size_t)slot + 1] = old_coords[3 * (size_t)src + 1];
        out_coords[3 * (size_t)slot + 2] = old_coords[3 * (size_t)src + 2];
      }
    } else if (add_coords) {  // an entry of the merged list
      const unsigned long long p = src - s;
      if (add_meaning >= 0) {
        atomicOr(ov + (add_meaning >> 5), 1u << (add_meaning & 31));
      } else {
        const BitVoxelWords v = ld_voxel_stream(add_voxels + (size_t)p * kBitVoxelBytes);
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (v.w[k]) atomicOr(ov + k, v.w[k]);
      }
      if (head) {
        out_ids[slot] = key;
        out_coords[3 * (size_t)slot] = add_coords[3 * p] + (uint32_t)add_coord_off.x;
        out_coords[3 * (size_t)slot + 1] = add_coords[3 * p + 1] + (uint32_t)add_coord_off.y;
        out_coords[3 * (size_t)slot + 2] = add_coords[3 * p + 2] + (uint32_t)add_coord_off.z;
      }
    } else {
      const unsigned long long p = src - s;
      const uint32_t bit = new_meaning[p];
      atomicOr(ov + (bit >> 5), 1u << (bit & 31));
      if (head) {
        out_ids[slot] = key;
        out_coords[3 * (size_t)slot] = coord_to_voxel(__ldg(xyz + 3 * p), side);
        out_coords[3 * (size_t)slot + 1] = coord_to_voxel(__ldg(xyz + 3 * p + 1), side);
        out_coords[3 * (size_t)slot + 2] = coord_to_voxel(__ldg(xyz + 3 * p + 2), side);
      }
    }
    if (i == total - 1) {  // number of unique entries -> result block (device + mapped host)
      out_d->count = slot + 1ull;
      out_h->count = slot + 1ull;
      __threadfence_system();
    }
  }
}

static int list_grid(const LaunchCtx& c, unsigned long long n) {
  unsigned long long g = (n + 255) / 256;
  const unsigned long long cap = (unsigned long long)c.sm_count * 8;
  if (g > cap) g = cap;
  return g < 1 ? 1 : (int)g;
}

size_t list_sort_temp_bytes(uint64_t n) {
  size_t a = 0, b = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, a, (const uint32_t*)nullptr, (uint32_t*)nullptr, (const uint32_t*)nullptr,
                                  (uint32_t*)nullptr, (long long)n);
  cub::DeviceScan::InclusiveSum(nullptr, b, (const uint32_t*)nullptr, (uint32_t*)nullptr, (long long)n);
  return a > b ? a : b;
}

// out[i] = sum of in[0, i) (CUB, library), tmp >= list_sort_temp_bytes(n)
cudaError_t launch_exclusive_sum(const LaunchCtx& c, const uint32_t* in_d, uint32_t* out_d, uint64_t n, void* tmp,
                                 size_t tmp_bytes) {
  ++*c.launches;
  return cub::DeviceScan::ExclusiveSum(tmp, tmp_bytes, in_d, out_d, (long long)n, c.stream);
}

cudaError_t launch_list_build(const LaunchCtx& c, const ListBuild& b) {
  const unsigned long long total = b.old_size + b.n_new;
  if (total == 0) return cudaSuccess;
  const int grid = list_grid(c, total);
  *c.launches += 3;
  k_list_keys<<<grid, 256, 0, c.stream>>>(b.old_ids, b.old_size, b.xyz_d, b.n_new, b.dx, b.dy, b.side, b.meta,
                                          b.single_meaning, b.keys_in, b.vals_in, b.new_meaning, b.add_ids, b.add_id_off);
  size_t tmp = b.tmp_bytes;
  cudaError_t e = cub::DeviceRadixSort::SortPairs(b.tmp, tmp, b.keys_in, b.keys_out, b.vals_in, b.vals_out,
                                                  (long long)total, 0, 32, c.stream);
  if (e != cudaSuccess) return e;
  k_list_heads<<<grid, 256, 0, c.stream>>>(b.keys_out, total, b.keys_in);  // keys_in is free again: head flags
  tmp = b.tmp_bytes;
  e = cub::DeviceScan::InclusiveSum(b.tmp, tmp, b.keys_in, b.vals_in, (long long)total, c.stream);  // vals_in: slots
  if (e != cudaSuccess) return e;
  e = cudaMemsetAsync(b.out_voxels, 0, (size_t)total * kBitVoxelBytes, c.stream);
  if (e != cudaSuccess) return e;
  k_list_merge<<<grid, 256, 0, c.stream>>>(b.keys_out, b.vals_out, b.vals_in, total, b.old_size, b.old_coords,
                                           b.old_voxels, b.xyz_d, b.new_meaning, b.side, b.out_ids, b.out_coords,
                                           (uint32_t*)b.out_voxels, c.result_d, c.result_hmap_d, b.add_coords, b.add_voxels,
                                           make_int3(b.add_coord_off[0], b.add_coord_off[1], b.add_coord_off[2]),
                                           b.add_meaning);
  return cudaGetLastError();
}

// ---- list x map collisions ------------------------------------------------------------------------------------------
// MODE 0  collideWith (DefaultCollider, GV/voxellist/kernels/VoxelListOperations.hpp:126-177)
// MODE 1  collideWithTypes: count + OR of the colliding list voxels (:182-240); map test is isOccupied(float)
// MODE 2  collideWithTypeMask: count entries with (mask & list voxel) != 0 (:260-314); map test is isOccupied(float)
// MODE 3  per-meaning counts over MODE 0's colliding set (semantics of BitVoxelList::collideCountingPerMeaning,
//         GV/voxellist/BitVoxelList.hpp:300-344): every set bit j of a colliding entry adds one to counts[j] and to
//         the total.
// thr_i: DefaultCollider's int8 threshold (modes 0/3); thr_f: the raw float the types/mask kernels compare a
// ProbabilisticVoxel against (GV/voxel/ProbabilisticVoxel.hpp:94-98).
template <int MAPT, int MODE, bool MARK>
__global__ void __launch_bounds__(kThreads, 4)
k_collide_list_map(const uint32_t* __restrict__ ids, uint8_t* __restrict__ voxels, unsigned long long size,
                   const uint8_t* __restrict__ map, long long n_total, long long map_first, long long map_n,
                   int thr_i, float thr_f, long long off, BitVoxelWords mask_in, const uint32_t* __restrict__ occ,
                   ReduceScratch* scratch,
                   ResultBlock* out_d, ResultBlock* out_h, unsigned long long* count_out_d, PeerReduce pr) {
  __shared__ uint32_t sh_hist[MODE == 3 ? 256 : 1];
  if (MODE == 3) {
    sh_hist[threadIdx.x] = 0u;  // kThreads == 256
    __syncthreads();
  }
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  // U entries per thread and step in three stages -- ids, then the map probe of every entry, then the (rare) hits --
  // so that a thread has U independent gathers in flight instead of one dependent id -> map -> list-voxel chain
  constexpr int U = 8;
  const unsigned long long stride = (unsigned long long)gridDim.x * kThreads * U;
  for (unsigned long long base = (unsigned long long)blockIdx.x * kThreads * U; base < size; base += stride) {
    long long v[U];  // index into this slab's part of the map, < 0: skip
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const unsigned long long i = base + (unsigned long long)u * kThreads + threadIdx.x;
      v[u] = -1;
      if (i < size) {
        const long long j = (long long)__ldg(ids + i) + off;
        // pointer range test of the reference (:139-140); then: another z-slab's voxel
        if (j >= 0 && j < n_total && j >= map_first && j < map_first + map_n) v[u] = j - map_first;
      }
    }
    uint32_t sw[U];
    if (occ != nullptr) {
#pragma unroll
      for (int u = 0; u < U; ++u) sw[u] = v[u] >= 0 ? __ldg(occ + (v[u] >> 5)) : 0u;
    }
    bool hit_map[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      hit_map[u] = false;
      if (v[u] < 0) continue;
      if (MAPT == kMapProb) {
        // "known" summary: an UNKNOWN voxel never collides, skip its byte
        if (occ != nullptr && !((sw[u] >> (v[u] & 31)) & 1u)) continue;
        const int p = (int)ld_stream_i8(map + v[u]);
        hit_map[u] = (MODE == 0 || MODE == 3) ? (p >= thr_i) : ((float)p >= thr_f);
      } else if (occ != nullptr) {  // occupancy summary of the map: one bit instead of a 32-byte voxel
        hit_map[u] = (sw[u] >> (v[u] & 31)) & 1u;
      } else {
        hit_map[u] = bit_occupied(ld_voxel_stream(map + (size_t)v[u] * kBitVoxelBytes));
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!hit_map[u]) continue;
      const unsigned long long i = base + (unsigned long long)u * kThreads + threadIdx.x;
      uint8_t* lp = voxels + (size_t)i * kBitVoxelBytes;
      const BitVoxelWords lv = ld_voxel(lp);
      bool hit;
      if (MODE == 0 || MODE == 3) {
        hit = bit_occupied(lv);
      } else if (MODE == 2) {
        uint32_t any = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) any |= lv.w[k] & mask_in.w[k];
        hit = any != 0u;
      } else {
        hit = true;
      }
      if (!hit) continue;
      if (MODE == 1) {
#pragma unroll
        for (int k = 0; k < 8; ++k) acc[k] |= lv.w[k];
      }
      if (MODE == 3) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          uint32_t w = lv.w[k];
          cnt += __popc(w);
          while (w) {
            const int b = __ffs(w) - 1;
            w &= w - 1;
            atomicAdd(&sh_hist[k * 32 + b], 1u);
          }
        }
      } else {
        ++cnt;
      }
      if (MARK && !(lv.w[0] & 4u)) *(uint32_t*)lp = lv.w[0] | 4u;  // list voxel .insert(eBVM_COLLISION)
    }
  }
  if (MODE == 3) {
    __syncthreads();
    const uint32_t h = sh_hist[threadIdx.x];
    if (h) atomicAdd(&scratch->pm_acc[threadIdx.x], (unsigned long long)h);  // zero at launch: the last CTA of the launch before reset it
    __threadfence();
    __syncthreads();
  }
  const bool last = finish_block<MODE == 1>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
  if (MODE == 3 && last) {
    __threadfence();  // acquire side of the arrival ticket: the other CTAs' histogram adds are in L2
    unsigned long long h = __ldcg(&scratch->pm_acc[threadIdx.x]);
    scratch->pm_acc[threadIdx.x] = 0ull;  // self-resetting, like the count accumulator
    if (pr.world > 1 && pr.vec_epoch != 0)  // sharded map: the 256 per-meaning counts are all-reduced here, over NVLink peer memory, like the
      h = peer_allreduce_word(h, threadIdx.x + 1, pr);  // count in finish_block (word 0 of the vector area stays unused)
    out_d->per_meaning[threadIdx.x] = h;
    // a device destination holds 257 words in this mode: the count (finish_block) and the per-meaning counts behind it --
    // no copy kernel follows the launch
    if (count_out_d) count_out_d[1 + threadIdx.x] = h;
    if (out_h) {
      out_h->per_meaning[threadIdx.x] = h;
      __threadfence_system();
    }
  }
}

// lower bounds of lo and hi in the sorted ids: lanes 0..15 search lo, lanes 16..31 search hi, 16-ary (each step narrows
// the range 17-fold with ONE round of loads instead of a chain of dependent probes). Results: out->count = i0,
// out->per_meaning[0] = i1 (device block and mapped host block).
__global__ void k_list_bounds(const uint32_t* __restrict__ ids, unsigned long long size, unsigned long long lo,
                              unsigned long long hi, ResultBlock* out_d, ResultBlock* out_h) {
  const int lane = threadIdx.x & 31, half = lane >> 4, sub = lane & 15;
  const unsigned long long key = half ? hi : lo;
  unsigned long long a = 0, b = size;  // answer in [a, b]
  while (a < b) {
    const unsigned long long len = b - a, step = (len + 16) / 17;  // probes at a + (sub + 1) * step - 1
    const unsigned long long pos = a + (unsigned long long)(sub + 1) * step - 1;
    const bool valid = pos < b;
    const bool less = valid && (unsigned long long)__ldg(ids + pos) < key;  // ids[pos] < key: the bound is right of pos
    const unsigned m = __ballot_sync(0xffffffffu, less) >> (16 * half) & 0xFFFFu;
    const int k = __popc(m);  // ids sorted: the probes below the bound are a prefix
    const unsigned long long na = k ? a + (unsigned long long)k * step : a;
    unsigned long long nb = k < 16 ? a + (unsigned long long)(k + 1) * step - 1 : b;  // nothing was probed beyond probe 15
    if (nb > b) nb = b;
    a = na; b = nb;
  }
  if (sub == 0) {
    if (half == 0) { out_d->count = a; if (out_h) out_h->count = a; }
    else { out_d->per_meaning[0] = a; if (out_h) out_h->per_meaning[0] = a; }
    __threadfence_system();
  }
}

cudaError_t launch_list_bounds(const LaunchCtx& c, const uint32_t* ids, uint64_t size, uint64_t lo, uint64_t hi) {
  ++*c.launches;
  k_list_bounds<<<1, 32, 0, c.stream>>>(ids, size, lo, hi, c.result_d, c.result_hmap_d);
  return cudaGetLastError();
}

template <int MAPT, int MODE>
static void go_list(const LaunchCtx& c, int grid, bool mark, const uint32_t* ids, uint8_t* voxels, uint64_t size,
                    const uint8_t* map, long long n_total, long long map_first, long long map_n, int thr_i, float thr_f,
                    long long off, const BitVoxelWords& mk, const uint32_t* occ, unsigned long long* out) {
  if (mark)
    k_collide_list_map<MAPT, MODE, true><<<grid, kThreads, 0, c.stream>>>(ids, voxels, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);
  else
    k_collide_list_map<MAPT, MODE, false><<<grid, kThreads, 0, c.stream>>>(ids, voxels, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);
}

cudaError_t launch_collide_list_map(const LaunchCtx& c, int mode, const uint32_t* ids, void* voxels, uint64_t size,
                                    const void* map_, int map_type, uint64_t n_total, uint64_t map_first,
                                    uint64_t map_n, int8_t thr_i, float thr_f, int64_t off, bool mark,
                                    const uint32_t mask_in8[8], unsigned long long* count_out_d, const uint32_t* occ) {
  BitVoxelWords mk;
  for (int k = 0; k < 8; ++k) mk.w[k] = mask_in8 ? mask_in8[k] : 0u;
  // "known" is a superset of "occupied at the threshold" only while UNKNOWN (-128) fails the test: DefaultCollider's
  // int8 threshold -128 (modes 0/3) or a raw float threshold <= -128 (modes 1/2, GV/voxel/ProbabilisticVoxel.hpp:94-98)
  // make UNKNOWN voxels collide in the reference -- read the bytes then
  if (map_type == kMapProb && (((mode == 0 || mode == 3) && thr_i == -128) || ((mode == 1 || mode == 2) && thr_f <= -128.0f)))
    occ = nullptr;
  unsigned long long work = (size + kThreads * 8 - 1) / (kThreads * 8);
  const unsigned long long max_grid = (unsigned long long)(c.sm_count * 4 < kMaxGrid ? c.sm_count * 4 : kMaxGrid);
  const int grid = (int)(work < 1 ? 1 : (work > max_grid ? max_grid : work));
  const uint8_t* map = (const uint8_t*)map_;
  uint8_t* vx = (uint8_t*)voxels;
  // mode 3: count_out_d, when given, receives 257 words (count, then 256 per-meaning counts); the histogram
  // accumulators live in ReduceScratch and reset themselves -- ONE launch, no memset before, no copy after
  ++*c.launches;
#define GVX_LM(MT)                                                                                                      \
  switch (mode) {                                                                                                       \
    case 0: go_list<MT, 0>(c, grid, mark, ids, vx, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, count_out_d); break; \
    case 1: go_list<MT, 1>(c, grid, mark, ids, vx, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, count_out_d); break; \
    case 2: go_list<MT, 2>(c, grid, mark, ids, vx, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, count_out_d); break; \
    default: go_list<MT, 3>(c, grid, mark, ids, vx, size, map, n_total, map_first, map_n, thr_i, thr_f, off, mk, occ, count_out_d); break; \
  }
  if (map_type == kMapProb) { GVX_LM(kMapProb) } else { GVX_LM(kMapBit) }
#undef GVX_LM
  return cudaGetLastError();
}

}  // namespace gvx
