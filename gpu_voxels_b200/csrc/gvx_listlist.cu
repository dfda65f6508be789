// gvx_listlist.cu -- BitVectorVoxelList x BitVectorVoxelList: the list operations built on "find the entries with
// equal voxel id in both lists" (GV/voxellist/BitVoxelList.hpp:408-508 findMatchingVoxels, TemplateVoxelList.hpp:
// 227-282 collideVoxellists). The reference materialises two temporary lists of matching entries with four thrust
// passes (binary_search, count, copy_if, twice) and reduces those; here every entry of `this` binary-searches the
// other list's sorted ids once (the upper levels of the search stay in L2) and is reduced in place.
//
//   mode 0  collideWith(list)                  count of ids present in both            BitVoxelList.hpp:86-99
//   mode 1  collideWithTypes(list)             + OR of (this | other) voxels           BitVoxelList.hpp:101-125
//   mode 2  collideWithBitcheck(list, margin)  matching pairs passing the bit test     BitVoxelList.hpp:267-296
//   mode 3  collideCountingPerMeaning(list)    per-bit histogram of this' voxels       BitVoxelList.hpp:299-344
//   mode 4  overlap stencil for subtract       keep[i] = id not in other               TemplateVoxelList.hpp:609-633
#include "gvx_internal.h"
#include "gvx_reduce.cuh"
#include "gvx_sv.cuh"

namespace gvx {

// index of `key` in the sorted, duplicate-free ids[0, n), or -1
__device__ __forceinline__ long long find_id(const uint32_t* __restrict__ ids, unsigned long long n, uint32_t key) {
  unsigned long long lo = 0, hi = n;
  while (lo < hi) {
    const unsigned long long mid = (lo + hi) >> 1;
    const uint32_t v = __ldg(ids + mid);
    if (v < key) lo = mid + 1; else hi = mid;
  }
  return (lo < n && __ldg(ids + lo) == key) ? (long long)lo : -1;
}

template <int MODE>
__global__ void __launch_bounds__(kThreads)
k_list_list(const uint32_t* __restrict__ a_ids, const uint8_t* __restrict__ a_vox, unsigned long long na,
            const uint32_t* __restrict__ b_ids, const uint8_t* __restrict__ b_vox, unsigned long long nb,
            uint32_t margin, uint32_t* __restrict__ keep, ReduceScratch* scratch, ResultBlock* out_d,
            ResultBlock* out_h, PeerReduce pr) {
  __shared__ uint32_t sh_hist[MODE == 3 ? 256 : 1];
  if (MODE == 3) {
    sh_hist[threadIdx.x] = 0u;  // kThreads == 256
    __syncthreads();
  }
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const unsigned long long stride = (unsigned long long)gridDim.x * kThreads;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kThreads + threadIdx.x; i < na; i += stride) {
    const long long j = find_id(b_ids, nb, __ldg(a_ids + i));
    if (MODE == 4) { keep[i] = j < 0 ? 1u : 0u; if (j >= 0) ++cnt; continue; }
    if (j < 0) continue;
    if (MODE == 0) { ++cnt; continue; }
    VoxelBytes va;
    const uint4* pa = reinterpret_cast<const uint4*>(a_vox + i * kBitVoxelBytes);
    va.q[0] = __ldg(pa); va.q[1] = __ldg(pa + 1);
    if (MODE == 3) {
#pragma unroll
      for (int k = 0; k < 8; ++k) {
        uint32_t w = va.w[k];
        cnt += __popc(w);
        while (w) {
          const int b = __ffs(w) - 1;
          w &= w - 1;
          atomicAdd(&sh_hist[k * 32 + b], 1u);
        }
      }
      continue;
    }
    VoxelBytes vb;
    const uint4* pb = reinterpret_cast<const uint4*>(b_vox + (unsigned long long)j * kBitVoxelBytes);
    vb.q[0] = __ldg(pb); vb.q[1] = __ldg(pb + 1);
    if (MODE == 1) {
#pragma unroll
      for (int k = 0; k < 8; ++k) acc[k] |= va.w[k] | vb.w[k];
      ++cnt;
    } else {  // MODE == 2
      bool hit;
      if (margin == 0) {  // BitvectorCollision, GV/voxellist/BitVoxelList.h:40-48: ALL bits, not only swept-volume ones
        uint32_t both = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) both |= va.w[k] & vb.w[k];
        hit = both != 0u;
      } else {  // BitvectorCollisionWithBitshift(margin, 0), BitVoxelList.h:56-73
        VoxelBytes out;
        out.q[0] = make_uint4(0, 0, 0, 0); out.q[1] = make_uint4(0, 0, 0, 0);
        hit = margin_check(va.b, vb.b, out.b, margin, 0u);
      }
      if (hit) ++cnt;
    }
  }
  if (MODE == 3) {
    __syncthreads();
    const uint32_t h = sh_hist[threadIdx.x];
    if (h) atomicAdd(&out_d->per_meaning[threadIdx.x], (unsigned long long)h);  // zeroed by the launcher
    __threadfence();
    __syncthreads();
  }
  const bool last = finish_block<MODE == 1>(cnt, acc, scratch, out_d, out_h, nullptr, pr);
  if (MODE == 3 && last) {
    __threadfence();  // acquire side of the arrival ticket: the other CTAs' histogram adds are in L2
    out_h->per_meaning[threadIdx.x] = __ldcg(&out_d->per_meaning[threadIdx.x]);
  }
}

static int ll_grid(const LaunchCtx& c, unsigned long long n) {
  unsigned long long g = (n + kThreads - 1) / kThreads;
  const unsigned long long cap = (unsigned long long)c.sm_count * 8;
  if (g > cap) g = cap;
  if (g > (unsigned long long)kMaxGrid) g = kMaxGrid;
  return g < 1 ? 1 : (int)g;
}

cudaError_t launch_list_list(const LaunchCtx& c, int mode, const uint32_t* a_ids, const void* a_vox, uint64_t na,
                             const uint32_t* b_ids, const void* b_vox, uint64_t nb, uint32_t margin, uint32_t* keep_d) {
  const int grid = ll_grid(c, na);
  if (mode == 3) {
    cudaError_t e = cudaMemsetAsync(c.result_d->per_meaning, 0, sizeof(c.result_d->per_meaning), c.stream);
    if (e != cudaSuccess) return e;
  }
  ++*c.launches;
#define GVX_LL(M) k_list_list<M><<<grid, kThreads, 0, c.stream>>>(a_ids, (const uint8_t*)a_vox, na, b_ids, \
      (const uint8_t*)b_vox, nb, margin, keep_d, c.scratch_d, c.result_d, c.result_hmap_d, c.peer)
  switch (mode) {
    case 0: GVX_LL(0); break;
    case 1: GVX_LL(1); break;
    case 2: GVX_LL(2); break;
    case 3: GVX_LL(3); break;
    default: GVX_LL(4); break;
  }
#undef GVX_LL
  return cudaGetLastError();
}

// entries with keep != 0 move to position pos[i] (exclusive prefix sum of keep) of the output arrays
__global__ void __launch_bounds__(kThreads)
k_list_compact(const uint32_t* __restrict__ ids, const uint32_t* __restrict__ coords, const uint8_t* __restrict__ vox,
               unsigned long long n, const uint32_t* __restrict__ keep, const uint32_t* __restrict__ pos,
               uint32_t* __restrict__ o_ids, uint32_t* __restrict__ o_coords, uint8_t* __restrict__ o_vox) {
  const unsigned long long stride = (unsigned long long)gridDim.x * kThreads;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    if (!keep[i]) continue;
    const unsigned long long o = pos[i];
    o_ids[o] = ids[i];
    o_coords[3 * o] = coords[3 * i]; o_coords[3 * o + 1] = coords[3 * i + 1]; o_coords[3 * o + 2] = coords[3 * i + 2];
    const uint4* s = reinterpret_cast<const uint4*>(vox + i * kBitVoxelBytes);
    uint4* d = reinterpret_cast<uint4*>(o_vox + o * kBitVoxelBytes);
    d[0] = s[0]; d[1] = s[1];
  }
}

cudaError_t launch_list_compact(const LaunchCtx& c, const uint32_t* ids, const uint32_t* coords, const void* vox,
                                uint64_t n, const uint32_t* keep_d, const uint32_t* pos_d, uint32_t* o_ids,
                                uint32_t* o_coords, void* o_vox) {
  ++*c.launches;
  k_list_compact<<<ll_grid(c, n), kThreads, 0, c.stream>>>(ids, coords, (const uint8_t*)vox, n, keep_d, pos_d, o_ids,
                                                          o_coords, (uint8_t*)o_vox);
  return cudaGetLastError();
}

// TemplateVoxelList::equals, GV/voxellist/TemplateVoxelList.hpp:806-825: number of entries that differ in id, coordinate
// or voxel (the caller has compared the sizes)
__global__ void __launch_bounds__(kThreads)
k_list_diff(const uint32_t* __restrict__ a_ids, const uint32_t* __restrict__ a_co, const uint8_t* __restrict__ a_vox,
            const uint32_t* __restrict__ b_ids, const uint32_t* __restrict__ b_co, const uint8_t* __restrict__ b_vox,
            unsigned long long n, ReduceScratch* scratch, ResultBlock* out_d, ResultBlock* out_h, PeerReduce pr) {
  uint32_t cnt = 0;
  uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const unsigned long long stride = (unsigned long long)gridDim.x * kThreads;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kThreads + threadIdx.x; i < n; i += stride) {
    bool same = a_ids[i] == b_ids[i] && a_co[3 * i] == b_co[3 * i] && a_co[3 * i + 1] == b_co[3 * i + 1] &&
                a_co[3 * i + 2] == b_co[3 * i + 2];
    const uint4* pa = reinterpret_cast<const uint4*>(a_vox + i * kBitVoxelBytes);
    const uint4* pb = reinterpret_cast<const uint4*>(b_vox + i * kBitVoxelBytes);
    const uint4 x0 = __ldg(pa), x1 = __ldg(pa + 1), y0 = __ldg(pb), y1 = __ldg(pb + 1);
    same = same && x0.x == y0.x && x0.y == y0.y && x0.z == y0.z && x0.w == y0.w && x1.x == y1.x && x1.y == y1.y &&
           x1.z == y1.z && x1.w == y1.w;
    if (!same) ++cnt;
  }
  finish_block<false>(cnt, acc, scratch, out_d, out_h, nullptr, pr);
}

cudaError_t launch_list_diff(const LaunchCtx& c, const uint32_t* a_ids, const uint32_t* a_co, const void* a_vox,
                             const uint32_t* b_ids, const uint32_t* b_co, const void* b_vox, uint64_t n) {
  ++*c.launches;
  k_list_diff<<<ll_grid(c, n), kThreads, 0, c.stream>>>(a_ids, a_co, (const uint8_t*)a_vox, b_ids, b_co,
                                                       (const uint8_t*)b_vox, n, c.scratch_d, c.result_d,
                                                       c.result_hmap_d, c.peer);
  return cudaGetLastError();
}

}  // namespace gvx
