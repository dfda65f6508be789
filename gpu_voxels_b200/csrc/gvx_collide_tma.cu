// gvx_collide_tma.cu -- collideWith map x map, voxel-streaming variant fed by TMA bulk tiles (sm_100a).
//
// Same result as k_collide_voxelwise (gvx_collide.cu; replaces kernelCollideVoxelMapsDebug,
// GV/voxelmap/kernels/VoxelMapOperations.hpp:129-184): instead of one LDG.E.256 per voxel and thread, one elected
// thread per CTA issues cp.async.bulk copies of whole tiles (TILE BitVectorVoxels = TILE x 32 B, TILE
// ProbabilisticVoxels = TILE B) into a ring of shared-memory stages, completion counted by an mbarrier per stage
// (complete_tx::bytes); all threads then read the tile from shared memory with conflict-free LDS.128.
// BASELINE.json north_star: "TMA bulk tiles where they beat plain loads" -- selected with
// gvx_set_option(GVX_OPT_COLLIDE_TMA, 1); the A/B measurement is benchmarks/tma_probe.py.
//
// Work split inside a tile: the Bit side of a tile is TILE * 2 chunks of 16 B; thread t reads chunks t, t + 256, ...
// (neighbouring lanes = neighbouring chunks: no bank conflicts). A voxel's two halves sit in lanes 2k and 2k + 1, so
// "any bit but bit 0 set" is the OR of the two lanes' partial results (one shuffle); the even lane then tests the
// Prob byte of that voxel, counts, and writes the eBVM_COLLISION mark back to GLOBAL memory where it changes the voxel.
// Requires off == 0, i_begin == 0, 16-byte aligned bases; no TYPES (those calls keep the plain kernel).
#include <cstdlib>
#include "gvx_internal.h"
#include "gvx_reduce.cuh"
#include "gvx_tma.cuh"

namespace gvx {

namespace {


template <int T> struct VoxelBytes { static constexpr int value = T == kMapBit ? kBitVoxelBytes : 1; };

// "any bit other than bit 0" of one 16-byte half of a BitVectorVoxel (half 0 holds bit 0)
__device__ __forceinline__ uint32_t half_nonzero(const uint4& q, int half) {
  return ((half == 0 ? (q.x & 0xFFFFFFFEu) : q.x) | q.y | q.z | q.w);
}

}  // namespace

template <int AT, int BT, bool MARK, int TILE, int kStages>
__global__ void __launch_bounds__(kThreads)
k_collide_tma(uint8_t* __restrict__ a, const uint8_t* __restrict__ b, long long n, int thr, ReduceScratch* scratch,
              ResultBlock* out_d, ResultBlock* out_h, unsigned long long* count_out_d, PeerReduce pr) {
  static_assert(AT == kMapBit || BT == kMapBit, "Prob x Prob has its own 2 B/voxel kernel");
  constexpr int A_BYTES = TILE * VoxelBytes<AT>::value, B_BYTES = TILE * VoxelBytes<BT>::value;
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int CHUNKS = TILE * 2 / kThreads;  // 16-byte chunks of a Bit tile per thread
  static_assert(TILE * 2 % kThreads == 0 && A_BYTES % 16 == 0 && B_BYTES % 16 == 0, "tile shape");
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ __align__(8) uint64_t full[kStages];

  const long long n_tiles = n / TILE;
  const int t = threadIdx.x;
  uint64_t policy = 0;
  if (t == 0) {
    for (int s = 0; s < kStages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  }
  __syncthreads();

  auto issue = [&](long long tile, int s) {  // thread 0 only
    uint8_t* st = smem + (size_t)s * STAGE_BYTES;
    mbar_expect_tx(&full[s], STAGE_BYTES);
    bulk_load(st, a + (size_t)tile * A_BYTES, A_BYTES, &full[s], policy);
    bulk_load(st + A_BYTES, b + (size_t)tile * B_BYTES, B_BYTES, &full[s], policy);
  };
  if (t == 0) {
    for (int s = 0; s < kStages; ++s) {
      const long long tile = (long long)blockIdx.x + (long long)s * gridDim.x;
      if (tile < n_tiles) issue(tile, s);
    }
  }

  uint32_t cnt = 0;
  const uint32_t acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
  const int half = t & 1;
  long long it = 0;
  for (long long tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++it) {
    const int s = (int)(it % kStages);
    mbar_wait(&full[s], (uint32_t)((it / kStages) & 1));
    const uint8_t* st = smem + (size_t)s * STAGE_BYTES;
    uint4 qa[CHUNKS], qb[CHUNKS];
    int pa[CHUNKS], pb[CHUNKS];
#pragma unroll
    for (int j = 0; j < CHUNKS; ++j) {
      const int c = t + j * kThreads, v = c >> 1;
      if (AT == kMapBit) qa[j] = reinterpret_cast<const uint4*>(st)[c];
      else pa[j] = (int)reinterpret_cast<const int8_t*>(st)[v];
      if (BT == kMapBit) qb[j] = reinterpret_cast<const uint4*>(st + A_BYTES)[c];
      else pb[j] = (int)reinterpret_cast<const int8_t*>(st + A_BYTES)[v];
    }
    __syncthreads();  // every thread holds its part of the stage in registers: the stage may be refilled
    if (t == 0) {
      const long long next = tile + (long long)kStages * gridDim.x;
      if (next < n_tiles) issue(next, s);
    }
#pragma unroll
    for (int j = 0; j < CHUNKS; ++j) {
      const int c = t + j * kThreads;
      const long long i = tile * TILE + (c >> 1);
      bool occ_a, occ_b;
      if (AT == kMapBit) {
        const uint32_t nz = half_nonzero(qa[j], half);
        occ_a = (nz | __shfl_xor_sync(0xffffffffu, nz, 1)) != 0u;
      } else occ_a = pa[j] >= (int)(int8_t)(thr & 0xFF);
      if (BT == kMapBit) {
        const uint32_t nz = half_nonzero(qb[j], half);
        occ_b = (nz | __shfl_xor_sync(0xffffffffu, nz, 1)) != 0u;
      } else occ_b = pb[j] >= (int)(int8_t)((thr >> 8) & 0xFF);
      if (half == 0 && occ_a && occ_b) {  // the even lane owns the voxel (and holds word 0 of a Bit voxel)
        ++cnt;
        if (MARK) {
          if (AT == kMapBit) {
            if (!(qa[j].x & 4u)) *reinterpret_cast<uint32_t*>(a + (size_t)i * kBitVoxelBytes) = qa[j].x | 4u;
          } else if (pa[j] != 127) reinterpret_cast<int8_t*>(a)[i] = 127;
        }
      }
    }
  }
  // the last n % TILE voxels: plain loads, CTA 0
  if (blockIdx.x == 0) {
    for (long long i = n_tiles * TILE + t; i < n; i += kThreads) {
      BitVoxelWords va, vb;
      int xa = 0, xb = 0;
      if (AT == kMapBit) va = ld_voxel(a + (size_t)i * kBitVoxelBytes); else xa = (int)reinterpret_cast<const volatile int8_t*>(a)[i];
      if (BT == kMapBit) vb = ld_voxel_stream(b + (size_t)i * kBitVoxelBytes); else xb = (int)ld_stream_i8(b + i);
      const bool occ_a = AT == kMapBit ? bit_occupied(va) : xa >= (int)(int8_t)(thr & 0xFF);
      const bool occ_b = BT == kMapBit ? bit_occupied(vb) : xb >= (int)(int8_t)((thr >> 8) & 0xFF);
      if (occ_a && occ_b) {
        ++cnt;
        if (MARK) {
          if (AT == kMapBit) {
            if (!(va.w[0] & 4u)) *reinterpret_cast<uint32_t*>(a + (size_t)i * kBitVoxelBytes) = va.w[0] | 4u;
          } else if (xa != 127) reinterpret_cast<int8_t*>(a)[i] = 127;
        }
      }
    }
  }
  finish_block<false>(cnt, acc, scratch, out_d, out_h, count_out_d, pr);
}

template <int AT, int BT, bool MARK, int TILE, int kStages>
static cudaError_t launch_tma_one(const LaunchCtx& c, uint8_t* a, const uint8_t* b, long long n, int thr,
                                  unsigned long long* out) {
  constexpr int stage = TILE * (VoxelBytes<AT>::value + VoxelBytes<BT>::value);
  constexpr int smem = kStages * stage;
  auto kern = k_collide_tma<AT, BT, MARK, TILE, kStages>;
  static bool configured = false;  // per instantiation
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int per_sm = (220 * 1024) / (smem + 1024) > 6 ? 6 : (220 * 1024) / (smem + 1024);
  const long long n_tiles = n / TILE;
  long long grid = (long long)c.sm_count * (per_sm < 1 ? 1 : per_sm);
  if (grid > kMaxGrid) grid = kMaxGrid;
  if (grid > n_tiles) grid = n_tiles;
  if (grid < 1) grid = 1;
  kern<<<(int)grid, kThreads, smem, c.stream>>>(a, b, n, thr, c.scratch_d, c.result_d, c.result_hmap_d, out, c.peer);
  return cudaGetLastError();
}

bool collide_tma_applicable(const void* a, int a_type, const void* b, int b_type, int64_t i_begin, int64_t off,
                            bool types) {
  if (types || off != 0 || i_begin != 0) return false;
  if (a_type != kMapBit && b_type != kMapBit) return false;
  return (((uintptr_t)a | (uintptr_t)b) & 15) == 0;
}

cudaError_t launch_collide_tma(const LaunchCtx& c, void* a_, int a_type, const void* b_, int b_type, int64_t n,
                               int thr, bool mark, unsigned long long* out) {  // thr: both int8 thresholds, packed
  uint8_t* a = (uint8_t*)a_;
  const uint8_t* b = (const uint8_t*)b_;
  // tile / ring shape by option value (c.tma): 1 = 512 voxels x 4 stages (66 KiB, 3 CTAs per SM); the others exist for
  // the sweep of benchmarks/tma_probe.py: 2 = 1024 x 3, 3 = 256 x 6, 4 = 512 x 2, 5 = 1024 x 2
  const int variant = c.tma - 1;
#define GVX_TMA(AT, BT, TILE, ST) return mark ? launch_tma_one<AT, BT, true, TILE, ST>(c, a, b, n, thr, out) \
                                              : launch_tma_one<AT, BT, false, TILE, ST>(c, a, b, n, thr, out)
#define GVX_TMA_PAIR(AT, BT)                \
  switch (variant) {                        \
    case 1: GVX_TMA(AT, BT, 1024, 3);       \
    case 2: GVX_TMA(AT, BT, 256, 6);        \
    case 3: GVX_TMA(AT, BT, 512, 2);        \
    case 4: GVX_TMA(AT, BT, 1024, 2);       \
    default: GVX_TMA(AT, BT, 512, 4);       \
  }
  if (a_type == kMapBit && b_type == kMapBit) GVX_TMA(kMapBit, kMapBit, 256, 4);
  if (a_type == kMapBit) GVX_TMA_PAIR(kMapBit, kMapProb)
  GVX_TMA_PAIR(kMapProb, kMapBit)
#undef GVX_TMA_PAIR
#undef GVX_TMA
}

}  // namespace gvx
