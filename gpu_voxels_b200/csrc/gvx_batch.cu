// gvx_batch.cu -- a batch of robot configurations against one environment map (BASELINE.json configs[2]).
//
// Per configuration the reference's planner does (gvl_ompl_planning/gvl_ompl_planner_helper.cpp:169-202):
//   robot_map.clearMap(); per link transformSubCloud (GV/helpers/MetaPointCloud.cu:624-649, one launch + 64-B H2D +
//   2 syncs each); robot_map.insertMetaPointCloud(clouds, meanings); n = robot_map.collideWith(env)
// i.e. a memset and two full streams over a robot grid for every configuration. Nothing in that answer needs the
// robot grid: n is the number of DISTINCT voxels that (a) receive at least one robot point whose meaning passes the
// collision mask and (b) are occupied in the environment. So:
//   1. k_env_bitmap: the environment map is reduced ONCE per call to 1 bit per voxel (2 MiB for 256^3): L2-resident.
//   2. k_batch_collide: one pass over (configuration, point): fused transform -> voxel index -> one bit test in the
//      L2-resident bitmap -> for the (rare) points inside obstacles an atomicOr on that configuration's "seen" bitmap;
//      the thread that flips the bit counts the voxel. No robot grid, no per-configuration launch, no host sync.
// Configurations are processed in waves sized to the "seen" scratch; results are 64-bit counts per configuration.
#include "gvx_internal.h"
#include "gvx_reduce.cuh"

namespace gvx {

// 1 bit per voxel: DefaultCollider's test of the `other` side (GV/voxel/DefaultCollider.hpp:51-80):
// BitVectorVoxel -> !noneButEmpty(), ProbabilisticVoxel -> occupancy >= threshold.
template <int MAPT>
__global__ void __launch_bounds__(256)
k_env_bitmap(const uint8_t* __restrict__ env, unsigned long long n, int thr, uint32_t* __restrict__ bits) {
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  const unsigned long long stride = (unsigned long long)gridDim.x * 256;
  for (unsigned long long i = (unsigned long long)blockIdx.x * 256 + threadIdx.x; i < n_round; i += stride) {
    bool occ = false;
    if (i < n) {
      if (MAPT == kMapProb) occ = (int)ld_stream_i8(env + i) >= thr;
      else occ = bit_occupied(ld_voxel_stream(env + (size_t)i * kBitVoxelBytes));
    }
    const uint32_t w = __ballot_sync(0xffffffffu, occ);
    if ((threadIdx.x & 31) == 0) bits[i >> 5] = w;
  }
}

constexpr int kBatchMaxClouds = 256;

// grid = (chunks, configurations of this wave). Each CTA walks a strided share of the base cloud under the
// configuration's per-link transforms.
__global__ void __launch_bounds__(kThreads)
k_batch_collide(const float* __restrict__ xyz, BatchDesc d, const uint32_t* __restrict__ env_bits,
                uint32_t* __restrict__ seen, unsigned long long seen_words, uint32_t cfg0,
                unsigned long long* __restrict__ counts) {
  __shared__ float sh_T[kBatchMaxClouds * 12];
  __shared__ uint32_t sh_prefix[kBatchMaxClouds + 1];
  __shared__ uint8_t sh_elig[kBatchMaxClouds];
  __shared__ uint32_t sh_cnt[kThreads / 32];
  const uint32_t cfg = cfg0 + blockIdx.y;
  const float* T = d.transforms_d + (size_t)cfg * d.n_clouds * 16;
  for (uint32_t k = threadIdx.x; k < d.n_clouds * 12; k += kThreads) sh_T[k] = T[(k / 12) * 16 + (k % 12)];
  for (uint32_t k = threadIdx.x; k <= d.n_clouds; k += kThreads) sh_prefix[k] = d.prefix_d[k];
  for (uint32_t k = threadIdx.x; k < d.n_clouds; k += kThreads) sh_elig[k] = d.eligible_d[k];
  __syncthreads();
  uint32_t* my_seen = seen + (size_t)blockIdx.y * seen_words;
  const uint32_t n = sh_prefix[d.n_clouds];
  uint32_t cnt = 0;
  for (uint32_t p = blockIdx.x * kThreads + threadIdx.x; p < n; p += gridDim.x * kThreads) {
    const uint32_t c = find_cloud(sh_prefix, d.n_clouds, p);
    if (!sh_elig[c]) continue;
    const float x = __ldg(xyz + 3 * (size_t)p), y = __ldg(xyz + 3 * (size_t)p + 1), z = __ldg(xyz + 3 * (size_t)p + 2);
    const float* M = sh_T + 12 * c;
    const float tx = affine_row(M, x, y, z), ty = affine_row(M + 4, x, y, z), tz = affine_row(M + 8, x, y, z);
    const uint32_t cx = coord_to_voxel(tx, d.side), cy = coord_to_voxel(ty, d.side), cz = coord_to_voxel(tz, d.side);
    if (cx >= d.dx || cy >= d.dy || cz >= d.dz) continue;
    const unsigned long long idx = ((unsigned long long)cz * d.dy + cy) * d.dx + cx;
    const uint32_t bit = 1u << (idx & 31);
    if (!(__ldg(env_bits + (idx >> 5)) & bit)) continue;
    const uint32_t old = atomicOr(my_seen + (idx >> 5), bit);
    cnt += !(old & bit);
  }
  const uint32_t wsum = __reduce_add_sync(0xffffffffu, cnt);
  if ((threadIdx.x & 31) == 0) sh_cnt[threadIdx.x >> 5] = wsum;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tot = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) tot += sh_cnt[w];
    if (tot) atomicAdd(counts + cfg, tot);
  }
}

cudaError_t launch_env_bitmap(const LaunchCtx& c, const void* env, int env_type, uint64_t n, int8_t thr,
                              uint32_t* bits_d) {
  unsigned long long g = (n + 255) / 256;
  const unsigned long long cap = (unsigned long long)c.sm_count * 8;
  const int grid = (int)(g < 1 ? 1 : (g > cap ? cap : g));
  ++*c.launches;
  if (env_type == kMapProb) k_env_bitmap<kMapProb><<<grid, 256, 0, c.stream>>>((const uint8_t*)env, n, thr, bits_d);
  else k_env_bitmap<kMapBit><<<grid, 256, 0, c.stream>>>((const uint8_t*)env, n, thr, bits_d);
  return cudaGetLastError();
}

cudaError_t launch_batch_collide(const LaunchCtx& c, const float* xyz_d, const BatchDesc& d, uint32_t n_points,
                                 const uint32_t* env_bits_d, uint32_t* seen_d, uint64_t seen_words, uint32_t cfg0,
                                 uint32_t n_cfg_wave, unsigned long long* counts_d) {
  if (d.n_clouds > (uint32_t)kBatchMaxClouds) return cudaErrorInvalidValue;
  // enough CTAs per configuration to fill the machine when the wave is small; at least 8 points per thread
  uint32_t chunks = (n_points + kThreads * 8 - 1) / (kThreads * 8);
  const uint32_t want = (uint32_t)((c.sm_count * 32 + n_cfg_wave - 1) / n_cfg_wave);  // ~4 waves of 8 CTAs/SM
  if (chunks > want && want >= 1) chunks = want > 1 ? want : 1;
  if (chunks < 1) chunks = 1;
  ++*c.launches;
  k_batch_collide<<<dim3(chunks, n_cfg_wave), kThreads, 0, c.stream>>>(xyz_d, d, env_bits_d, seen_d, seen_words, cfg0,
                                                                       counts_d);
  return cudaGetLastError();
}

}  // namespace gvx
