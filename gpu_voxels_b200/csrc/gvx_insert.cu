// gvx_insert.cu -- point-cloud voxelisation for sm_100a.
//
// Replaces kernelInsertGlobalPointCloud / kernelInsertMetaPointCloud x2
// (GV/voxelmap/kernels/VoxelMapOperations.hpp:241-264, 302-340, 386-437), kernelTransformCloud
// (GV/helpers/kernels/MetaPointCloudOperations.cu:77-91), kernelTransformSensorData
// (GV/voxelmap/kernels/VoxelMapOperations.cu:27-71), kernelInsertSensorData without ray casting
// (GV/voxelmap/kernels/VoxelMapOperations.hpp:559-613) and kernelClearVoxelMap(bits) (:45-72).
//
// The fp32 point transform, the float->voxel index computation and the scatter are ONE kernel. The
// reference's non-atomic byte read-modify-write (BitVector::setBit) loses bits when two meanings of the
// same byte hit one voxel concurrently; here a BitVectorVoxel insert is an atomicOr on the 32-bit word
// that holds the meaning bit, issued once per (voxel, meaning) group of a warp (__match_any_sync), and the
// probabilistic log-odds update is a saturating add applied with one CAS per (voxel) group of a warp, so
// results equal the sequential (race-free) semantics and are deterministic.
#include <cmath>

#include "gvx_internal.h"
#include "gvx_sensor.cuh"
#include "gvx_tma.cuh"

namespace gvx {

__device__ __forceinline__ int8_t prob_insert_value(int meaning) {
  // ProbabilisticVoxel::insert, GV/voxel/ProbabilisticVoxel.hpp:77-92
  return meaning == 0 ? (int8_t)-127 : ((meaning == 1 || meaning == 2) ? (int8_t)127 : (int8_t)-128);
}

// Publish the "points outside the map" flag: OR into flag_accum, last CTA moves it to the result block and
// resets it (self-cleaning, no memset between launches).
__device__ __forceinline__ void finish_flag(bool outside, ReduceScratch* s, ResultBlock* out_d, ResultBlock* out_h) {
  __shared__ bool sh_last;
  const bool any = __syncthreads_or(outside);
  if (threadIdx.x == 0) {
    if (any) atomicOr(&s->flag_accum, 1u);
    __threadfence();
    const unsigned int t = atomicInc(&s->ticket, gridDim.x - 1);
    sh_last = (t == gridDim.x - 1);
    if (sh_last) {
      __threadfence();
      const unsigned int v = atomicExch(&s->flag_accum, 0u);
      out_d->flag = v;
      out_h->flag = v;
      __threadfence_system();
    }
  }
}

template <bool PROB, bool META>
__global__ void __launch_bounds__(kInsThreads)
k_insert(MapView m, const float* __restrict__ xyz, unsigned long long n, int meaning, MetaDesc d,
         bool want_flag, ReduceScratch* s, ResultBlock* out_d, ResultBlock* out_h) {
  extern __shared__ uint32_t sh_prefix[];  // META: n_clouds + 1 entries when they fit, else unused
  const uint32_t* prefix = d.prefix_d;
  if (META && d.n_clouds + 1 <= 1024) {
    for (uint32_t k = threadIdx.x; k <= d.n_clouds; k += kInsThreads) sh_prefix[k] = d.prefix_d[k];
    __syncthreads();
    prefix = sh_prefix;
  }
  const uint64_t slab = (uint64_t)m.dx * m.dy;
  bool outside = false;
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  // uniform trip count per warp so the warp-collective match below is always convergent
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n_round; i += stride) {
    bool live = i < n;
    unsigned long long key = ~0ull - (threadIdx.x & 31);  // unique per lane => no peers
    uint64_t idx = 0;
    int mg = meaning;
    if (live) {
      float x = __ldg(xyz + 3 * i), y = __ldg(xyz + 3 * i + 1), z = __ldg(xyz + 3 * i + 2);
      if (META) {
        const uint32_t c = find_cloud(prefix, d.n_clouds, (uint32_t)i);
        mg = d.meanings_d[d.n_meanings == 1 ? 0 : c];
        if (d.transforms_d) {
          const float* M = d.transforms_d + 16 * (size_t)c;
          const float tx = affine_row(M, x, y, z), ty = affine_row(M + 4, x, y, z), tz = affine_row(M + 8, x, y, z);
          x = tx; y = ty; z = tz;
        }
      }
      const uint32_t cx = coord_to_voxel(x, m.side), cy = coord_to_voxel(y, m.side), cz = coord_to_voxel(z, m.side);
      if (cx < m.dx && cy < m.dy && cz < m.dz) {
        live = cz >= m.z0 && cz < m.z1;  // another slab's voxel: not outside, not ours
        idx = (uint64_t)(cz - m.z0) * slab + (uint64_t)cy * m.dx + cx;
        key = (idx << 8) | (uint32_t)(mg & 255);
      } else {
        outside = true;
        live = false;
      }
    }
    if (PROB) {
      if (live) {
        const int8_t v = prob_insert_value(mg);
        ((int8_t*)m.data)[idx] = v;  // same value from every writer
        if (m.occ && v != (int8_t)-128) summary_set(m.occ, idx);  // "known" summary
      }
    } else {
      const unsigned peers = __match_any_sync(0xffffffffu, key);
      if (live && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
        const uint32_t bit = (uint32_t)(mg & 255);
        atomicOr((uint32_t*)m.data + idx * 8 + (bit >> 5), 1u << (bit & 31));
        // occupancy summary: every meaning except eBVM_FREE (bit 0) makes the voxel occupied
        if (m.occ && bit != 0u) summary_set(m.occ, idx);
      }
    }
  }
  if (want_flag) finish_flag(outside, s, out_d, out_h);
}

__global__ void __launch_bounds__(kInsThreads)
k_transform(const float* in, float* out, unsigned long long n, const Affine12 Mv) {  // in == out is allowed
  const float* M = Mv.a;
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n; i += stride) {
    const float x = in[3 * i], y = in[3 * i + 1], z = in[3 * i + 2];
    const float tx = affine_row(M, x, y, z), ty = affine_row(M + 4, x, y, z), tz = affine_row(M + 8, x, y, z);
    out[3 * i] = tx; out[3 * i + 1] = ty; out[3 * i + 2] = tz;
  }
}

__device__ __forceinline__ bool robot_bit_set(const uint8_t* __restrict__ robot, uint64_t voxel, uint32_t bit) {
  return (__ldg(robot + voxel * kBitVoxelBytes + (bit >> 3)) >> (bit & 7)) & 1u;
}

// insertSensorData without ray casting: sensor-frame point -> world (pose = [R|t], re-projection onto z = 0
// with the reference's double-precision factor when the world point is below the floor), NaN filter, bounds
// test on the point and on the sensor origin, updateOccupancy(+delta) at the hit voxel.
__global__ void __launch_bounds__(kInsThreads)
k_update_logodds(MapView m, const float* __restrict__ xyz, unsigned long long n, const Affine12 pose, int delta,
                 const uint8_t* __restrict__ robot, uint32_t robot_bit) {
  pdl_wait();               // the kernel before this one (a collide on this map, as a rule) is complete
  pdl_launch_dependents();  // every CTA of this grid is resident at once: the next kernel may move in behind it
  const float* P = pose.a;
  const uint32_t sx = coord_to_voxel(P[3], m.side), sy = coord_to_voxel(P[7], m.side), sz = coord_to_voxel(P[11], m.side);
  const bool sensor_inside = sx < m.dx && sy < m.dy && sz < m.dz;
  const uint64_t slab = (uint64_t)m.dx * m.dy;
  const unsigned long long nvox = slab * m.dz;
  if (n > nvox) n = nvox;  // the reference's loop bound, VoxelMapOperations.hpp:569
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n_round; i += stride) {
    bool live = false;
    uint64_t idx = 0;
    if (i < n) {
      const float rx = __ldg(xyz + 3 * i), ry = __ldg(xyz + 3 * i + 1), rz = __ldg(xyz + 3 * i + 2);
      float wx, wy, wz;
      sensor_to_world(P, rx, ry, rz, wx, wy, wz);
      if (!(isnan(wx) || isnan(wy) || isnan(wz))) {
        const uint32_t cx = coord_to_voxel(wx, m.side), cy = coord_to_voxel(wy, m.side), cz = coord_to_voxel(wz, m.side);
        if (cx < m.dx && cy < m.dy && cz < m.dz && sensor_inside && cz >= m.z0 && cz < m.z1) {
          live = true;
          idx = (uint64_t)(cz - m.z0) * slab + (uint64_t)cy * m.dx + cx;
          // cut_real_robot, VoxelMapOperations.hpp:578-592: the sensor sees the robot there, no update
          if (robot) live = !robot_bit_set(robot, (uint64_t)cz * slab + (uint64_t)cy * m.dx + cx, robot_bit);
        }
      }
    }
    const unsigned long long key = live ? idx : (~0ull - (threadIdx.x & 31));
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    if (live && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
      const bool first = logodds_add((int8_t*)m.data, idx, delta * __popc(peers));
      if (m.occ && first) summary_set(m.occ, idx);  // first touch: the voxel becomes "known" (and never UNKNOWN again)
    }
  }
}

// Free-space pass of insertSensorData with ray casting (RayCaster::rayCast, GV/voxelmap/kernels/VoxelMapOperations.h:
// 199-323): every live point walks the integer line from the sensor's voxel to its hit voxel, 1 + |dx| + |dy| + |dz|
// cells with both ends, and applies updateOccupancy(delta_free < 0) to each. This pass holds ONLY decrements, which
// commute under the saturating clamp; the hits follow in k_update_logodds (the defined order).
// The lanes of a warp are neighbouring pixels: their rays leave the sensor through the same voxels, so each step folds
// equal voxels into one update (match_any), and a voxel already at MIN_PROBABILITY (-127) is skipped after one L2 read
// -- the cone in front of the sensor saturates after a few frames and then costs no atomics at all.
__global__ void __launch_bounds__(kInsThreads)
k_raycast_free(MapView m, const float* __restrict__ xyz, unsigned long long n, const Affine12 pose, int delta_free,
               const uint8_t* __restrict__ robot, uint32_t robot_bit) {
  const float* P = pose.a;
  const uint32_t sx = coord_to_voxel(P[3], m.side), sy = coord_to_voxel(P[7], m.side), sz = coord_to_voxel(P[11], m.side);
  if (!(sx < m.dx && sy < m.dy && sz < m.dz)) return;
  const uint64_t slab = (uint64_t)m.dx * m.dy;
  const unsigned long long nvox = slab * m.dz;
  if (n > nvox) n = nvox;
  const unsigned lane = threadIdx.x & 31;
  int8_t* map = (int8_t*)m.data;
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n_round; i += stride) {
    int32_t px = (int32_t)sx, py = (int32_t)sy, pz = (int32_t)sz, steps = 0;
    int32_t dx2 = 0, dy2 = 0, dz2 = 0, ix = 0, iy = 0, iz = 0, e_xy = 0, e_xz = 0, e_yz = 0;
    if (i < n) {
      const float rx = __ldg(xyz + 3 * i), ry = __ldg(xyz + 3 * i + 1), rz = __ldg(xyz + 3 * i + 2);
      float wx, wy, wz;
      sensor_to_world(P, rx, ry, rz, wx, wy, wz);
      if (!(isnan(wx) || isnan(wy) || isnan(wz))) {
        const uint32_t cx = coord_to_voxel(wx, m.side), cy = coord_to_voxel(wy, m.side), cz = coord_to_voxel(wz, m.side);
        if (cx < m.dx && cy < m.dy && cz < m.dz &&
            !(robot && robot_bit_set(robot, (uint64_t)cz * slab + (uint64_t)cy * m.dx + cx, robot_bit))) {
          int32_t ddx, ddy, ddz;
          if (cx > sx) { ddx = (int32_t)(cx - sx); ix = 1; } else { ddx = (int32_t)(sx - cx); ix = -1; }
          if (cy > sy) { ddy = (int32_t)(cy - sy); iy = 1; } else { ddy = (int32_t)(sy - cy); iy = -1; }
          if (cz > sz) { ddz = (int32_t)(cz - sz); iz = 1; } else { ddz = (int32_t)(sz - cz); iz = -1; }
          steps = 1 + ddx + ddy + ddz;
          e_xy = ddx - ddy; e_xz = ddx - ddz; e_yz = ddy - ddz;
          dx2 = 2 * ddx; dy2 = 2 * ddy; dz2 = 2 * ddz;
        }
      }
    }
    const int32_t max_steps = __reduce_max_sync(0xffffffffu, steps);
    for (int32_t s = 0; s < max_steps; ++s) {
      const bool act = s < steps;
      bool need = false;
      uint64_t idx = 0;
      int v = 0;
      if (act && (uint32_t)pz >= m.z0 && (uint32_t)pz < m.z1) {
        idx = (uint64_t)((uint32_t)pz - m.z0) * slab + (uint64_t)py * m.dx + (uint64_t)px;
        v = (int)__ldcg(map + idx);
        need = v != -127;  // at the floor already: further decrements change nothing
      }
      if (__any_sync(0xffffffffu, need)) {
        const unsigned long long key = need ? idx : (~0ull - lane);
        const unsigned peers = __match_any_sync(0xffffffffu, key);
        if (need && lane == (unsigned)(__ffs(peers) - 1)) {
          logodds_add(map, idx, delta_free * __popc(peers));
          if (m.occ && v == -128) summary_set(m.occ, idx);  // first touch: the voxel becomes "known"
        }
      }
      if (act) {
        if (e_xy > 0 && e_xz > 0) { px += ix; e_xy -= dy2; e_xz -= dz2; }
        else if (e_yz > 0)        { py += iy; e_xy += dx2; e_yz -= dz2; }
        else                      { pz += iz; e_xz += dx2; e_yz += dy2; }
      }
    }
  }
}

// Several sensor frames (each with its own pose) in ONE launch: frame f owns points [prefix[f], prefix[f+1]).
// Same per-point arithmetic as k_update_logodds. Used when a z-slab rank folds the whole multi-sensor cloud and keeps
// only its slab's points: one launch instead of one per frame.
constexpr int kMaxFramesSmem = 64;
constexpr int kFoldStages = 2;  // k_update_logodds_frames: per-warp ring of 1.5 KB point chunks filled by bulk copies
constexpr int kFoldDrain = 128;  // k_update_logodds_frames: queued hits a warp applies at once (4 per lane)
__global__ void __launch_bounds__(kInsThreads)
k_update_logodds_multi(MapView m, const float* __restrict__ xyz, const uint32_t* __restrict__ prefix_g,
                       const float* __restrict__ poses_g, uint32_t n_frames, int delta) {
  __shared__ float sh_pose[kMaxFramesSmem * 12];
  __shared__ uint32_t sh_prefix[kMaxFramesSmem + 1];
  __shared__ uint8_t sh_inside[kMaxFramesSmem];
  pdl_wait();
  pdl_launch_dependents();
  const bool in_smem = n_frames <= (uint32_t)kMaxFramesSmem;
  if (in_smem) {
    for (uint32_t k = threadIdx.x; k < n_frames * 12; k += kInsThreads) sh_pose[k] = poses_g[k];
    for (uint32_t k = threadIdx.x; k <= n_frames; k += kInsThreads) sh_prefix[k] = prefix_g[k];
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < n_frames; k += kInsThreads) {
      const float* P = sh_pose + 12 * k;
      sh_inside[k] = coord_to_voxel(P[3], m.side) < m.dx && coord_to_voxel(P[7], m.side) < m.dy &&
                     coord_to_voxel(P[11], m.side) < m.dz;
    }
    __syncthreads();
  }
  const uint32_t* prefix = in_smem ? sh_prefix : prefix_g;
  const float* poses = in_smem ? sh_pose : poses_g;
  const uint64_t slab = (uint64_t)m.dx * m.dy;
  const unsigned long long n = prefix[n_frames];
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n_round; i += stride) {
    bool live = false;
    uint64_t idx = 0;
    if (i < n && i - prefix[find_cloud(prefix, n_frames, (uint32_t)i)] < slab * m.dz) {  // :569: a frame is cut at the voxel count
      const uint32_t f = find_cloud(prefix, n_frames, (uint32_t)i);
      const float* P = poses + 12 * (size_t)f;
      const bool sensor_inside = in_smem ? (bool)sh_inside[f]
                                         : (coord_to_voxel(P[3], m.side) < m.dx && coord_to_voxel(P[7], m.side) < m.dy &&
                                            coord_to_voxel(P[11], m.side) < m.dz);
      const float rx = __ldg(xyz + 3 * i), ry = __ldg(xyz + 3 * i + 1), rz = __ldg(xyz + 3 * i + 2);
      float wx = affine_row(P, rx, ry, rz), wy = affine_row(P + 4, rx, ry, rz), wz = affine_row(P + 8, rx, ry, rz);
      if (wz < 0.0f) {
        const double factor = (double)__fdiv_rn(P[11], __fsub_rn(P[11], wz));
        const float qx = (float)((double)rx * factor), qy = (float)((double)ry * factor), qz = (float)((double)rz * factor);
        wx = affine_row(P, qx, qy, qz); wy = affine_row(P + 4, qx, qy, qz); wz = affine_row(P + 8, qx, qy, qz);
      }
      if (sensor_inside && !(isnan(wx) || isnan(wy) || isnan(wz))) {
        // z first: on a z-slab rank most points of the rig's cloud belong to other slabs and skip two divisions
        const uint32_t cz = coord_to_voxel(wz, m.side);
        if (cz >= m.z0 && cz < m.z1) {
          const uint32_t cx = coord_to_voxel(wx, m.side), cy = coord_to_voxel(wy, m.side);
          if (cx < m.dx && cy < m.dy) {
            live = true;
            idx = (uint64_t)(cz - m.z0) * slab + (uint64_t)cy * m.dx + cx;
          }
        }
      }
    }
    const unsigned long long key = live ? idx : (~0ull - (threadIdx.x & 31));
    const unsigned peers = __match_any_sync(0xffffffffu, key);
    if (live && (threadIdx.x & 31) == (unsigned)(__ffs(peers) - 1)) {
      const bool first = logodds_add((int8_t*)m.data, idx, delta * __popc(peers));
      if (m.occ && first) summary_set(m.occ, idx);  // first touch: the voxel becomes "known" (and never UNKNOWN again)
    }
  }
}

// Up to kMaxFramesByValue frames with their poses in the kernel PARAMETERS (nothing for the host to stage): grid.y is
// the frame, so a CTA works with one pose held in registers exactly like k_update_logodds. The world z of a point is
// computed first: on a z-slab rank most points of a multi-sensor cloud belong to other slabs and are dropped after
// three FMAs and one division.
__global__ void __launch_bounds__(kInsThreads)
k_update_logodds_frames(MapView m, const float* __restrict__ xyz, const __grid_constant__ FramesByValue fv, int delta,
                        float z_lo, float z_hi) {
  pdl_wait();
  pdl_launch_dependents();
  const uint32_t f = fv.order[blockIdx.y];
  float P[12];
#pragma unroll
  for (int k = 0; k < 12; ++k) P[k] = fv.poses[12 * f + k];
  if (!(coord_to_voxel(P[3], m.side) < m.dx && coord_to_voxel(P[7], m.side) < m.dy && coord_to_voxel(P[11], m.side) < m.dz))
    return;  // sensor origin outside the map: the frame contributes nothing (VoxelMapOperations.hpp:577-580)
  const uint64_t slab = (uint64_t)m.dx * m.dy;
  const unsigned long long first = fv.prefix[f], n_frame = fv.prefix[f + 1] - fv.prefix[f], nvox = slab * m.dz;
  const unsigned long long n = n_frame < nvox ? n_frame : nvox;  // the reference's loop bound, VoxelMapOperations.hpp:569
  // U points per thread and step, their loads issued together: with several frames per launch a thread runs many steps,
  // and one dependent load -> divide -> atomic chain per step would leave the SMs waiting on latency
  constexpr int U = 4;
  // Every WARP owns a ring of kFoldStages shared-memory stages of 128 points (1.5 KB): lane 0 issues one bulk copy per
  // stage (cp.async.bulk, SASS UBLKCP) kFoldStages chunks ahead, an mbarrier counts the bytes, the lanes then take their
  // U points each (x/y/z sit 12 bytes apart: conflict-free stride-3 reads). HBM keeps streaming while the warp classifies
  // the chunk before, and there is no __syncthreads anywhere: a warp that drains its hit queue (a multi-microsecond chain
  // of DRAM misses) never stalls the other warps of its CTA. (A CTA-wide load -> __syncthreads -> compute loop filtered
  // the frame at 1.7 TB/s on a z-slab rank, where 7 of 8 points are only read and dropped.)
  constexpr int kWarps = kInsThreads / 32;
  constexpr uint32_t kChunkPts = 32 * U, kChunkBytes = kChunkPts * 12;
  __shared__ __align__(128) float sh_xyz[kWarps][kFoldStages][kChunkPts * 3];
  __shared__ __align__(8) uint64_t sh_full[kWarps][kFoldStages];
  const unsigned lane = threadIdx.x & 31, lt = (1u << lane) - 1u;
  const int warp = threadIdx.x >> 5;
  // bulk copies need 16-byte aligned source chunks: the frame must start on a multiple of 4 points of a 16-byte aligned array
  const bool staged = ((first & 3ull) == 0) && ((reinterpret_cast<uintptr_t>(xyz) & 15) == 0);
  const unsigned long long n_chunks = (n + kChunkPts - 1) / kChunkPts, n_full = staged ? n / kChunkPts : 0;
  const unsigned long long gw = (unsigned long long)blockIdx.x * kWarps + warp, n_warps = (unsigned long long)gridDim.x * kWarps;
  uint64_t policy = 0;
  if (lane == 0) {
    for (int st = 0; st < kFoldStages; ++st) mbar_init(&sh_full[warp][st], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  }
  __syncwarp();
  auto issue = [&](unsigned long long chunk, int st) {  // lane 0 only; full chunks only
    mbar_expect_tx(&sh_full[warp][st], kChunkBytes);
    bulk_load(sh_xyz[warp][st], xyz + 3 * (first + chunk * kChunkPts), kChunkBytes, &sh_full[warp][st], policy);
  };
  if (lane == 0) {
    for (int st = 0; st < kFoldStages; ++st) {
      const unsigned long long chunk = gw + (unsigned long long)st * n_warps;
      if (chunk < n_full) issue(chunk, st);
    }
  }
  // Candidates are not classified where they are found. On a z-slab rank most lanes of a warp hold points of OTHER slabs:
  // the cheap test (world z against the slab's two float bounds) runs for every point, but the rest -- two more IEEE
  // divisions, bounds, index, and the read-modify-write chain (DRAM miss -> CAS -> maybe the summary word) -- costs a warp
  // the same with 4 live lanes as with 32. Every warp appends its candidate POINTS to a small queue in shared memory and
  // drains it 128 at a time, 4 per lane with all lanes busy and their loads issued together: instructions and latency
  // chains follow the number of hits, not the number of points read.
  __shared__ float sh_qx[kWarps][kFoldDrain + 32], sh_qy[kWarps][kFoldDrain + 32], sh_qz[kWarps][kFoldDrain + 32];
  float *qx = sh_qx[warp], *qy = sh_qy[warp], *qz = sh_qz[warp];  // per warp: < kFoldDrain waiting + one append
  int qn = 0;  // warp-uniform number of queued candidates
  int8_t* map = (int8_t*)m.data;
  auto drain = [&](int cnt) {  // the last cnt (<= kFoldDrain) queued candidates: entry j * 32 + lane of them for j = 0..3
    __syncwarp();
    constexpr int J = kFoldDrain / 32;
    bool lead[J];
    uint64_t id[J];
    int tot[J];
    uint32_t oldw[J], got[J];
#pragma unroll
    for (int j = 0; j < J; ++j) {
      lead[j] = false; id[j] = 0; tot[j] = 0;
      if (j * 32 >= cnt) continue;  // warp-uniform
      const int k = j * 32 + (int)lane;
      bool live = false;
      if (k < cnt) {  // the arithmetic of k_update_logodds, point by point
        const float rx = qx[qn - cnt + k], ry = qy[qn - cnt + k], rz = qz[qn - cnt + k];
        float wx, wy, wz;
        sensor_to_world(P, rx, ry, rz, wx, wy, wz);
        if (!(isnan(wx) || isnan(wy) || isnan(wz))) {
          const uint32_t cx = coord_to_voxel(wx, m.side), cy = coord_to_voxel(wy, m.side), cz = coord_to_voxel(wz, m.side);
          if (cx < m.dx && cy < m.dy && cz >= m.z0 && cz < m.z1) {
            live = true;
            id[j] = (uint64_t)(cz - m.z0) * slab + (uint64_t)cy * m.dx + cx;
          }
        }
      }
      // equal voxels of one round fold into one saturating update (exact for same-sign deltas)
      const unsigned peers = __match_any_sync(0xffffffffu, live ? id[j] : (~0ull - lane));
      lead[j] = live && lane == (unsigned)(__ffs(peers) - 1);
      tot[j] = delta * __popc(peers);
    }
    // the read-modify-writes of a lane go out together: loads of the old words, then the CAS of each, then (rare) the
    // retry loop for a CAS that lost a race. One latency chain instead of J.
#pragma unroll
    for (int j = 0; j < J; ++j)
      if (lead[j]) oldw[j] = __ldcg(reinterpret_cast<const uint32_t*>(map + (id[j] & ~3ull)));
#pragma unroll
    for (int j = 0; j < J; ++j) {
      if (!lead[j]) continue;
      const int shift = (int)(id[j] & 3ull) * 8;
      int v = (int)(int8_t)((oldw[j] >> shift) & 0xFFu) + tot[j];
      v = v < -127 ? -127 : (v > 127 ? 127 : v);
      const uint32_t neww = (oldw[j] & ~(0xFFu << shift)) | (((uint32_t)v & 0xFFu) << shift);
      got[j] = oldw[j];
      if (neww != oldw[j]) got[j] = atomicCAS(reinterpret_cast<uint32_t*>(map + (id[j] & ~3ull)), oldw[j], neww);
    }
    // first touch of a voxel (its byte was UNKNOWN in the word this lane read): only then can the "known" bit be missing
    // from the summary -- every other hit skips the summary word, the second random DRAM sector of an update
#pragma unroll
    for (int j = 0; j < J; ++j) {
      if (!lead[j]) continue;
      bool first_touch = ((oldw[j] >> ((int)(id[j] & 3ull) * 8)) & 0xFFu) == 0x80u;
      if (got[j] != oldw[j]) first_touch |= logodds_add(map, id[j], tot[j]);  // another update got in between: redo
      if (m.occ && first_touch) atomicOr(m.occ + (id[j] >> 5), 1u << (id[j] & 31));
    }
    qn -= cnt;
    __syncwarp();
  };
  unsigned long long it = 0;
  for (unsigned long long chunk = gw; chunk < n_chunks; chunk += n_warps, ++it) {  // warp-uniform
    const unsigned long long base = chunk * kChunkPts;
    float rx[U], ry[U], rz[U];
    bool have[U];
    if (chunk < n_full) {
      const int st = (int)(it % kFoldStages);
      mbar_wait(&sh_full[warp][st], (uint32_t)((it / kFoldStages) & 1));
      const float* sp = sh_xyz[warp][st];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t k = u * 32 + lane;
        have[u] = true;
        rx[u] = sp[3 * k]; ry[u] = sp[3 * k + 1]; rz[u] = sp[3 * k + 2];
      }
      __syncwarp();  // every lane holds its points in registers: the stage may be refilled
      if (lane == 0) {
        const unsigned long long next = chunk + (unsigned long long)kFoldStages * n_warps;
        if (next < n_full) issue(next, st);
      }
    } else {  // the partial last chunk of a frame, or an unaligned frame: plain loads
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const unsigned long long k = base + (unsigned long long)u * 32 + lane;
        have[u] = k < n;
        if (have[u]) {
          const float* p = xyz + 3 * (first + k);
          rx[u] = __ldg(p); ry[u] = __ldg(p + 1); rz[u] = __ldg(p + 2);
        }
      }
    }
    // the cheap test, for every point read: a point whose world z is negative goes through the re-projection (it can end
    // anywhere), every other point is a candidate iff z_lo <= wz < z_hi, which is (cz >= m.z0 && cz < m.z1) exactly
    // (fold_z_bound); NaN fails both
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const float wz = affine_row(P + 8, rx[u], ry[u], rz[u]);
      const bool cand = have[u] && (wz < 0.0f || (wz >= z_lo && wz < z_hi));
      const unsigned bal = __ballot_sync(0xffffffffu, cand);
      if (!bal) continue;  // a warp of another slab's points
      if (cand) {
        const int pos = qn + __popc(bal & lt);
        qx[pos] = rx[u]; qy[pos] = ry[u]; qz[pos] = rz[u];
      }
      qn += __popc(bal);
      if (qn >= kFoldDrain) drain(kFoldDrain);
    }
  }
  while (qn > 0) drain(qn < kFoldDrain ? qn : kFoldDrain);
}

// BitVoxelMap::clearBits / clearBit: voxel &= ~mask, written back only where it changes. When the map carries an
// occupancy summary the warp re-derives its 32 voxels' summary word from the result (warps cover aligned 32-voxel
// groups: block size and stride are multiples of 32).
__global__ void __launch_bounds__(kInsThreads)
k_clear_bits(uint8_t* data, unsigned long long n, BitVoxelWords mask, uint32_t* occ) {
  const unsigned long long stride = (unsigned long long)gridDim.x * kInsThreads;
  const unsigned long long n_round = (n + 31ull) & ~31ull;
  for (unsigned long long i = (unsigned long long)blockIdx.x * kInsThreads + threadIdx.x; i < n_round; i += stride) {
    bool occupied = false;
    if (i < n) {
      BitVoxelWords v = ld_voxel(data + i * kBitVoxelBytes);
      uint32_t hit = 0;
#pragma unroll
      for (int k = 0; k < 8; ++k) { hit |= v.w[k] & mask.w[k]; v.w[k] &= ~mask.w[k]; }
      if (hit) {
        uint4* p = reinterpret_cast<uint4*>(data + i * kBitVoxelBytes);
        p[0] = make_uint4(v.w[0], v.w[1], v.w[2], v.w[3]);
        p[1] = make_uint4(v.w[4], v.w[5], v.w[6], v.w[7]);
      }
      occupied = bit_occupied(v);
    }
    if (occ) {
      const uint32_t w = __ballot_sync(0xffffffffu, occupied);
      if ((threadIdx.x & 31) == 0) occ[i >> 5] = w;
    }
  }
}

static int grid_for(const LaunchCtx& c, unsigned long long n, int per_sm) {
  unsigned long long g = (n + kInsThreads - 1) / kInsThreads;
  const unsigned long long cap = (unsigned long long)c.sm_count * per_sm;
  if (g > cap) g = cap;
  if (g > (unsigned long long)kMaxGrid) g = kMaxGrid;
  return g < 1 ? 1 : (int)g;
}

cudaError_t launch_insert_points(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n, int meaning,
                                 bool want_flag) {
  const int grid = grid_for(c, n, 8);
  MetaDesc d{};
  ++*c.launches;
  if (m.type == kMapProb) k_insert<true, false><<<grid, kInsThreads, 0, c.stream>>>(m, xyz_d, n, meaning, d, want_flag, c.scratch_d, c.result_d, c.result_hmap_d);
  else k_insert<false, false><<<grid, kInsThreads, 0, c.stream>>>(m, xyz_d, n, meaning, d, want_flag, c.scratch_d, c.result_d, c.result_hmap_d);
  return cudaGetLastError();
}

cudaError_t launch_insert_meta(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n, const MetaDesc& d,
                               bool want_flag) {
  const int grid = grid_for(c, n, 8);
  const size_t smem = d.n_clouds + 1 <= 1024 ? (d.n_clouds + 1) * sizeof(uint32_t) : 0;
  ++*c.launches;
  if (m.type == kMapProb) k_insert<true, true><<<grid, kInsThreads, smem, c.stream>>>(m, xyz_d, n, 0, d, want_flag, c.scratch_d, c.result_d, c.result_hmap_d);
  else k_insert<false, true><<<grid, kInsThreads, smem, c.stream>>>(m, xyz_d, n, 0, d, want_flag, c.scratch_d, c.result_d, c.result_hmap_d);
  return cudaGetLastError();
}

cudaError_t launch_transform(const LaunchCtx& c, const float* in_d, float* out_d, uint64_t n, const Affine12& M) {
  ++*c.launches;
  k_transform<<<grid_for(c, n, 8), kInsThreads, 0, c.stream>>>(in_d, out_d, n, M);
  return cudaGetLastError();
}

cudaError_t launch_update_logodds(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                  const Affine12& pose, int delta, const void* robot_d, uint32_t robot_bit) {
  ++*c.launches;
  return pdl_launch(k_update_logodds, grid_for(c, n, 8), kInsThreads, c.stream, c.pdl, m, xyz_d, n, pose, delta,
                    (const uint8_t*)robot_d, robot_bit);
}

cudaError_t launch_raycast_free(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                const Affine12& pose, int delta_free, const void* robot_d, uint32_t robot_bit) {
  ++*c.launches;
  k_raycast_free<<<grid_for(c, n, 8), kInsThreads, 0, c.stream>>>(m, xyz_d, n, pose, delta_free, (const uint8_t*)robot_d,
                                                                  robot_bit);
  return cudaGetLastError();
}

// Smallest float w >= 0 whose voxel layer floor(w / side) is >= z. The device computes a layer as
// (uint32)floorf(__fdiv_rn(w, side)) (coord_to_voxel); IEEE division, floorf and the conversion are all monotone in w, so
// for every non-negative, non-NaN w:  layer(w) >= z  <=>  w >= fold_z_bound(z). The host's float division is the same
// correctly rounded operation (this file is built with -ffp-contract=off, no fast-math), so the bound is found by
// stepping single ulps around z * side. A z-slab rank tests lo <= wz < hi instead of dividing every point of the frame.
static float fold_z_bound(uint32_t z, float side) {
  if (z == 0) return 0.0f;
  auto layer_ge = [&](float w) { return (double)floorf(w / side) >= (double)z; };
  float w = (float)z * side;
  for (int k = 0; k < 64 && layer_ge(w); ++k) w = nextafterf(w, -INFINITY);
  for (int k = 0; k < 128 && !layer_ge(w); ++k) w = nextafterf(w, INFINITY);
  return w;
}

cudaError_t launch_update_logodds_multi(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                        const uint32_t* prefix_d, const float* poses_d, uint32_t n_frames, int delta,
                                        const FramesByValue* by_value) {
  ++*c.launches;
  if (by_value) {
    uint32_t longest = 0;
    for (uint32_t f = 0; f < n_frames; ++f) {
      const uint32_t len = by_value->prefix[f + 1] - by_value->prefix[f];
      if (len > longest) longest = len;
    }
    // 4 points per thread and step. The grid is NOT divided among the frames: on a z-slab rank most frames belong to
    // other slabs and their CTAs retire after one read, leaving the SMs to the frame(s) that do the atomics
    // ... but the whole grid is ONE wave of co-resident CTAs (4 per SM: 62 registers, 46 KB of shared memory), so that a CTA
    // runs many chunks: its bulk-copy ring stays full and its warps' hit queues fill up before they are drained
    int gx = grid_for(c, (longest + 3) / 4, 4);
    const int per_frame = (c.sm_count * 4 + (int)n_frames - 1) / (int)n_frames;
    if (gx > per_frame) gx = per_frame < 1 ? 1 : per_frame;
    return pdl_launch(k_update_logodds_frames, dim3((unsigned)gx, n_frames), kInsThreads, c.stream, c.pdl, m, xyz_d,
                      *by_value, delta, fold_z_bound(m.z0, m.side), fold_z_bound(m.z1, m.side));
  }
  return pdl_launch(k_update_logodds_multi, dim3((unsigned)grid_for(c, n, 8)), kInsThreads, c.stream, c.pdl, m, xyz_d,
                    prefix_d, poses_d, n_frames, delta);
}

cudaError_t launch_clear_bits(const LaunchCtx& c, void* data, uint64_t n_voxels, const uint32_t mask8[8], uint32_t* occ) {
  BitVoxelWords mk;
  for (int k = 0; k < 8; ++k) mk.w[k] = mask8[k];
  ++*c.launches;
  k_clear_bits<<<grid_for(c, n_voxels, 8), kInsThreads, 0, c.stream>>>((uint8_t*)data, n_voxels, mk, occ);
  return cudaGetLastError();
}

}  // namespace gvx
