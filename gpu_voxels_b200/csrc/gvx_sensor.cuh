// gvx_sensor.cuh -- device functions shared by the sensor-model kernels (gvx_insert.cu, gvx_dispatch.cu).
#pragma once
#include "gvx_device.cuh"

namespace gvx {

constexpr int kInsThreads = 256;

// Saturating int8 add on one byte of the map with a 32-bit CAS. For same-sign deltas,
// clamp(occ + k*delta, -127, 127) equals k sequential ProbabilisticVoxel::updateOccupancy calls
// (GV/voxel/ProbabilisticVoxel.hpp:51-57), so a warp's duplicates are folded into one update.
// Returns true when the voxel was UNKNOWN (-128) at the first read: only then can its bit be missing from a valid
// "known" summary (bit == (occupancy != UNKNOWN), kept exact by every mutating kernel), so callers touch the summary
// word -- a second random DRAM sector per point -- only on the first touch of a voxel. clamp(.., -127, 127) never
// returns to -128, so an UNKNOWN voxel always changes and a known one stays known.
__device__ __forceinline__ bool logodds_add(int8_t* map, uint64_t idx, int total_delta) {
  uint32_t* wp = reinterpret_cast<uint32_t*>(map + (idx & ~3ull));
  const int shift = (int)(idx & 3ull) * 8;
  uint32_t old = __ldcg(wp), assumed;
  const bool was_unknown = ((old >> shift) & 0xFFu) == 0x80u;
  do {
    assumed = old;
    int v = (int)(int8_t)((assumed >> shift) & 0xFFu) + total_delta;
    v = v < -127 ? -127 : (v > 127 ? 127 : v);
    const uint32_t nw = (assumed & ~(0xFFu << shift)) | (((uint32_t)v & 0xFFu) << shift);
    if (nw == assumed) break;
    old = atomicCAS(wp, assumed, nw);
  } while (old != assumed);
  return was_unknown;
}

// kernelTransformSensorData, GV/voxelmap/kernels/VoxelMapOperations.cu:27-71: pose * raw point; a point that lands below
// z = 0 is pulled back along its ray onto the floor with a DOUBLE factor
__device__ __forceinline__ void sensor_to_world(const float* __restrict__ P, float rx, float ry, float rz, float& wx,
                                                float& wy, float& wz) {
  wx = affine_row(P, rx, ry, rz); wy = affine_row(P + 4, rx, ry, rz); wz = affine_row(P + 8, rx, ry, rz);
  if (wz < 0.0f) {
    const double factor = (double)__fdiv_rn(P[11], __fsub_rn(P[11], wz));
    const float qx = (float)((double)rx * factor), qy = (float)((double)ry * factor), qz = (float)((double)rz * factor);
    wx = affine_row(P, qx, qy, qz); wy = affine_row(P + 4, qx, qy, qz); wz = affine_row(P + 8, qx, qy, qz);
  }
}

}  // namespace gvx
