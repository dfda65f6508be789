// gvx_internal.h -- host-side launcher prototypes shared between the kernel files and gvx_api.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "gvx_device.cuh"

namespace gvx {

struct LaunchCtx {
  cudaStream_t stream;
  int sm_count;
  ReduceScratch* scratch_d;   // device
  ResultBlock* result_d;      // device copy of the result block
  ResultBlock* result_hmap_d; // device alias of the pinned, mapped host result block
  uint64_t* launches;         // kernel launch counter (host)
};

// collideWith map x map. a/b: device base pointers of the voxel arrays; a is indexed at i + off.
// [i_begin, i_end) is the range of i for which 0 <= i + off < n. count_out_d: optional extra device
// destination for the 64-bit count (async API), may be nullptr.
cudaError_t launch_collide_map_map(const LaunchCtx& c, void* a, int a_type, const void* b, int b_type,
                                   int64_t i_begin, int64_t i_end, int64_t off, int8_t thr, bool mark,
                                   bool types, unsigned long long* count_out_d);

// voxelisation
struct MetaDesc {             // device-resident description of a MetaPointCloud insert
  const uint32_t* prefix_d;   // n_clouds + 1 exclusive prefix of cloud sizes
  const int32_t* meanings_d;  // n_meanings
  const float* transforms_d;  // n_clouds * 16 or nullptr
  uint32_t n_clouds;
  uint32_t n_meanings;
};
struct MapView {
  void* data;
  int type;
  uint32_t dx, dy, dz, z0, z1;
  float side;
};
cudaError_t launch_insert_points(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n, int meaning,
                                 bool want_flag);
cudaError_t launch_insert_meta(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                               const MetaDesc& d, bool want_flag);
struct Affine12 { float a[12]; };  // row-major [R|t], passed to kernels by value (no H2D copy, no sync)
cudaError_t launch_transform(const LaunchCtx& c, const float* in_d, float* out_d, uint64_t n, const Affine12& M);
cudaError_t launch_update_logodds(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                  const Affine12& pose, int delta);
cudaError_t launch_clear_bits(const LaunchCtx& c, void* data, uint64_t n_voxels, const uint32_t mask8[8]);

}  // namespace gvx
