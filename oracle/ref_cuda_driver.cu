// oracle/ref_cuda_driver.cu -- TEST INFRASTRUCTURE, not product code.
//
// Launches the REFERENCE's own CUDA kernels, included unmodified by include path from
// /root/reference (never copied into this repo), with the reference's launch shapes
// (computeLinearLoad: <=65535 blocks x 1024 threads, GV/helpers/MathHelpers.cpp:29-77) and the
// reference's result path (D2H copy of per-block uint16 counters + host sum,
// GV/voxelmap/TemplateVoxelMap.hpp:486-519). Built only where /root/reference exists, into
// oracle/_ref/libgvref_cuda.so (git-ignored, travels to the GPU box). Used by tests/ as the
// on-GPU oracle for float->index edge cases and by bench.py as the "ref_cuda" reported baseline.
//
// All pointers are DEVICE pointers unless a name ends in _h.
#include <cstdint>
#include <cassert>
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>
#include <gpu_voxels/voxel/BitVoxel.hpp>
#include <gpu_voxels/voxel/ProbabilisticVoxel.hpp>
#include <gpu_voxels/voxel/DefaultCollider.hpp>
#include <gpu_voxels/voxel/SVCollider.hpp>
#include <gpu_voxels/voxelmap/kernels/VoxelMapOperations.hpp>
#include <gpu_voxels/voxellist/kernels/VoxelListOperations.hpp>

using namespace gpu_voxels;

// kernelTransformCloud lives in a .cu of the reference (GV/helpers/kernels/MetaPointCloudOperations.cu:77-91);
// include the translation unit itself so the kernel body is the reference's.
#include <gpu_voxels/helpers/kernels/MetaPointCloudOperations.cu>
// kernelTransformSensorData: GV/voxelmap/kernels/VoxelMapOperations.cu:27-71
#include <gpu_voxels/voxelmap/kernels/VoxelMapOperations.cu>

static void linear_load(uint32_t n, uint32_t* blocks, uint32_t* threads) {
  // restates GV/helpers/MathHelpers.cpp:29-77 (host-only launch-shape arithmetic)
  if (n == 0) { *blocks = 1; *threads = 1; return; }
  if (n <= cMAX_NR_OF_BLOCKS * cMAX_THREADS_PER_BLOCK) {
    *blocks = (n + cMAX_THREADS_PER_BLOCK - 1) / cMAX_THREADS_PER_BLOCK; *threads = cMAX_THREADS_PER_BLOCK;
  } else { *blocks = cMAX_NR_OF_BLOCKS; *threads = cMAX_THREADS_PER_BLOCK; }
}

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "ref_cuda: %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); return -1; } } while (0)

static uint16_t* g_results_d = nullptr;
static uint16_t* g_results_h = nullptr;
static bool* g_outside_d = nullptr;
static int ensure_scratch() {
  if (!g_results_d) {
    CK(cudaMalloc(&g_results_d, cMAX_NR_OF_BLOCKS * sizeof(uint16_t)));
    g_results_h = (uint16_t*)malloc(cMAX_NR_OF_BLOCKS * sizeof(uint16_t));
    CK(cudaMalloc(&g_outside_d, sizeof(bool)));
  }
  return 0;
}

extern "C" {

// GV/voxelmap/TemplateVoxelMap.hpp:587-607
int gvref_insert_points(void* map, int is_prob, uint32_t dx, uint32_t dy, uint32_t dz, float side,
                        const float* xyz, uint32_t n, int meaning, int* any_outside_h) {
  if (ensure_scratch()) return -1;
  CK(cudaMemset(g_outside_d, 0, sizeof(bool)));
  uint32_t b, t; linear_load(n, &b, &t);
  CK(cudaDeviceSynchronize());
  if (is_prob)
    voxelmap::kernelInsertGlobalPointCloud<<<b, t>>>((ProbabilisticVoxel*)map, Vector3ui(dx, dy, dz), side,
        (const Vector3f*)xyz, (std::size_t)n, (BitVoxelMeaning)meaning, g_outside_d);
  else
    voxelmap::kernelInsertGlobalPointCloud<<<b, t>>>((BitVectorVoxel*)map, Vector3ui(dx, dy, dz), side,
        (const Vector3f*)xyz, (std::size_t)n, (BitVoxelMeaning)meaning, g_outside_d);
  CK(cudaGetLastError());
  bool o = false;
  CK(cudaMemcpy(&o, g_outside_d, sizeof(bool), cudaMemcpyDeviceToHost));
  CK(cudaDeviceSynchronize());
  if (any_outside_h) *any_outside_h = o ? 1 : 0;
  return 0;
}

// GV/helpers/MetaPointCloud.cu:624-649 -> kernelTransformCloud
int gvref_transform_cloud(const float* M16_h, const float* in_xyz, float* out_xyz, uint32_t n) {
  Matrix4f* m_d;
  CK(cudaMalloc(&m_d, sizeof(Matrix4f)));
  CK(cudaMemcpy(m_d, M16_h, sizeof(Matrix4f), cudaMemcpyHostToDevice));
  uint32_t b, t; linear_load(n, &b, &t);
  CK(cudaDeviceSynchronize());
  kernelTransformCloud<<<b, t>>>(m_d, (const Vector3f*)in_xyz, (Vector3f*)out_xyz, n);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaFree(m_d));
  return 0;
}

} // extern "C"
