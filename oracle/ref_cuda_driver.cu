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
// VoxelListOperations.hpp:137 calls getHighestVoxelPtr unqualified from namespace voxellist (accepted by the CUDA 7.5
// era front end); make the name visible there without touching the reference source.
namespace gpu_voxels { namespace voxellist { using voxelmap::getHighestVoxelPtr; } }
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


static uint32_t sum_block_counters(uint32_t blocks) {
  // host tail of collisionCheckWithCounterRelativeTransform (GV/voxelmap/TemplateVoxelMap.hpp:507-518)
  cudaMemcpy(g_results_h, g_results_d, cMAX_NR_OF_BLOCKS * sizeof(uint16_t), cudaMemcpyDeviceToHost);
  uint32_t n = 0;
  for (uint32_t i = 0; i < blocks; i++) n += g_results_h[i];
  return n;
}

// collisionCheckWithCounterRelativeTransform, offset 0 (GV/voxelmap/TemplateVoxelMap.hpp:486-519):
// kernelCollideVoxelMapsDebug + 128 KiB D2H + host sum. a/b types: 0 = BitVectorVoxel, 4 = ProbabilisticVoxel.
// Marks a (the kernel always does). Returns the count or -1.
long long gvref_collide_map_map(void* a, int a_type, void* b, int b_type, uint32_t n, float thr) {
  if (ensure_scratch()) return -1;
  uint32_t bl, t; linear_load(n, &bl, &t);
  DefaultCollider col(thr);
  CK(cudaDeviceSynchronize());
  if (a_type == 0 && b_type == 4)
    voxelmap::kernelCollideVoxelMapsDebug<<<bl, t>>>((BitVectorVoxel*)a, n, (ProbabilisticVoxel*)b, col, g_results_d);
  else if (a_type == 4 && b_type == 0)
    voxelmap::kernelCollideVoxelMapsDebug<<<bl, t>>>((ProbabilisticVoxel*)a, n, (BitVectorVoxel*)b, col, g_results_d);
  else if (a_type == 4 && b_type == 4)
    voxelmap::kernelCollideVoxelMapsDebug<<<bl, t>>>((ProbabilisticVoxel*)a, n, (ProbabilisticVoxel*)b, col, g_results_d);
  else
    voxelmap::kernelCollideVoxelMapsDebug<<<bl, t>>>((BitVectorVoxel*)a, n, (BitVectorVoxel*)b, col, g_results_d);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  return (long long)sum_block_counters(bl);
}

// insertMetaPointCloud with per-cloud meanings (GV/voxelmap/TemplateVoxelMap.hpp:631-664): builds the
// MetaPointCloudStruct the way MetaPointCloud.cu does (struct + cloud_sizes + base addresses in device memory) and
// uploads the meanings array per call like the reference.
int gvref_insert_meta(void* map, int is_prob, uint32_t dx, uint32_t dy, uint32_t dz, float side, const float* xyz,
                      const uint32_t* cloud_sizes_h, uint32_t n_clouds, const int32_t* meanings_h, int* any_outside_h) {
  if (ensure_scratch()) return -1;
  std::vector<Vector3f*> bases(n_clouds);
  uint32_t acc = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) { bases[c] = (Vector3f*)xyz + acc; acc += cloud_sizes_h[c]; }
  MetaPointCloudStruct st;
  st.num_clouds = (uint16_t)n_clouds;
  st.accumulated_cloud_size = acc;
  MetaPointCloudStruct* st_d; BitVoxelMeaning* mg_d;
  CK(cudaMalloc(&st.cloud_sizes, n_clouds * sizeof(uint32_t)));
  CK(cudaMalloc(&st.clouds_base_addresses, n_clouds * sizeof(Vector3f*)));
  CK(cudaMalloc(&st_d, sizeof(MetaPointCloudStruct)));
  CK(cudaMalloc(&mg_d, n_clouds * sizeof(BitVoxelMeaning)));
  CK(cudaMemcpy(st.cloud_sizes, cloud_sizes_h, n_clouds * sizeof(uint32_t), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(st.clouds_base_addresses, bases.data(), n_clouds * sizeof(Vector3f*), cudaMemcpyHostToDevice));
  CK(cudaMemcpy(st_d, &st, sizeof(MetaPointCloudStruct), cudaMemcpyHostToDevice));
  std::vector<BitVoxelMeaning> mg(n_clouds);
  for (uint32_t c = 0; c < n_clouds; ++c) mg[c] = (BitVoxelMeaning)meanings_h[c];
  CK(cudaMemcpy(mg_d, mg.data(), n_clouds * sizeof(BitVoxelMeaning), cudaMemcpyHostToDevice));
  CK(cudaMemset(g_outside_d, 0, sizeof(bool)));
  uint32_t b, t; linear_load(acc, &b, &t);
  CK(cudaDeviceSynchronize());
  if (is_prob)
    voxelmap::kernelInsertMetaPointCloud<<<b, t>>>((ProbabilisticVoxel*)map, st_d, mg_d, Vector3ui(dx, dy, dz), side, g_outside_d);
  else
    voxelmap::kernelInsertMetaPointCloud<<<b, t>>>((BitVectorVoxel*)map, st_d, mg_d, Vector3ui(dx, dy, dz), side, g_outside_d);
  CK(cudaGetLastError());
  bool o = false;
  CK(cudaMemcpy(&o, g_outside_d, sizeof(bool), cudaMemcpyDeviceToHost));
  CK(cudaDeviceSynchronize());
  if (any_outside_h) *any_outside_h = o ? 1 : 0;
  cudaFree(st.cloud_sizes); cudaFree(st.clouds_base_addresses); cudaFree(st_d); cudaFree(mg_d);
  return 0;
}

// ProbVoxelMap::insertSensorData without ray casting (GV/voxelmap/ProbVoxelMap.hpp:52-102): kernelTransformSensorData
// + kernelInsertSensorData<DummyRayCaster>. pose12_h = row-major [R|t]; raw_d: n sensor-frame points (device);
// tmp_d: n points of scratch (device).
int gvref_insert_sensor_data(void* map, uint32_t dx, uint32_t dy, uint32_t dz, float side, const float* pose12_h,
                             float* raw_d, float* tmp_d, uint32_t n) {
  Sensor s(Vector3f(pose12_h[3], pose12_h[7], pose12_h[11]),
           Matrix3f(pose12_h[0], pose12_h[1], pose12_h[2], pose12_h[4], pose12_h[5], pose12_h[6], pose12_h[8],
                    pose12_h[9], pose12_h[10]), n, 1);
  Sensor* s_d;
  CK(cudaMalloc(&s_d, sizeof(Sensor)));
  CK(cudaMemcpy(s_d, &s, sizeof(Sensor), cudaMemcpyHostToDevice));
  uint32_t b, t; linear_load(n, &b, &t);
  voxelmap::kernelTransformSensorData<<<b, t>>>(s_d, (Vector3f*)raw_d, (Vector3f*)tmp_d);
  CK(cudaGetLastError());
  voxelmap::kernelInsertSensorData<<<b, t>>>((ProbabilisticVoxel*)map, dx * dy * dz, Vector3ui(dx, dy, dz), side, s_d,
                                            (const Vector3f*)tmp_d, false, (BitVoxel<BIT_VECTOR_LENGTH>*)NULL,
                                            (uint32_t)eBVM_OCCUPIED, voxelmap::DummyRayCaster());
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaFree(s_d));
  return 0;
}

// ProbVoxelMap::insertSensorData complete (GV/voxelmap/ProbVoxelMap.hpp:52-102): raycast != 0 selects RayCaster,
// robot_map_d != NULL enables cut_real_robot with bit `robot_bit`. serial != 0 launches <<<1, 1>>>: the reference's
// kernel run by ONE thread walks the points in index order without races (its updates are non-atomic), which is the
// deterministic result the parity tests compare against.
int gvref_insert_sensor_data_full(void* map, uint32_t dx, uint32_t dy, uint32_t dz, float side, const float* pose12_h,
                                  float* raw_d, float* tmp_d, uint32_t n, int raycast, void* robot_map_d,
                                  uint32_t robot_bit, int serial) {
  Sensor s(Vector3f(pose12_h[3], pose12_h[7], pose12_h[11]),
           Matrix3f(pose12_h[0], pose12_h[1], pose12_h[2], pose12_h[4], pose12_h[5], pose12_h[6], pose12_h[8],
                    pose12_h[9], pose12_h[10]), n, 1);
  Sensor* s_d;
  CK(cudaMalloc(&s_d, sizeof(Sensor)));
  CK(cudaMemcpy(s_d, &s, sizeof(Sensor), cudaMemcpyHostToDevice));
  uint32_t b, t; linear_load(n, &b, &t);
  voxelmap::kernelTransformSensorData<<<b, t>>>(s_d, (Vector3f*)raw_d, (Vector3f*)tmp_d);
  CK(cudaGetLastError());
  if (serial) { b = 1; t = 1; }
  const bool cut = robot_map_d != NULL;
  if (raycast)
    voxelmap::kernelInsertSensorData<<<b, t>>>((ProbabilisticVoxel*)map, dx * dy * dz, Vector3ui(dx, dy, dz), side, s_d,
                                              (const Vector3f*)tmp_d, cut, (BitVoxel<BIT_VECTOR_LENGTH>*)robot_map_d,
                                              robot_bit, voxelmap::RayCaster());
  else
    voxelmap::kernelInsertSensorData<<<b, t>>>((ProbabilisticVoxel*)map, dx * dy * dz, Vector3ui(dx, dy, dz), side, s_d,
                                              (const Vector3f*)tmp_d, cut, (BitVoxel<BIT_VECTOR_LENGTH>*)robot_map_d,
                                              robot_bit, voxelmap::DummyRayCaster());
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  CK(cudaFree(s_d));
  return 0;
}

// BitVoxelList::collideWith(map) (GV/voxellist/TemplateVoxelList.hpp:731-774): kernelCollideWithVoxelMap + D2H + host sum
long long gvref_collide_list_map(const uint32_t* ids, void* voxels, uint32_t size, const void* map, int map_type,
                                 uint32_t dx, uint32_t dy, uint32_t dz, float thr, int ox, int oy, int oz) {
  if (ensure_scratch()) return -1;
  uint32_t bl, t; linear_load(size, &bl, &t);
  DefaultCollider col(thr);
  CK(cudaDeviceSynchronize());
  if (map_type == 4)
    voxellist::kernelCollideWithVoxelMap<<<bl, t>>>((const MapVoxelID*)ids, (BitVectorVoxel*)voxels, size,
        (const ProbabilisticVoxel*)map, Vector3ui(dx, dy, dz), col, Vector3i(ox, oy, oz), g_results_d);
  else
    voxellist::kernelCollideWithVoxelMap<<<bl, t>>>((const MapVoxelID*)ids, (BitVectorVoxel*)voxels, size,
        (const BitVectorVoxel*)map, Vector3ui(dx, dy, dz), col, Vector3i(ox, oy, oz), g_results_d);
  CK(cudaGetLastError());
  CK(cudaDeviceSynchronize());
  return (long long)sum_block_counters(bl);
}

// clearMap (GV/voxelmap/TemplateVoxelMap.hpp:205-246)
int gvref_clear_map(void* map, int is_prob, uint64_t n) {
  CK(cudaMemset(map, is_prob ? 0x80 : 0x00, n * (is_prob ? 1 : 32)));
  return 0;
}

} // extern "C"
