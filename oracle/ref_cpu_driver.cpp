// oracle/ref_cpu_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The reference has no CPU path, but its voxel value types and colliders are __host__ __device__
// (GV/voxel/BitVoxel.hpp, ProbabilisticVoxel.hpp, DefaultCollider.hpp, SVCollider.hpp,
// GV/helpers/BitVector.h, GV/voxelmap/kernels/VoxelMapOperations.h). This file includes those
// headers UNMODIFIED by include path from /root/reference and calls them from OpenMP host loops
// shaped like the reference kernels, giving (a) an independent check of oracle/gvoracle.c that can run
// in the CPU-only container and (b) the host-core baseline `bench.py --impl reference` times.
// Built only where /root/reference exists, into oracle/_ref/libgvref_cpu.so.
//
// NOT used for float->voxel mapping of negative/NaN input: a host compile of mapToVoxels is not the
// device semantics there (see oracle/gvoracle.c header).
#define THRUST_DEVICE_SYSTEM THRUST_DEVICE_SYSTEM_CPP
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <cassert>
#include <cmath>
#include <omp.h>
#include <cuda_runtime.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>
#include <gpu_voxels/voxel/BitVoxel.hpp>
#include <gpu_voxels/voxel/ProbabilisticVoxel.hpp>
#include <gpu_voxels/voxel/DefaultCollider.hpp>
#include <gpu_voxels/voxel/SVCollider.hpp>
#include <gpu_voxels/voxelmap/kernels/VoxelMapOperations.h>

// DHParameters::convertDHtoM: the reference's own translation unit, compiled for the host as it is there (nvcc hands
// it to the same g++); no FMA contraction (-ffp-contract=off and no -mfma, like the reference's x86-64 build)
#include <gpu_voxels/robot/dh_robot/KinematicLink.cu>

using namespace gpu_voxels;
static_assert(sizeof(BitVectorVoxel) == 32, "BitVectorVoxel layout");
static_assert(sizeof(ProbabilisticVoxel) == 1, "ProbabilisticVoxel layout");

// floatToProbability is protected in the reference; a derived class exposes it unchanged.
struct ColliderProbe : public DefaultCollider {
  static Probability f2p(float v) { return DefaultCollider::floatToProbability(v); }
};

template <class VA, class VB>
static uint64_t collide_loop(VA* a, const VB* b, uint64_t n, DefaultCollider collider, int64_t off, int mark) {
  uint64_t cnt = 0;
  // body of kernelCollideVoxelMapsDebug, GV/voxelmap/kernels/VoxelMapOperations.hpp:141-156
#pragma omp parallel for schedule(static) reduction(+ : cnt)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    int64_t j = i + off;
    if (j < 0 || j >= (int64_t)n) continue;
    if (collider.collide(a[j], b[i])) {
      if (mark) a[j].insert(eBVM_COLLISION);
      ++cnt;
    }
  }
  return cnt;
}

extern "C" {

int gvrefcpu_num_threads() { return omp_get_max_threads(); }
void gvrefcpu_set_num_threads(int n) { if (n > 0) omp_set_num_threads(n); }

// DefaultCollider(float) ctor -> floatToProbability, GV/voxel/DefaultCollider.hpp:39-43,94-98
int gvrefcpu_float_to_probability(float thr) { return (int)ColliderProbe::f2p(thr); }

// GV/voxelmap/kernels/VoxelMapOperations.h:61-74
int64_t gvrefcpu_voxel_index_signed(const uint32_t* dims, const int32_t* off) {
  return (int64_t)voxelmap::getVoxelIndexSigned(Vector3ui(dims[0], dims[1], dims[2]), Vector3i(off[0], off[1], off[2]));
}

// host evaluation of mapToVoxels (valid for non-negative finite coordinates only)
void gvrefcpu_map_to_voxels(float side, const float* xyz, uint32_t* out3) {
  Vector3ui c = voxelmap::mapToVoxels(side, Vector3f(xyz[0], xyz[1], xyz[2]));
  out3[0] = c.x; out3[1] = c.y; out3[2] = c.z;
}

// Matrix4f * Vector3f on the host (g++ -ffp-contract=off: NO fma contraction -> differs from the device
// in the last bit; exported so the tests can show exactly that)
void gvrefcpu_transform_point_host(const float* M16, const float* p, float* o) {
  Matrix4f m;
  memcpy(&m, M16, sizeof(Matrix4f));
  Vector3f r = m * Vector3f(p[0], p[1], p[2]);
  o[0] = r.x; o[1] = r.y; o[2] = r.z;
}

// voxel.insert(meaning) on a given voxel index: BitVoxel.hpp:60-65 / ProbabilisticVoxel.hpp:77-92
void gvrefcpu_insert_at(void* map, int is_prob, uint64_t idx, int meaning) {
  if (is_prob) ((ProbabilisticVoxel*)map)[idx].insert((BitVoxelMeaning)meaning);
  else ((BitVectorVoxel*)map)[idx].insert((BitVoxelMeaning)meaning);
}
int gvrefcpu_update_occupancy(int occ, int delta) {
  ProbabilisticVoxel v; v.occupancy() = (Probability)occ;
  return (int)v.updateOccupancy((Probability)delta);
}

// a_type/b_type: 0 = BitVectorVoxel, 4 = ProbabilisticVoxel (MapType enum values)
static uint64_t collide_any(void* a, int a_type, const void* b, int b_type, uint64_t n, DefaultCollider c, int64_t off, int mark) {
  if (a_type == 0 && b_type == 0) return collide_loop((BitVectorVoxel*)a, (const BitVectorVoxel*)b, n, c, off, mark);
  if (a_type == 0 && b_type == 4) return collide_loop((BitVectorVoxel*)a, (const ProbabilisticVoxel*)b, n, c, off, mark);
  if (a_type == 4 && b_type == 0) return collide_loop((ProbabilisticVoxel*)a, (const BitVectorVoxel*)b, n, c, off, mark);
  return collide_loop((ProbabilisticVoxel*)a, (const ProbabilisticVoxel*)b, n, c, off, mark);
}
uint64_t gvrefcpu_collide_map_map(void* a, int a_type, const void* b, int b_type, uint64_t n, float thr,
                                  int64_t off, int mark) {
  return collide_any(a, a_type, b, b_type, n, DefaultCollider(thr), off, mark);
}
// DefaultCollider(Probability threshold1, Probability threshold2), GV/voxel/DefaultCollider.hpp:45-49
uint64_t gvrefcpu_collide_map_map_thresholds(void* a, int a_type, const void* b, int b_type, uint64_t n, int t1, int t2,
                                             int64_t off, int mark) {
  return collide_any(a, a_type, b, b_type, n, DefaultCollider((Probability)t1, (Probability)t2), off, mark);
}

// body of kernelCollideVoxelMapsBitvector with SVCollider(thr) Bit x Prob, VoxelMapOperations.hpp:202-214
uint64_t gvrefcpu_collide_bit_prob_types(void* a_, const void* b_, uint64_t n, float thr, int mark, uint8_t* mask32) {
  BitVectorVoxel* a = (BitVectorVoxel*)a_;
  const ProbabilisticVoxel* b = (const ProbabilisticVoxel*)b_;
  SVCollider collider(thr);
  BitVector<BIT_VECTOR_LENGTH> acc;
  memcpy(&acc, mask32, 32);
  uint64_t cnt = 0;
  for (uint64_t i = 0; i < n; ++i) {
    BitVector<BIT_VECTOR_LENGTH> temp;
    if (collider.collide(a[i], b[i], &temp, 0)) {
      if (mark) a[i].insert(eBVM_COLLISION);
      acc |= temp;
      ++cnt;
    }
  }
  memcpy(mask32, &acc, 32);
  return cnt;
}

// SVCollider Bit x Bit windowed check on two single voxels: bitMarginCollisionCheck, GV/helpers/BitVector.h:415-471
int gvrefcpu_bit_margin_check(const uint8_t* v1, const uint8_t* v2, uint8_t* collisions32, int margin, uint32_t sv_offset) {
  BitVector<BIT_VECTOR_LENGTH> a, b, c;
  memcpy(&a, v1, 32); memcpy(&b, v2, 32); memcpy(&c, collisions32, 32);
  bool r = bitMarginCollisionCheck<BIT_VECTOR_LENGTH>(a, b, &c, (uint8_t)margin, sv_offset);
  memcpy(collisions32, &c, 32);
  return r ? 1 : 0;
}
// performLeftShift, GV/helpers/BitVector.h:361-402
void gvrefcpu_shift_left(uint8_t* v32, int shift) {
  BitVector<BIT_VECTOR_LENGTH> a;
  memcpy(&a, v32, 32);
  performLeftShift(a, (uint8_t)shift);
  memcpy(v32, &a, 32);
}

// list-entry predicates of the list x map kernels, GV/voxellist/kernels/VoxelListOperations.hpp:139-150,207-216,282-292
int gvrefcpu_list_entry_hit(int mode, const uint8_t* list_voxel32, const void* map_voxel, int map_type, float thr,
                            const uint8_t* mask32) {
  BitVectorVoxel lv; memcpy(&lv, list_voxel32, 32);
  if (mode == 0) {
    DefaultCollider collider(thr);
    if (map_type == 4) return collider.collide(lv, *(const ProbabilisticVoxel*)map_voxel);
    return collider.collide(lv, *(const BitVectorVoxel*)map_voxel);
  }
  bool occ = map_type == 4 ? ((const ProbabilisticVoxel*)map_voxel)->isOccupied(thr)
                           : ((const BitVectorVoxel*)map_voxel)->isOccupied(thr);
  if (mode == 1) return occ;
  BitVectorVoxel m; memcpy(&m, mask32, 32);
  return occ && !(m.bitVector() & lv.bitVector()).isZero();
}

// KinematicChain::setConfiguration's loop, GV/robot/dh_robot/KinematicChain.cu:104-124, with the reference's
// DHParameters::convertDHtoM and Matrix4f operator*; out[cfg][link] = the matrix handed to transformSubCloud
void gvrefcpu_dh_chain(const float* dh4, const int32_t* joint_types, uint32_t n_links, const float* joint_values,
                       uint32_t n_cfg, float* out) {
  for (uint32_t c = 0; c < n_cfg; ++c) {
    Matrix4f transformation = gpu_voxels::Matrix4f::createIdentity();
    for (uint32_t l = 0; l < n_links; ++l) {
      memcpy(out + ((size_t)c * n_links + l) * 16, &transformation, 64);
      robot::KinematicLink link(robot::DHParameters(dh4[4 * l], dh4[4 * l + 1], dh4[4 * l + 2], dh4[4 * l + 3], 0.0f,
                                                    (robot::DHJointType)joint_types[l]));
      link.setJointValue(joint_values[(size_t)c * n_links + l]);
      Matrix4f m;
      link.getMatrixRepresentation(m);
      transformation = transformation * m;
    }
  }
}

} // extern "C"
