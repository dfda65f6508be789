/* oracle/gvoracle.c -- CPU ORACLE. TEST INFRASTRUCTURE ONLY, NOT PRODUCT CODE.
 *
 * A plain-C restatement of the reference's algorithm for the hot path (point-cloud
 * voxelisation and collision counting of gpu-voxels). Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load this library; the product
 * (gpu_voxels_b200/) never links, imports or calls it.
 *
 * Every function cites the reference file:line it follows; paths are relative to
 * /root/reference/packages/gpu_voxels/src/gpu_voxels/ ("GV/").
 *
 * Parity pinning: this oracle is checked (tests/test_oracle_golden.py) against every
 * known-answer test the reference's own Boost suites hold for the path
 * (GV/test/testing_voxelmap.cu, testing_voxellist.cu, testing_bitvector.cu) and, where
 * /root/reference is mounted, against oracle/_ref/libgvref_cpu.so, which is the reference's own
 * __host__ __device__ voxel/collider headers compiled with g++ (oracle/ref_cpu_driver.cpp).
 *
 * Semantics that differ from a naive host compile of the reference and are restated on purpose:
 *  - float -> voxel coordinate uses DEVICE conversion rules (SASS F2I.U32.FLOOR.NTZ): NaN and
 *    negatives give 0, values >= 2^32 give 0xFFFFFFFF (GV/voxelmap/kernels/VoxelMapOperations.h:123-133).
 *  - the affine transform uses the FMA contraction nvcc emits for GV/helpers/cuda_matrices.h:740-746:
 *    fadd(a14, fma(a13,z, fma(a11,x, fmul(a12,y)))) (checked in SASS of oracle/_ref/libgvref_cuda.so).
 *  - inserts are race-free (sequential semantics); indices are 64-bit.
 *
 * Build: make -C oracle   (gcc -O2 -fopenmp -ffp-contract=off; no -ffast-math)
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

#define GVO_BITVOXEL_BYTES 32 /* BIT_VECTOR_LENGTH=256, GV/helpers/common_defines.h:46 */
#define GVO_EBVM_FREE 0
#define GVO_EBVM_OCCUPIED 1
#define GVO_EBVM_COLLISION 2
#define GVO_EBVM_UNKNOWN 3
#define GVO_EBVM_UNDEFINED 255
#define GVO_MAP_BIT 0  /* MT_BITVECTOR_VOXELMAP */
#define GVO_MAP_PROB 4 /* MT_PROBAB_VOXELMAP, enum order GV/helpers/common_defines.h:59-73 */

int gvo_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void gvo_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ---- float -> voxel coordinate: GV/voxelmap/kernels/VoxelMapOperations.h:123-133 -------------
 * static_cast<uint32_t>(floor(c / side)) evaluated with DEVICE rules: IEEE div.rn.f32, floor,
 * then a saturating convert where NaN -> 0. */
static inline uint32_t f2u_floor_sat(float q) {
  float f = floorf(q);
  if (!(f > 0.0f)) return 0u; /* NaN, negatives, -0, +0 */
  if (f >= 4294967296.0f) return 0xFFFFFFFFu;
  return (uint32_t)f;
}
void gvo_map_to_voxels(float side, const float* xyz, uint32_t* out3) {
  out3[0] = f2u_floor_sat(xyz[0] / side);
  out3[1] = f2u_floor_sat(xyz[1] / side);
  out3[2] = f2u_floor_sat(xyz[2] / side);
}

/* GV/voxelmap/kernels/VoxelMapOperations.h:61-74 (int32 arithmetic in the reference; 64-bit here,
 * identical wherever the reference does not overflow) */
int64_t gvo_voxel_index_signed(const uint32_t* dims, const int32_t* off) {
  return (int64_t)off[2] * (int64_t)dims[0] * (int64_t)dims[1] + (int64_t)off[1] * (int64_t)dims[0] + (int64_t)off[0];
}

/* ---- affine transform: GV/helpers/cuda_matrices.h:740-746, contraction as in the SASS of
 * kernelTransformCloud (GV/helpers/kernels/MetaPointCloudOperations.cu:77-91) */
static inline void xform(const float* M, const float* p, float* o) {
  const float x = p[0], y = p[1], z = p[2];
  for (int r = 0; r < 3; ++r) {
    const float* a = M + 4 * r;
    float t = a[1] * y;
    t = fmaf(a[0], x, t);
    t = fmaf(a[2], z, t);
    o[r] = a[3] + t;
  }
}
void gvo_transform_points(const float* M16, const float* in_xyz, float* out_xyz, uint64_t n) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    float o[3];
    xform(M16, in_xyz + 3 * i, o);
    out_xyz[3 * i] = o[0]; out_xyz[3 * i + 1] = o[1]; out_xyz[3 * i + 2] = o[2];
  }
}

/* ---- voxel value types --------------------------------------------------------------------- */
/* BitVector::setBit, GV/helpers/BitVector.h:223-228 */
static inline void bit_set(uint8_t* v, uint32_t m) { v[m >> 3] |= (uint8_t)(1u << (m & 7)); }
/* BitVector::noneButEmpty, GV/helpers/BitVector.h:183-198 */
static inline int bit_none_but_empty(const uint8_t* v) {
  if (v[0] & 0xFE) return 0;
  for (int i = 1; i < GVO_BITVOXEL_BYTES; ++i) if (v[i]) return 0;
  return 1;
}
/* ProbabilisticVoxel::insert, GV/voxel/ProbabilisticVoxel.hpp:77-92 */
static inline int8_t prob_insert_value(int meaning) {
  switch (meaning) {
    case GVO_EBVM_FREE: return -127;
    case GVO_EBVM_COLLISION:
    case GVO_EBVM_OCCUPIED: return 127;
    default: return -128;
  }
}
/* ProbabilisticVoxel::updateOccupancy, GV/voxel/ProbabilisticVoxel.hpp:51-57 */
static inline int8_t prob_update(int8_t occ, int delta) {
  int v = (int)occ + delta;
  if (v < -127) v = -127;
  if (v > 127) v = 127;
  return (int8_t)v;
}
/* DefaultCollider::floatToProbability, GV/voxel/DefaultCollider.hpp:94-98 (C truncation toward 0) */
int gvo_float_to_probability(float thr) {
  float tmp = (thr * ((float)127 - (float)(-127))) + (float)(-127);
  /* host conversion in the reference (g++/x86): float -> int32 truncation, then low 8 bits */
  return (int)(int8_t)(int32_t)tmp;
}

/* ---- insertPointCloud / insertMetaPointCloud ------------------------------------------------
 * kernelInsertGlobalPointCloud  GV/voxelmap/kernels/VoxelMapOperations.hpp:241-264
 * kernelInsertMetaPointCloud    GV/voxelmap/kernels/VoxelMapOperations.hpp:302-340 (one meaning)
 *                               GV/voxelmap/kernels/VoxelMapOperations.hpp:386-437 (per-cloud meanings)
 * Sequential, hence race-free. The map may be a z-slab [z0,z1) of the full grid (multi-GPU
 * sharding, not in the reference): `map` then holds voxels z0*dx*dy .. z1*dx*dy only, and
 * any_outside still means "outside the WHOLE grid". z0=0, z1=dz is the reference case. */
static inline int insert_one(uint8_t* map, int map_type, const uint32_t* dims, uint32_t z0, uint32_t z1,
                             float side, const float* p, int meaning) {
  uint32_t c[3];
  gvo_map_to_voxels(side, p, c);
  if (c[0] < dims[0] && c[1] < dims[1] && c[2] < dims[2]) {
    if (c[2] >= z0 && c[2] < z1) {
      uint64_t idx = (uint64_t)(c[2] - z0) * dims[0] * dims[1] + (uint64_t)c[1] * dims[0] + c[0];
      if (map_type == GVO_MAP_BIT) bit_set(map + idx * GVO_BITVOXEL_BYTES, (uint32_t)meaning & 255u);
      else ((int8_t*)map)[idx] = prob_insert_value(meaning);
    }
    return 0;
  }
  return 1;
}

/* meanings: n_meanings == 1 (all clouds) or == n_clouds. transforms: NULL or n_clouds x 16 floats
 * (row-major Matrix4f, GV/helpers/cuda_matrices.h:429-434), applied as
 * MetaPointCloud::transformSubCloud would before the insert (GV/helpers/MetaPointCloud.cu:624-649). */
int gvo_insert_meta(uint8_t* map, int map_type, const uint32_t* dims, uint32_t z0, uint32_t z1, float side,
                    const float* xyz, const uint32_t* cloud_sizes, uint32_t n_clouds,
                    const int32_t* meanings, uint32_t n_meanings, const float* transforms, int* any_outside) {
  int outside = 0;
  uint64_t base = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) {
    const int meaning = meanings[n_meanings == 1 ? 0 : c];
    for (uint64_t i = 0; i < cloud_sizes[c]; ++i) {
      const float* p = xyz + 3 * (base + i);
      float q[3];
      if (transforms) { xform(transforms + 16 * c, p, q); p = q; }
      outside |= insert_one(map, map_type, dims, z0, z1, side, p, meaning);
    }
    base += cloud_sizes[c];
  }
  if (any_outside) *any_outside = outside;
  return 0;
}

int gvo_insert_points(uint8_t* map, int map_type, const uint32_t* dims, uint32_t z0, uint32_t z1, float side,
                      const float* xyz, uint64_t n, int meaning, int* any_outside) {
  int outside = 0;
  for (uint64_t i = 0; i < n; ++i) outside |= insert_one(map, map_type, dims, z0, z1, side, xyz + 3 * i, meaning);
  if (any_outside) *any_outside = outside;
  return 0;
}

/* ---- insertSensorData without ray casting (the log-odds update) -------------------------------
 * kernelTransformSensorData  GV/voxelmap/kernels/VoxelMapOperations.cu:27-71
 * kernelInsertSensorData<DummyRayCaster>, cut_real_robot=false  GV/voxelmap/kernels/VoxelMapOperations.hpp:559-613
 * pose12 = row-major 3x4 [R | t] (Sensor::orientation, Sensor::position). delta = cSENSOR_MODEL_OCCUPIED (72,
 * GV/voxelmap/kernels/VoxelMapOperations.h:38-39). Sequential => no lost updates. n is the frame size
 * (the reference additionally caps at the voxel count, :569). */
/* one sensor point -> slab-local voxel index of its hit, or -1 (NaN, outside the map, another slab) */
static inline int64_t logodds_hit_index(const uint32_t* dims, uint32_t z0, uint32_t z1, float side, const float* M,
                                        const float* sensor_pos, int sensor_inside, const float* raw) {
  float w[3];
  xform(M, raw, w);
  if (w[2] < 0) { /* VoxelMapOperations.cu:56-64: re-project onto z=0 with a DOUBLE factor */
    double factor = (double)((float)(sensor_pos[2] / (sensor_pos[2] - w[2])));
    float r2[3];
    r2[0] = (float)((double)raw[0] * factor);
    r2[1] = (float)((double)raw[1] * factor);
    r2[2] = (float)((double)raw[2] * factor);
    xform(M, r2, w);
  }
  if (isnan(w[0]) || isnan(w[1]) || isnan(w[2])) return -1;
  uint32_t c[3];
  gvo_map_to_voxels(side, w, c);
  if (!(c[0] < dims[0] && c[1] < dims[1] && c[2] < dims[2] && sensor_inside)) return -1;
  if (!(c[2] >= z0 && c[2] < z1)) return -1;
  return (int64_t)((uint64_t)(c[2] - z0) * dims[0] * dims[1] + (uint64_t)c[1] * dims[0] + c[0]);
}

int gvo_update_logodds(int8_t* map, const uint32_t* dims, uint32_t z0, uint32_t z1, float side,
                       const float* xyz_sensor, uint64_t n, const float* pose12, int delta) {
  float M[16];
  memcpy(M, pose12, 12 * sizeof(float));
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
  const float sensor_pos[3] = {pose12[3], pose12[7], pose12[11]};
  uint32_t sc[3];
  gvo_map_to_voxels(side, sensor_pos, sc);
  const int sensor_inside = sc[0] < dims[0] && sc[1] < dims[1] && sc[2] < dims[2];
  uint64_t nvox = (uint64_t)dims[0] * dims[1] * dims[2];
  if (n > nvox) n = nvox;
  for (uint64_t i = 0; i < n; ++i) {
    const int64_t idx = logodds_hit_index(dims, z0, z1, side, M, sensor_pos, sensor_inside, xyz_sensor + 3 * i);
    if (idx >= 0) map[idx] = prob_update(map[idx], delta);
  }
  return 0;
}

/* The same update as an OpenMP loop over the points on all host threads (the timed "straightforward OpenMP host loop of
 * identical semantics" of bench.py's CPU baseline): the read-modify-write of a voxel is a compare-and-swap on its
 * byte, so no update is lost, and equal-sign saturating adds commute (clamp is monotone), so the map is byte-identical
 * to the sequential loop above for every thread schedule (tests/test_oracle_golden.py). */
int gvo_update_logodds_omp(int8_t* map, const uint32_t* dims, uint32_t z0, uint32_t z1, float side,
                           const float* xyz_sensor, uint64_t n, const float* pose12, int delta) {
  float M[16];
  memcpy(M, pose12, 12 * sizeof(float));
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
  const float sensor_pos[3] = {pose12[3], pose12[7], pose12[11]};
  uint32_t sc[3];
  gvo_map_to_voxels(side, sensor_pos, sc);
  const int sensor_inside = sc[0] < dims[0] && sc[1] < dims[1] && sc[2] < dims[2];
  uint64_t nvox = (uint64_t)dims[0] * dims[1] * dims[2];
  if (n > nvox) n = nvox;
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    const int64_t idx = logodds_hit_index(dims, z0, z1, side, M, sensor_pos, sensor_inside, xyz_sensor + 3 * i);
    if (idx < 0) continue;
    int8_t old = __atomic_load_n(&map[idx], __ATOMIC_RELAXED);
    for (;;) {
      const int8_t nv = prob_update(old, delta);
      if (nv == old) break;
      if (__atomic_compare_exchange_n(&map[idx], &old, nv, 1, __ATOMIC_RELAXED, __ATOMIC_RELAXED)) break;
    }
  }
  return 0;
}

/* ---- DefaultCollider::collide, GV/voxel/DefaultCollider.hpp:51-80 ----------------------------- */
static inline int collide_default(const uint8_t* a, int a_type, const uint8_t* b, int b_type, int8_t t1, int8_t t2) {
  if (a_type == GVO_MAP_PROB && b_type == GVO_MAP_PROB) return (int8_t)a[0] >= t1 && (int8_t)b[0] >= t2;
  if (a_type == GVO_MAP_PROB) return (int8_t)a[0] >= t1 && !bit_none_but_empty(b);
  if (b_type == GVO_MAP_PROB) return (int8_t)b[0] >= t2 && !bit_none_but_empty(a);
  return !bit_none_but_empty(a) && !bit_none_but_empty(b);
}
static inline size_t vox_bytes(int t) { return t == GVO_MAP_PROB ? 1 : GVO_BITVOXEL_BYTES; }

/* ---- collideWith map x map ---------------------------------------------------------------------
 * collisionCheckWithCounterRelativeTransform  GV/voxelmap/TemplateVoxelMap.hpp:486-519
 * kernelCollideVoxelMapsDebug                 GV/voxelmap/kernels/VoxelMapOperations.hpp:129-184
 * count of i in [0,n) with collide(a[i+off], b[i]); `off` = signed LINEAR offset of a's base pointer
 * (getVoxelPtrSignedOffset, VoxelMapOperations.h:92-106). Where i+off leaves [0,n) the reference reads
 * outside its allocation (undefined); here those i are skipped. mark != 0 reproduces the side effect
 * a[i+off].insert(eBVM_COLLISION) (:149): bit 2 set / occupancy := 127. */
int gvo_collide_map_map(uint8_t* a, int a_type, const uint8_t* b, int b_type, uint64_t n, float thr,
                        int64_t off, int mark, uint64_t* count) {
  const int8_t t = (int8_t)gvo_float_to_probability(thr);
  const size_t sa = vox_bytes(a_type), sb = vox_bytes(b_type);
  uint64_t cnt = 0;
#pragma omp parallel for schedule(static) reduction(+ : cnt)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    int64_t j = i + off;
    if (j < 0 || j >= (int64_t)n) continue;
    uint8_t* va = a + (size_t)j * sa;
    if (collide_default(va, a_type, b + (size_t)i * sb, b_type, t, t)) {
      ++cnt;
      if (mark) {
        if (a_type == GVO_MAP_PROB) ((int8_t*)va)[0] = 127;
        else bit_set(va, GVO_EBVM_COLLISION);
      }
    }
  }
  *count = cnt;
  return 0;
}

/* the same with DefaultCollider(Probability threshold1, Probability threshold2) (GV/voxel/DefaultCollider.hpp:45-49):
 * t1 applies to a ProbabilisticVoxel on side a, t2 to one on side b (collide_default above, :51-80) */
int gvo_collide_map_map_thresholds(uint8_t* a, int a_type, const uint8_t* b, int b_type, uint64_t n, int t1, int t2,
                                   int64_t off, int mark, uint64_t* count) {
  const size_t sa = vox_bytes(a_type), sb = vox_bytes(b_type);
  uint64_t cnt = 0;
#pragma omp parallel for schedule(static) reduction(+ : cnt)
  for (int64_t i = 0; i < (int64_t)n; ++i) {
    int64_t j = i + off;
    if (j < 0 || j >= (int64_t)n) continue;
    uint8_t* va = a + (size_t)j * sa;
    if (collide_default(va, a_type, b + (size_t)i * sb, b_type, (int8_t)t1, (int8_t)t2)) {
      ++cnt;
      if (mark) {
        if (a_type == GVO_MAP_PROB) ((int8_t*)va)[0] = 127;
        else bit_set(va, GVO_EBVM_COLLISION);
      }
    }
  }
  *count = cnt;
  return 0;
}

/* collisionCheck (boolean), GV/voxelmap/TemplateVoxelMap.hpp:329-414, kernelCollideVoxelMaps :78-120 */
int gvo_collide_map_map_any(const uint8_t* a, int a_type, const uint8_t* b, int b_type, uint64_t n, float thr, int* flag) {
  uint64_t c = 0;
  gvo_collide_map_map((uint8_t*)a, a_type, b, b_type, n, thr, 0, 0, &c);
  *flag = c != 0;
  return 0;
}

/* ---- collideWithTypes BitVectorVoxelMap x ProbVoxelMap -----------------------------------------
 * BitVoxelMap::collideWithTypes         GV/voxelmap/BitVoxelMap.hpp:212-217
 * collisionCheckBitvector               GV/voxelmap/BitVoxelMap.hpp:139-188
 * kernelCollideVoxelMapsBitvector       GV/voxelmap/kernels/VoxelMapOperations.hpp:187-239
 * SVCollider::collide(Bit,Prob,*mask)   GV/voxel/SVCollider.hpp:96-106
 * mask32 is OR-ed into, never reset (BitVoxelMap.hpp:178); offset is ignored by the reference. */
int gvo_collide_bit_prob_types(uint8_t* a, const int8_t* b, uint64_t n, float thr, int mark,
                               uint8_t* mask32, uint64_t* count) {
  const int8_t t2 = (int8_t)gvo_float_to_probability(thr); /* SVCollider::floatToProbability, SVCollider.hpp:131-135 */
  uint64_t cnt = 0;
  for (uint64_t i = 0; i < n; ++i) {
    uint8_t* va = a + i * GVO_BITVOXEL_BYTES;
    if (b[i] >= t2 && !bit_none_but_empty(va)) {
      for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) mask32[k] |= va[k];
      ++cnt;
      if (mark) bit_set(va, GVO_EBVM_COLLISION);
    }
  }
  *count = cnt;
  return 0;
}

/* ---- clearBit / clearBits: GV/voxelmap/kernels/VoxelMapOperations.hpp:45-72 ------------------- */
int gvo_clear_bits(uint8_t* a, uint64_t n, const uint8_t* mask32) {
#pragma omp parallel for schedule(static)
  for (int64_t i = 0; i < (int64_t)n; ++i)
    for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) a[(size_t)i * GVO_BITVOXEL_BYTES + k] &= (uint8_t)~mask32[k];
  return 0;
}

/* ---- BitVectorVoxelList -------------------------------------------------------------------------
 * Storage: SoA id[uint32] / coord[3 x uint32] / voxel[32 B], GV/voxellist/TemplateVoxelList.h:250-252.
 * Insert appends one entry per point WITHOUT a bounds check (ids wrap in uint32 arithmetic,
 * kernels GV/voxellist/kernels/VoxelListOperations.hpp:41-120), then make_unique
 * (GV/voxellist/TemplateVoxelList.hpp:141-208): stable sort by id, OR-merge equal ids into the first
 * entry of each run (Merge, :41-65), drop the rest (unique_by_key keeps the first => its coord).
 * The caller provides capacity for old_size + n_points entries; *size is updated. */
typedef struct { uint32_t id; uint32_t src; } gvo_key_t;
static int key_cmp(const void* pa, const void* pb) {
  const gvo_key_t* a = (const gvo_key_t*)pa; const gvo_key_t* b = (const gvo_key_t*)pb;
  if (a->id != b->id) return a->id < b->id ? -1 : 1;
  return a->src < b->src ? -1 : (a->src > b->src); /* src tie-break == stable sort */
}
int gvo_list_insert_meta(uint32_t* ids, uint32_t* coords, uint8_t* voxels, uint64_t* size,
                         const uint32_t* dims, float side, const float* xyz, const uint32_t* cloud_sizes,
                         uint32_t n_clouds, const int32_t* meanings, uint32_t n_meanings) {
  uint64_t old = *size, pos = old, base = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) {
    const int meaning = meanings[n_meanings == 1 ? 0 : c];
    for (uint64_t i = 0; i < cloud_sizes[c]; ++i, ++pos) {
      uint32_t v[3];
      gvo_map_to_voxels(side, xyz + 3 * (base + i), v);
      coords[3 * pos] = v[0]; coords[3 * pos + 1] = v[1]; coords[3 * pos + 2] = v[2];
      memset(voxels + pos * GVO_BITVOXEL_BYTES, 0, GVO_BITVOXEL_BYTES);
      bit_set(voxels + pos * GVO_BITVOXEL_BYTES, (uint32_t)meaning & 255u);
      /* getVoxelIndexUnsigned in uint32 arithmetic, VoxelMapOperations.h:54-58 */
      ids[pos] = v[2] * dims[0] * dims[1] + v[1] * dims[0] + v[0];
    }
    base += cloud_sizes[c];
  }
  const uint64_t total = pos;
  if (total == 0) { *size = 0; return 0; }
  gvo_key_t* keys = (gvo_key_t*)malloc(total * sizeof(gvo_key_t));
  uint32_t* ids2 = (uint32_t*)malloc(total * 4);
  uint32_t* co2 = (uint32_t*)malloc(total * 12);
  uint8_t* vx2 = (uint8_t*)malloc(total * GVO_BITVOXEL_BYTES);
  if (!keys || !ids2 || !co2 || !vx2) { free(keys); free(ids2); free(co2); free(vx2); return -1; }
  for (uint64_t i = 0; i < total; ++i) { keys[i].id = ids[i]; keys[i].src = (uint32_t)i; }
  qsort(keys, total, sizeof(gvo_key_t), key_cmp);
  uint64_t out = 0;
  for (uint64_t i = 0; i < total;) {
    uint64_t j = i;
    ids2[out] = keys[i].id;
    memcpy(co2 + 3 * out, coords + 3 * (uint64_t)keys[i].src, 12);
    memset(vx2 + out * GVO_BITVOXEL_BYTES, 0, GVO_BITVOXEL_BYTES);
    for (; j < total && keys[j].id == keys[i].id; ++j)
      for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k)
        vx2[out * GVO_BITVOXEL_BYTES + k] |= voxels[(uint64_t)keys[j].src * GVO_BITVOXEL_BYTES + k];
    ++out; i = j;
  }
  memcpy(ids, ids2, out * 4); memcpy(coords, co2, out * 12); memcpy(voxels, vx2, out * GVO_BITVOXEL_BYTES);
  free(keys); free(ids2); free(co2); free(vx2);
  *size = out;
  return 0;
}

/* isOccupied(float) of the map voxel as the list "types"/"mask" kernels use it: raw float threshold
 * for ProbabilisticVoxel (GV/voxel/ProbabilisticVoxel.hpp:94-98), !noneButEmpty for BitVoxel
 * (GV/voxel/BitVoxel.h:77-81). */
static inline int map_is_occupied_raw(const uint8_t* v, int map_type, float thr) {
  if (map_type == GVO_MAP_PROB) return (float)(int8_t)v[0] >= thr;
  return !bit_none_but_empty(v);
}

/* ---- list x map collisions -----------------------------------------------------------------------
 * mode 0  collideWith              BitVoxelList.hpp:70-84 -> TemplateVoxelList.hpp:731-774,
 *                                  kernelCollideWithVoxelMap VoxelListOperations.hpp:126-177 (DefaultCollider)
 * mode 1  collideWithTypes         BitVoxelList.hpp:127-218, kernel VoxelListOperations.hpp:182-240
 *                                  (mask_out cleared first, BitVoxelList.hpp:163)
 * mode 2  collideWithTypeMask      BitVoxelList.hpp:220-265, kernel VoxelListOperations.hpp:260-314
 * mode 3  per-meaning counts       semantics of BitVoxelList::collideCountingPerMeaning (BitVoxelList.hpp:300-344)
 *                                  applied to mode 0's colliding set; *count = sum over set bits. The reference has
 *                                  this for list x list only; list x map is an addition.
 * A map voxel is addressed as map + id + signed linear offset and used only if it lies inside the map
 * (pointer range test, VoxelListOperations.hpp:139-140); map_first/map_n describe a z-slab shard
 * (map_first = 0, map_n = N unsharded; N = dx*dy*dz is the full grid size used for the range test).
 * Colliding LIST voxels get eBVM_COLLISION when mark != 0 (:147, :214, :289). */
int gvo_collide_list_map(int mode, const uint32_t* ids, uint8_t* voxels, uint64_t size,
                         const uint8_t* map, int map_type, uint64_t n_total, uint64_t map_first, uint64_t map_n,
                         float thr, int64_t off, int mark, const uint8_t* mask_in32, uint8_t* mask_out32,
                         uint64_t* per_meaning256, uint64_t* count) {
  const int8_t t = (int8_t)gvo_float_to_probability(thr);
  const size_t sm = vox_bytes(map_type);
  uint64_t cnt = 0;
  if (mode == 1 && mask_out32) memset(mask_out32, 0, GVO_BITVOXEL_BYTES);
  for (uint64_t i = 0; i < size; ++i) {
    int64_t j = (int64_t)ids[i] + off;
    if (j < 0 || j >= (int64_t)n_total) continue;
    if ((uint64_t)j < map_first || (uint64_t)j >= map_first + map_n) continue; /* other shard's voxel */
    const uint8_t* mv = map + ((uint64_t)j - map_first) * sm;
    uint8_t* lv = voxels + i * GVO_BITVOXEL_BYTES;
    int hit;
    if (mode == 0 || mode == 3) hit = collide_default(lv, GVO_MAP_BIT, mv, map_type, t, t);
    else hit = map_is_occupied_raw(mv, map_type, thr);
    if (hit && mode == 2) {
      int any = 0;
      for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) any |= mask_in32[k] & lv[k];
      hit = any != 0;
    }
    if (!hit) continue;
    if (mode == 1) for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) mask_out32[k] |= lv[k];
    if (mode == 3) {
      for (int bit = 0; bit < 256; ++bit)
        if (lv[bit >> 3] & (1u << (bit & 7))) { per_meaning256[bit]++; ++cnt; }
    } else ++cnt;
    if (mark) bit_set(lv, GVO_EBVM_COLLISION);
  }
  *count = cnt;
  return 0;
}

/* ---- batch of robot configurations vs one environment map (config 3) ---------------------------
 * Per configuration this is the planner's state check (gvl_ompl_planning/gvl_ompl_planner_helper.cpp:169-202):
 *   robot_map.clearMap(); per link: transformSubCloud(link, T) (GV/robot/dh_robot/KinematicChain.cu:93-126);
 *   robot_map.insertMetaPointCloud(clouds, meanings); count = robot_map.collideWith(env).
 * counts[c] = number of voxels v with env[v] occupied (DefaultCollider side for `env_type`, threshold thr)
 * and robot[v] "set": mask32 == NULL -> !noneButEmpty(robot[v]) (DefaultCollider Bit x *);
 * mask32 != NULL -> (robot[v] & mask) != 0 (the BitVoxelMeaning collision mask, semantics of
 * collideWithTypeMask, GV/voxellist/kernels/VoxelListOperations.hpp:286). No side effects survive the call.
 * transforms: n_cfg x n_clouds x 16 floats. robot_scratch: caller-provided N x 32 zeroed bytes (left zeroed). */
int gvo_batch_robot_collide(const uint8_t* env, int env_type, const uint32_t* dims, float side,
                            const float* base_xyz, const uint32_t* cloud_sizes, uint32_t n_clouds,
                            const int32_t* meanings, const float* transforms, uint32_t n_cfg,
                            const uint8_t* mask32, float thr, uint8_t* robot_scratch, uint64_t* counts) {
  const int8_t t = (int8_t)gvo_float_to_probability(thr);
  uint64_t n_pts = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) n_pts += cloud_sizes[c];
  uint64_t* touched = (uint64_t*)malloc((n_pts ? n_pts : 1) * sizeof(uint64_t));
  if (!touched) return -1;
  const size_t se = vox_bytes(env_type);
  for (uint32_t cfg = 0; cfg < n_cfg; ++cfg) {
    uint64_t nt = 0, base = 0;
    for (uint32_t c = 0; c < n_clouds; ++c) {
      const float* M = transforms + ((uint64_t)cfg * n_clouds + c) * 16;
      for (uint64_t i = 0; i < cloud_sizes[c]; ++i) {
        float q[3]; uint32_t v[3];
        xform(M, base_xyz + 3 * (base + i), q);
        gvo_map_to_voxels(side, q, v);
        if (v[0] < dims[0] && v[1] < dims[1] && v[2] < dims[2]) {
          uint64_t idx = (uint64_t)v[2] * dims[0] * dims[1] + (uint64_t)v[1] * dims[0] + v[0];
          bit_set(robot_scratch + idx * GVO_BITVOXEL_BYTES, (uint32_t)meanings[c] & 255u);
          touched[nt++] = idx;
        }
      }
      base += cloud_sizes[c];
    }
    uint64_t cnt = 0;
    for (uint64_t k = 0; k < nt; ++k) {
      uint8_t* rv = robot_scratch + touched[k] * GVO_BITVOXEL_BYTES;
      int set = 0, nonzero = 0;
      for (int b = 0; b < GVO_BITVOXEL_BYTES; ++b) nonzero |= rv[b];
      if (!nonzero) continue; /* already counted + cleared via an earlier duplicate */
      if (mask32) { for (int b = 0; b < GVO_BITVOXEL_BYTES; ++b) set |= rv[b] & mask32[b]; }
      else set = !bit_none_but_empty(rv);
      const uint8_t* ev = env + touched[k] * se;
      int env_occ = env_type == GVO_MAP_PROB ? ((int8_t)ev[0] >= t) : !bit_none_but_empty(ev);
      if (set && env_occ) ++cnt;
      memset(rv, 0, GVO_BITVOXEL_BYTES);
    }
    counts[cfg] = cnt;
  }
  free(touched);
  return 0;
}

/* ---- input generators of the reference's tests, restated literally (fp32 accumulation matters) ---
 * GV/helpers/GeometryGeneration.cpp:90-105 createBoxOfPoints: `for (float x = min; x <= max; x += delta)`.
 * Returns the number of points; writes at most cap points (call with cap = 0 to size). */
uint64_t gvo_create_box_of_points(const float* mn, const float* mx, float delta, float* out_xyz, uint64_t cap) {
  uint64_t n = 0;
  for (float x = mn[0]; x <= mx[0]; x += delta)
    for (float y = mn[1]; y <= mx[1]; y += delta)
      for (float z = mn[2]; z <= mx[2]; z += delta) {
        if (n < cap) { out_xyz[3 * n] = x; out_xyz[3 * n + 1] = y; out_xyz[3 * n + 2] = z; }
        ++n;
      }
  return n;
}
/* GV/helpers/GeometryGeneration.cpp:159-221: x = i*2*side + side/2.0 (float product, DOUBLE sum, float store).
 * white == NULL -> createEquidistantPointsInBox, else createNonOverlapping3dCheckerboard. */
uint64_t gvo_create_equidistant_points(uint64_t max_points, const uint32_t* max_coords, float side,
                                       float* black_xyz, float* white_xyz) {
  uint64_t n = 0;
  for (uint32_t i = 0; i < (max_coords[0] - 1) / 2; i++)
    for (uint32_t j = 0; j < (max_coords[1] - 1) / 2; j++)
      for (uint32_t k = 0; k < (max_coords[2] - 1) / 2; k++) {
        if (n >= max_points) return n;
        black_xyz[3 * n] = (float)((double)((float)(i * 2) * side) + (double)side / 2.0);
        black_xyz[3 * n + 1] = (float)((double)((float)(j * 2) * side) + (double)side / 2.0);
        black_xyz[3 * n + 2] = (float)((double)((float)(k * 2) * side) + (double)side / 2.0);
        if (white_xyz) {
          white_xyz[3 * n] = (float)((double)((float)(i * 2 + 1) * side) + (double)side / 2.0);
          white_xyz[3 * n + 1] = (float)((double)((float)(j * 2 + 1) * side) + (double)side / 2.0);
          white_xyz[3 * n + 2] = (float)((double)((float)(k * 2 + 1) * side) + (double)side / 2.0);
        }
        ++n;
      }
  return n;
}

/* ---- batched forward kinematics of a DH chain ---------------------------------------------------
 * DHParameters::convertDHtoM, GV/robot/dh_robot/KinematicLink.cu:30-84: sin/cos are evaluated in DOUBLE on the
 * float angle and rounded to float; every product/sum after that is float with one rounding per operation
 * (host g++, no FMA contraction); b == 0 is kept as written ("a * ct - b * st").
 * KinematicChain::setConfiguration, GV/robot/dh_robot/KinematicChain.cu:104-124: link i's cloud is transformed by
 * the product of the DH matrices of links 0..i-1 (identity for link 0): out[cfg][i] = T_0 * ... * T_{i-1}, with
 * Matrix4f operator* (GV/helpers/cuda_matrices.h:772-803) summing left to right.
 * dh4: n_links x {d, theta, a, alpha}; joint_types: 0 REVOLUTE, 1 PRISMATIC (KinematicLink.h:37-41);
 * joint_values: n_cfg x n_links; out: n_cfg x n_links x 16 row-major. */
static void dh_to_m(const float* p, int joint_type, float value, float* m) {
  float d_current = p[0], theta_current = p[1];
  const float a = p[2], alpha = p[3], b = 0;
  if (joint_type == 1) d_current = p[0] + value;
  else if (joint_type == 0) theta_current = p[1] + value;
  const float ca = (float)cos((double)alpha), sa = (float)sin((double)alpha);
  const float ct = (float)cos((double)theta_current), st = (float)sin((double)theta_current);
  m[0] = ct;   m[1] = -st * ca; m[2] = st * sa;   m[3] = a * ct - b * st;
  m[4] = st;   m[5] = ct * ca;  m[6] = -ct * sa;  m[7] = a * st + b * ct;
  m[8] = 0.0f; m[9] = sa;       m[10] = ca;       m[11] = d_current;
  m[12] = 0.0f; m[13] = 0.0f;   m[14] = 0.0f;     m[15] = 1.0f;
}

static void mat4_mul(const float* x, const float* y, float* r) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      r[4 * i + j] = x[4 * i + 0] * y[0 + j] + x[4 * i + 1] * y[4 + j] + x[4 * i + 2] * y[8 + j] + x[4 * i + 3] * y[12 + j];
}

int gvo_dh_chain(const float* dh4, const int32_t* joint_types, uint32_t n_links, const float* joint_values,
                 uint32_t n_cfg, float* out) {
  for (uint32_t c = 0; c < n_cfg; ++c) {
    float acc[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1}, m[16], t[16];
    for (uint32_t l = 0; l < n_links; ++l) {
      memcpy(out + ((size_t)c * n_links + l) * 16, acc, sizeof acc);
      dh_to_m(dh4 + 4 * l, joint_types[l], joint_values[(size_t)c * n_links + l], m);
      mat4_mul(acc, m, t);
      memcpy(acc, t, sizeof acc);
    }
  }
  return 0;
}

/* ---- swept-volume window check and ID shift ----------------------------------------------------
 * bitMarginCollisionCheck, GV/helpers/BitVector.h:416-471, restated with the same 64-bit window arithmetic:
 * the window holds 4 bytes of v2 above bit 32 and whatever has been shifted down below it; byte 0 of v2 loses its
 * four non-swept-volume bits, byte 0 of v1 does NOT. `low` = sv_offset % 8 moves the first visited bit (and so can
 * skip byte 0 and de-synchronise the refill), `high` = sv_offset / 8 is a bit shift of v1 -- the reference names
 * them the other way round. Every visited byte of `collisions` is overwritten, the others are left alone. */
int gvo_bit_margin_check(const uint8_t* v1, const uint8_t* v2, uint8_t* collisions, uint32_t margin, uint32_t sv_offset) {
  uint64_t window = 0;
  for (int k = 0; k < 4; ++k) window += (uint64_t)(k == 0 ? (v2[0] & 0xF0) : v2[k]) << (32 + 8 * k);
  const uint32_t low = (uint8_t)(sv_offset % 8), high = (uint8_t)(sv_offset / 8);
  for (uint32_t i = 4 /* eBVM_SWEPT_VOLUME_START */ + low; i < 254 /* eBVM_SWEPT_VOLUME_END, GV/helpers/common_defines.h:57 */; i += 8) {
    uint8_t out = 0;
    uint64_t probe = (uint64_t)v1[i >> 3] << (32 - margin + high);
    for (uint32_t s = 0; s <= 2 * margin; ++s) {
      out |= (uint8_t)((probe & window) >> (32 - margin + s));
      probe <<= 1;
    }
    collisions[i >> 3] = out;
    window >>= 8;
    if (i < 256 - 32) window += (uint64_t)v2[(i + 32) >> 3] << 56;
  }
  for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) if (collisions[k]) return 1;
  return 0;
}

/* performLeftShift, GV/helpers/BitVector.h:361-402: the vector moves `shift` (<= 56) bits towards bit 0 through a
 * 64-bit window; the four non-swept-volume bits of byte 0 are cleared */
void gvo_shift_left(uint8_t* v, uint32_t shift) {
  uint64_t window = 0;
  for (int k = 0; k < 8; ++k) window += (uint64_t)v[k] << (8 * k);
  for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) {
    uint8_t nb = (uint8_t)(window >> shift);
    if (k == 0) nb &= 0xF0;
    v[k] = nb;
    window >>= 8;
    if (k + 8 < GVO_BITVOXEL_BYTES) window += (uint64_t)v[k + 8] << 56;
  }
}

/* BitVoxelMap::collisionCheckBitvector(BitVoxelMap*, SVCollider(thr, margin), types, sv_offset),
 * GV/voxelmap/BitVoxelMap.hpp:91-135 + kernelCollideVoxelMapsBitvector, VoxelMapOperations.hpp:187-239, launched
 * with one thread per voxel (ceil(n/1024) blocks of 1024), so every thread's `temp` starts empty and the block
 * reduction ORs all of them: types32 |= union of the per-voxel collision bits, count = voxels with any.
 * collideWithTypes(BitVectorVoxelMap*) is this with margin 0, sv_offset 0 (BitVoxelMap.hpp:205-210). */
int gvo_collide_bitcheck(uint8_t* a, const uint8_t* b, uint64_t n, uint32_t margin, uint32_t sv_offset, int mark,
                         uint8_t* types32, uint64_t* count) {
  uint64_t cnt = 0;
  for (uint64_t i = 0; i < n; ++i) {
    uint8_t temp[GVO_BITVOXEL_BYTES] = {0};
    /* shortcut (keeps the CPU suite fast, changes nothing): an empty v1 probes with zeros, an empty v2 gives an empty
     * window -- every visited byte is then written as 0 */
    const uint64_t* qa = (const uint64_t*)(a + i * GVO_BITVOXEL_BYTES);
    const uint64_t* qb = (const uint64_t*)(b + i * GVO_BITVOXEL_BYTES);
    if ((qa[0] | qa[1] | qa[2] | qa[3]) == 0 || ((qb[0] & ~(uint64_t)0xF) | qb[1] | qb[2] | qb[3]) == 0) continue;
    if (gvo_bit_margin_check(a + i * GVO_BITVOXEL_BYTES, b + i * GVO_BITVOXEL_BYTES, temp, margin, sv_offset)) {
      if (mark) bit_set(a + i * GVO_BITVOXEL_BYTES, GVO_EBVM_COLLISION);
      for (int k = 0; k < GVO_BITVOXEL_BYTES; ++k) types32[k] |= temp[k];
      ++cnt;
    }
  }
  *count = cnt;
  return 0;
}

/* BitVoxelMap::shiftLeftSweptVolumeIDs, GV/voxelmap/BitVoxelMap.hpp:234-246 (kernelShiftBitVector,
 * VoxelMapOperations.hpp:615-624); shift > 56 is refused there with an error message and no change */
int gvo_shift_map(uint8_t* a, uint64_t n, uint32_t shift) {
  if (shift > 56) return 1;
  for (uint64_t i = 0; i < n; ++i) gvo_shift_left(a + i * GVO_BITVOXEL_BYTES, shift);
  return 0;
}

/* ---- insertSensorData with ray casting and robot cut -------------------------------------------
 * kernelInsertSensorData<RayCaster>, GV/voxelmap/kernels/VoxelMapOperations.hpp:559-613, and RayCaster::rayCast,
 * GV/voxelmap/kernels/VoxelMapOperations.h:199-323: an integer line walk from the sensor's voxel to the hit voxel
 * that visits 1 + |dx| + |dy| + |dz| cells (both ends included) and applies updateOccupancy(cSENSOR_MODEL_FREE = -10)
 * to each, then updateOccupancy(cSENSOR_MODEL_OCCUPIED = +72) at the hit voxel. With cut_real_robot a point whose hit
 * voxel has `robot_bit` set in the robot map is dropped altogether (:578-592).
 *
 * ORDER. The reference updates voxels with non-atomic read-modify-writes from one thread per point, so its result
 * depends on the schedule (lost updates, and clamp(-128 - 10) = -127 vs clamp(-128 + 72) = -56 do not commute). The
 * defined order here and in the CUDA path: ALL free-space decrements of the frame first, then ALL hits. Decrements
 * commute among themselves and so do hits, so the result is min(max(v - 10*kf, -127) + 72*ko, 127) per voxel
 * (v = -128 -> the first update clamps as in ProbabilisticVoxel::updateOccupancy). It equals the reference run
 * sequentially wherever a voxel is not both crossed and hit by DIFFERENT rays (tests pin exactly that).
 * kf / ko (nullable): per-voxel counts of decrements / hits of this call, for tests. */
static void ray_walk(int8_t* map, const uint32_t* dims, uint32_t z0, uint32_t z1, const uint32_t* from,
                     const uint32_t* to, int delta_free, uint32_t* kf) {
  int32_t d[3], inc[3], p[3];
  for (int k = 0; k < 3; ++k) {
    if (to[k] > from[k]) { d[k] = (int32_t)(to[k] - from[k]); inc[k] = 1; }
    else { d[k] = (int32_t)(from[k] - to[k]); inc[k] = -1; }
    p[k] = (int32_t)from[k];
  }
  int32_t n = 1 + d[0] + d[1] + d[2];
  int32_t e_xy = d[0] - d[1], e_xz = d[0] - d[2], e_yz = d[1] - d[2];
  d[0] *= 2; d[1] *= 2; d[2] *= 2;
  for (; n > 0; --n) {
    if ((uint32_t)p[2] >= z0 && (uint32_t)p[2] < z1) {
      const uint64_t idx = (uint64_t)((uint32_t)p[2] - z0) * dims[0] * dims[1] + (uint64_t)p[1] * dims[0] + (uint64_t)p[0];
      map[idx] = prob_update(map[idx], delta_free);
      if (kf) ++kf[idx];
    }
    if (e_xy > 0 && e_xz > 0) { p[0] += inc[0]; e_xy -= d[1]; e_xz -= d[2]; }
    else if (e_yz > 0)        { p[1] += inc[1]; e_xy += d[0]; e_yz -= d[2]; }
    else                      { p[2] += inc[2]; e_xz += d[0]; e_yz += d[1]; }
  }
}

int gvo_insert_sensor_data(int8_t* map, const uint32_t* dims, uint32_t z0, uint32_t z1, float side,
                           const float* xyz_sensor, uint64_t n, const float* pose12, int raycast,
                           const uint8_t* robot_map /* whole BitVectorVoxelMap or NULL */, int robot_bit,
                           int delta_free, int delta_occ, uint32_t* kf, uint32_t* ko) {
  float M[16];
  memcpy(M, pose12, 12 * sizeof(float));
  M[12] = 0; M[13] = 0; M[14] = 0; M[15] = 1;
  const float sensor_pos[3] = {pose12[3], pose12[7], pose12[11]};
  uint32_t sc[3];
  gvo_map_to_voxels(side, sensor_pos, sc);
  if (!(sc[0] < dims[0] && sc[1] < dims[1] && sc[2] < dims[2])) return 0;
  uint64_t nvox = (uint64_t)dims[0] * dims[1] * dims[2];
  if (n > nvox) n = nvox;
  for (int phase = raycast ? 0 : 1; phase < 2; ++phase) {
    for (uint64_t i = 0; i < n; ++i) {
      const float* raw = xyz_sensor + 3 * i;
      float w[3];
      xform(M, raw, w);
      if (w[2] < 0) {
        double factor = (double)((float)(sensor_pos[2] / (sensor_pos[2] - w[2])));
        float r2[3];
        r2[0] = (float)((double)raw[0] * factor);
        r2[1] = (float)((double)raw[1] * factor);
        r2[2] = (float)((double)raw[2] * factor);
        xform(M, r2, w);
      }
      if (isnan(w[0]) || isnan(w[1]) || isnan(w[2])) continue;
      uint32_t c[3];
      gvo_map_to_voxels(side, w, c);
      if (!(c[0] < dims[0] && c[1] < dims[1] && c[2] < dims[2])) continue;
      if (robot_map) {
        const uint64_t g = ((uint64_t)c[2] * dims[1] + c[1]) * dims[0] + c[0];
        const uint8_t* rv = robot_map + g * GVO_BITVOXEL_BYTES;
        if (rv[robot_bit >> 3] & (1u << (robot_bit & 7))) continue; /* the sensor sees the robot: no update */
      }
      if (phase == 0) ray_walk(map, dims, z0, z1, sc, c, delta_free, kf);
      else if (c[2] >= z0 && c[2] < z1) {
        const uint64_t idx = (uint64_t)(c[2] - z0) * dims[0] * dims[1] + (uint64_t)c[1] * dims[0] + c[0];
        map[idx] = prob_update(map[idx], delta_occ);
        if (ko) ++ko[idx];
      }
    }
  }
  return 0;
}
