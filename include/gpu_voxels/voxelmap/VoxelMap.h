// gpu_voxels/voxelmap/VoxelMap.h -- dense voxel maps of the reference's API over the C ABI:
//   voxelmap::BitVectorVoxelMap  (GV/voxelmap/BitVoxelMap.h, .hpp:57-246; GV/voxelmap/VoxelMap.h typedefs)
//   voxelmap::ProbVoxelMap       (GV/voxelmap/ProbVoxelMap.h, .hpp:52-155)
// common part = TemplateVoxelMap (GV/voxelmap/TemplateVoxelMap.h:52-260, .hpp:54-664). Every method is a few lines
// around one gvx_* call; all voxel work happens in libgvx.so (hand-written sm_100a CUDA).
#ifndef GPU_VOXELS_B200_VOXELMAP_VOXELMAP_H
#define GPU_VOXELS_B200_VOXELMAP_VOXELMAP_H

#include <climits>
#include <iostream>

#include <gpu_voxels/GpuVoxelsMap.h>
#include <gpu_voxels/voxel/DefaultCollider.h>
#include <gpu_voxels/voxel/ProbabilisticVoxel.h>
#include <gpu_voxels/voxel/SVCollider.h>

namespace gpu_voxels {
namespace voxelmap {

class ProbVoxelMap;

// non-template common base (TemplateVoxelMap<Voxel> in the reference)
class AbstractVoxelMap : public GpuVoxelsMap
{
public:
  virtual ~AbstractVoxelMap() { if (m_handle) gvx_map_destroy(m_handle); }

  virtual void insertPointCloud(const std::vector<Vector3f>& points, const BitVoxelMeaning voxel_meaning)
  {
    lock_guard guard(this->m_mutex);
    if (points.empty()) return;
    insertChecked(reinterpret_cast<const float*>(&points[0]), points.size(), GVX_MEM_HOST, voxel_meaning);
  }
  virtual void insertPointCloud(const PointCloud& pointcloud, const BitVoxelMeaning voxel_meaning)
  {
    insertPointCloud(pointcloud.getConstDevicePointer(), uint32_t(pointcloud.getPointCloudSize()), voxel_meaning);
  }
  // device-pointer overload (GV/voxelmap/TemplateVoxelMap.hpp:586-607)
  virtual void insertPointCloud(const Vector3f* points_d, uint32_t size, const BitVoxelMeaning voxel_meaning)
  {
    lock_guard guard(this->m_mutex);
    if (!size) return;
    insertChecked(reinterpret_cast<const float*>(points_d), size, GVX_MEM_DEVICE, voxel_meaning);
  }
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud, BitVoxelMeaning voxel_meaning)
  {
    std::vector<BitVoxelMeaning> one(1, voxel_meaning);
    insertMeta(meta_point_cloud, one, NULL);
  }
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud,
                                    const std::vector<BitVoxelMeaning>& voxel_meanings)
  {
    if (voxel_meanings.size() != meta_point_cloud.getNumberOfPointclouds())  // GV/voxelmap/TemplateVoxelMap.hpp:636-641
    {
      LOGGING_ERROR_C(VoxelmapLog, TemplateVoxelMap, "Number of voxel meanings differs from number of sub-clouds" << endl);
      return;
    }
    insertMeta(meta_point_cloud, voxel_meanings, NULL);
  }
  // extension: per-sub-cloud transforms fused into the insert (replaces transformSubCloud x n + insertMetaPointCloud)
  void insertMetaPointCloudTransformed(const MetaPointCloud& meta_point_cloud,
                                       const std::vector<BitVoxelMeaning>& voxel_meanings,
                                       const std::vector<Matrix4f>& transforms)
  {
    if (transforms.size() != meta_point_cloud.getNumberOfPointclouds() ||
        (voxel_meanings.size() != 1 && voxel_meanings.size() != transforms.size()))
    {
      LOGGING_ERROR_C(VoxelmapLog, TemplateVoxelMap, "Number of transforms / meanings differs from number of sub-clouds" << endl);
      return;
    }
    insertMeta(meta_point_cloud, voxel_meanings, transforms.empty() ? NULL : transforms[0].data());
  }

  virtual void clearMap()
  {
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_map_clear(m_handle), "clearMap");
  }
  virtual std::size_t getMemoryUsage() const
  {
    return std::size_t(m_dim.x) * m_dim.y * m_dim.z * (m_map_type == MT_PROBAB_VOXELMAP ? 1 : 32);
  }
  virtual Vector3ui getDimensions() const { return m_dim; }
  virtual Vector3f getMetricDimensions() const { return Vector3f(float(m_dim.x), float(m_dim.y), float(m_dim.z)) * m_voxel_side_length; }
  float getVoxelSideLength() const { return m_voxel_side_length; }
  uint32_t getVoxelMapSize() const { return m_dim.x * m_dim.y * m_dim.z; }
  void* getVoidDeviceDataPtr() { void* p = NULL; gvx_map_device_ptr(m_handle, &p, NULL, NULL); return p; }
  const void* getConstVoidDeviceDataPtr() const { const void* p = NULL; gvx_map_device_ptr_const(m_handle, &p, NULL, NULL); return p; }
  virtual gvx_handle_t getHandle() const { return m_handle; }

  // GV/voxelmap/AbstractVoxelMap.h:60-66, .cu:23-33; BitVoxelMap.hpp:220-226, ProbVoxelMap.hpp:104-108
  virtual MapType getTemplateType() const = 0;
  virtual bool needsRebuild() const
  {
    LOGGING_ERROR_C(VoxelmapLog, AbstractVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
    return false;
  }
  virtual bool rebuild()
  {
    LOGGING_ERROR_C(VoxelmapLog, AbstractVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
    return false;
  }
  virtual bool insertRobotConfiguration(const MetaPointCloud*, bool)
  {
    LOGGING_ERROR_C(VoxelmapLog, TemplateVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
    return false;
  }

  // the reference's binary file format (GV/voxelmap/TemplateVoxelMap.hpp:666-799)
  virtual bool writeToDisk(const std::string path)
  {
    lock_guard guard(this->m_mutex);
    return gvxCheck(gvx_map_write_file(m_handle, path.c_str()), "writeToDisk");
  }
  // TemplateVoxelMap::merge x2 (GV/voxelmap/TemplateVoxelMap.hpp:801-813): not supported by dense maps there either
  virtual bool merge(const GpuVoxelsMapSharedPtr, const Vector3i& = Vector3i(), const BitVoxelMeaning* = NULL)
  {
    LOGGING_ERROR_C(VoxelmapLog, TemplateVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_YET_SUPPORTED << endl);
    return false;
  }
  virtual bool merge(const GpuVoxelsMapSharedPtr, const Vector3f&, const BitVoxelMeaning* = NULL)
  {
    LOGGING_ERROR_C(VoxelmapLog, TemplateVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_YET_SUPPORTED << endl);
    return false;
  }
  virtual bool readFromDisk(const std::string path)
  {
    lock_guard guard(this->m_mutex);
    return gvxCheck(gvx_map_read_file(m_handle, path.c_str()), "readFromDisk");
  }

  // collisionCheck: boolean, no side effect (GV/voxelmap/TemplateVoxelMap.hpp:329-414)
  bool collisionCheck(const AbstractVoxelMap* other, float coll_threshold = 1.0)
  {
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    int flag = 0;
    // worker threads query on their own collision lane; GVX_SHARD=z (a collective) and failures fall back to the serial call
    const gvx_handle_t lane = detail::threadLane();
    if (lane && gvx_lane_collide_map_map_any(lane, m_handle, other->m_handle, coll_threshold, &flag) == GVX_OK)
      return flag != 0;
    gvxCheck(gvx_collide_map_map_any(m_handle, other->m_handle, coll_threshold, &flag), "collisionCheck");
    return flag != 0;
  }

  // TemplateVoxelMap::collisionCheckWithCounter / ...RelativeTransform (GV/voxelmap/TemplateVoxelMap.hpp:477-519) with
  // a collider object: number of colliding voxels, `this` gets eBVM_COLLISION marks. DefaultCollider(t1, t2) with two
  // different thresholds is honoured (t1 for this map's ProbabilisticVoxels, t2 for the other's).
  uint32_t collisionCheckWithCounter(AbstractVoxelMap* other, DefaultCollider collider = DefaultCollider())
  {
    return collisionCheckWithCounterRelativeTransform(other, collider);
  }
  uint32_t collisionCheckWithCounterRelativeTransform(AbstractVoxelMap* other, DefaultCollider collider = DefaultCollider(),
                                                      const Vector3i& offset = Vector3i())
  {
    // the collider's two int8 thresholds go to the kernels as they are (threshold1: this map, threshold2: other)
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    const int32_t off[3] = {offset.x, offset.y, offset.z};
    uint64_t n = 0;
    if (!gvxCheck(gvx_collide_map_map_thresholds(m_handle, other->m_handle, int(collider.getThreshold1()),
                                                 int(collider.getThreshold2()), off, 1, &n), "collisionCheckWithCounter"))
      return 0;
    return uint32_t(n);
  }

protected:
  AbstractVoxelMap(const Vector3ui dim, const float voxel_side_length, const MapType map_type, void* dev_data)
    : m_dim(dim), m_voxel_side_length(voxel_side_length), m_handle(0)
  {
    gvxEnsureInit();
    this->m_map_type = map_type;
    const int t = map_type == MT_PROBAB_VOXELMAP ? GVX_MT_PROBAB_VOXELMAP : GVX_MT_BITVECTOR_VOXELMAP;
    int rc = dev_data ? gvx_map_wrap(&m_handle, t, dev_data, dim.x, dim.y, dim.z, voxel_side_length, 0, dim.z)
                      : gvx_map_create(&m_handle, t, dim.x, dim.y, dim.z, voxel_side_length);
    gvxCheck(rc, "voxel map construction");
  }

  // collisionCheckWithCounterRelativeTransform (GV/voxelmap/TemplateVoxelMap.hpp:486-519); marks `this`
  size_t collideMapMap(const AbstractVoxelMap* other, float coll_threshold, const Vector3i& offset)
  {
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    const int32_t off[3] = {offset.x, offset.y, offset.z};
    uint64_t n = 0;
    if (!gvxCheck(gvx_collide_map_map(m_handle, other->m_handle, coll_threshold, off, 1, &n), "collideWith")) return SSIZE_MAX;
    return size_t(n);
  }

  void insertChecked(const float* xyz, uint64_t n, int mem_kind, BitVoxelMeaning voxel_meaning)
  {
    int outside = 0;
    const int rc = gvx_insert_points(m_handle, xyz, n, mem_kind, int(voxel_meaning), &outside);
    if (rc == GVX_ERR_UNSUPPORTED)  // ProbVoxelMap: only eBVM_OCCUPIED (GV/voxelmap/ProbVoxelMap.hpp:118-140)
    {
      LOGGING_ERROR_C(VoxelmapLog, ProbVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
      return;
    }
    gvxCheck(rc, "insertPointCloud");
    if (outside) LOGGING_WARNING_C(VoxelmapLog, TemplateVoxelMap, "You tried to insert points that lie outside the map dimensions!" << endl);
  }
  void insertMeta(const MetaPointCloud& mpc, const std::vector<BitVoxelMeaning>& meanings, const float* transforms)
  {
    lock_guard guard(this->m_mutex);
    if (mpc.getNumberOfPointclouds() == 0 || mpc.getAccumulatedPointcloudSize() == 0) return;
    const std::vector<int32_t> mg = detail::meaningsToInts(meanings);
    int outside = 0;
    const int rc = gvx_insert_meta(m_handle, reinterpret_cast<const float*>(mpc.getDeviceBasePointer()), GVX_MEM_DEVICE,
                                   &mpc.getPointcloudSizes()[0], mpc.getNumberOfPointclouds(), &mg[0], uint32_t(mg.size()),
                                   transforms, &outside);
    if (rc == GVX_ERR_UNSUPPORTED)
    {
      LOGGING_ERROR_C(VoxelmapLog, ProbVoxelMap, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
      return;
    }
    gvxCheck(rc, "insertMetaPointCloud");
    if (outside) LOGGING_WARNING_C(VoxelmapLog, TemplateVoxelMap, "You tried to insert points that lie outside the map dimensions!" << endl);
  }

  Vector3ui m_dim;
  float m_voxel_side_length;
  gvx_handle_t m_handle;
};

template <std::size_t length> class BitVoxelMap;

class ProbVoxelMap : public AbstractVoxelMap
{
public:
  ProbVoxelMap(const Vector3ui dim, const float voxel_side_length, const MapType map_type = MT_PROBAB_VOXELMAP)
    : AbstractVoxelMap(dim, voxel_side_length, map_type, NULL), m_init_sensor(false) {}
  ProbVoxelMap(int8_t* dev_data, const Vector3ui dim, const float voxel_side_length, const MapType map_type = MT_PROBAB_VOXELMAP)
    : AbstractVoxelMap(dim, voxel_side_length, map_type, dev_data), m_init_sensor(false) {}

  // typed view of the raw voxel array (GV/voxelmap/TemplateVoxelMap.h:getDeviceDataPtr): one int8 log-odds byte per voxel
  ProbabilisticVoxel* getDeviceDataPtr() { return static_cast<ProbabilisticVoxel*>(getVoidDeviceDataPtr()); }
  const ProbabilisticVoxel* getConstDeviceDataPtr() const { return static_cast<const ProbabilisticVoxel*>(getConstVoidDeviceDataPtr()); }

  virtual MapType getTemplateType() const { return MT_PROBAB_VOXELMAP; }
  // TemplateVoxelMap::printVoxelMapData (GV/voxelmap/TemplateVoxelMap.hpp:281-286): every voxel, one per line
  virtual void printVoxelMapData()
  {
    lock_guard guard(this->m_mutex);
    std::vector<int8_t> host(getVoxelMapSize());
    if (host.empty() || !gvxCheck(gvx_map_download(m_handle, &host[0]), "printVoxelMapData")) return;
    std::cout << "VoxelMap dump: " << std::endl;
    for (std::size_t i = 0; i < host.size(); ++i) std::cout << i << ": " << int(host[i]) << std::endl;
  }

  // ProbVoxelMap only knows eBVM_OCCUPIED; anything else is a logged no-op (GV/voxelmap/ProbVoxelMap.hpp:110-116)
  virtual void clearBitVoxelMeaning(BitVoxelMeaning voxel_meaning)
  {
    if (voxel_meaning != eBVM_OCCUPIED)
      LOGGING_ERROR_C(VoxelmapLog, ProbVoxelMap, GPU_VOXELS_MAP_ONLY_SUPPORTS_BVM_OCCUPIED << endl);
    else
      this->clearMap();
  }

  // sensor model (GV/voxelmap/TemplateVoxelMap.hpp:835-909)
  void initSensorSettings(const Sensor& sensor) { lock_guard guard(this->m_mutex); m_sensor = sensor; m_init_sensor = true; }
  void updateSensorPose(const Sensor& sensor)
  {
    lock_guard guard(this->m_mutex);
    if (!m_init_sensor) { LOGGING_ERROR_C(VoxelmapLog, VoxelMap, "Initialize Sensor first!" << endl); return; }
    m_sensor.position = sensor.position;
    m_sensor.orientation = sensor.orientation;
  }
  // insertSensorData (GV/voxelmap/ProbVoxelMap.hpp:52-102): points = sensor.data_size HOST points in the sensor frame;
  // enable_raycasting adds the free-space update along every ray (RayCaster, VoxelMapOperations.h:199-323);
  // cut_real_robot drops points whose hit voxel has bit `voxel_meaning` set in robot_map (a DEVICE pointer to the
  // voxels of a BitVectorVoxelMap of the same dimensions, as in the reference).
  template <std::size_t length>
  void insertSensorData(const Vector3f* points, const bool enable_raycasting, const bool cut_real_robot,
                        const BitVoxelMeaning voxel_meaning, BitVoxel<length>* robot_map = NULL)
  {
    lock_guard guard(this->m_mutex);
    if (!m_init_sensor) { LOGGING_ERROR_C(VoxelmapLog, EnvironmentMap, "Call initSensorSettings() first!" << endl); return; }
    const Matrix3f& R = m_sensor.orientation;
    const float pose[12] = {R.a11, R.a12, R.a13, m_sensor.position.x, R.a21, R.a22, R.a23, m_sensor.position.y,
                            R.a31, R.a32, R.a33, m_sensor.position.z};
    gvx_handle_t robot = 0;
    if (cut_real_robot)
    {
      if (!robot_map) { LOGGING_ERROR_C(VoxelmapLog, ProbVoxelMap, "cut_real_robot needs the robot map" << endl); return; }
      if (!gvxCheck(gvx_map_wrap(&robot, GVX_MT_BITVECTOR_VOXELMAP, robot_map, m_dim.x, m_dim.y, m_dim.z,
                                 m_voxel_side_length, 0, m_dim.z), "insertSensorData (robot map)")) return;
    }
    gvxCheck(gvx_insert_sensor_data(m_handle, reinterpret_cast<const float*>(points), m_sensor.data_size, GVX_MEM_HOST, pose,
                                    enable_raycasting ? 1 : 0, robot, int(voxel_meaning), cSENSOR_MODEL_FREE,
                                    cSENSOR_MODEL_OCCUPIED), "insertSensorData");
    gvx_sync();  // the reference returns after cudaDeviceSynchronize; `points` may be reused by the caller
    if (robot) gvx_map_destroy(robot);
  }
  void insertSensorData(const Vector3f* points, const bool enable_raycasting, const bool cut_real_robot,
                        const BitVoxelMeaning voxel_meaning)
  {
    insertSensorData<BIT_VECTOR_LENGTH>(points, enable_raycasting, cut_real_robot, voxel_meaning, NULL);
  }

  // GV/voxelmap/ProbVoxelMap.hpp:144-155
  size_t collideWith(const BitVoxelMap<BIT_VECTOR_LENGTH>* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i());
  size_t collideWith(const ProbVoxelMap* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    return collideMapMap(map, coll_threshold, offset);
  }

private:
  Sensor m_sensor;
  bool m_init_sensor;
};

template <std::size_t length>
class BitVoxelMap : public AbstractVoxelMap
{
public:
  // typed view of the raw voxel array (GV/voxelmap/TemplateVoxelMap.h:getDeviceDataPtr): 32 bytes per voxel
  BitVoxel<length>* getDeviceDataPtr() { return static_cast<BitVoxel<length>*>(getVoidDeviceDataPtr()); }
  const BitVoxel<length>* getConstDeviceDataPtr() const { return static_cast<const BitVoxel<length>*>(getConstVoidDeviceDataPtr()); }

  BitVoxelMap(const Vector3ui dim, const float voxel_side_length, const MapType map_type = MT_BITVECTOR_VOXELMAP)
    : AbstractVoxelMap(dim, voxel_side_length, map_type, NULL) {}
  BitVoxelMap(BitVoxel<length>* dev_data, const Vector3ui dim, const float voxel_side_length, const MapType map_type = MT_BITVECTOR_VOXELMAP)
    : AbstractVoxelMap(dim, voxel_side_length, map_type, dev_data) {}

  virtual MapType getTemplateType() const { return MT_BITVECTOR_VOXELMAP; }
  virtual void printVoxelMapData()
  {
    lock_guard guard(this->m_mutex);
    std::vector<BitVoxel<length> > host(getVoxelMapSize());
    if (host.empty() || !gvxCheck(gvx_map_download(m_handle, &host[0]), "printVoxelMapData")) return;
    std::cout << "VoxelMap dump: " << std::endl;
    for (std::size_t i = 0; i < host.size(); ++i) std::cout << i << ": " << host[i] << std::endl;
  }

  // GV/voxelmap/BitVoxelMap.hpp:57-81, 228-232
  virtual void clearBitVoxelMeaning(BitVoxelMeaning voxel_meaning) { clearBit(uint32_t(voxel_meaning)); }
  void clearBit(const uint32_t bit_index)
  {
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_map_clear_meaning(m_handle, int(bit_index)), "clearBit");
  }
  void clearBits(BitVector<length> bits)
  {
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_map_clear_bits(m_handle, bits.bytes()), "clearBits");
  }

  // GV/voxelmap/BitVoxelMap.hpp:191-203
  size_t collideWith(const BitVoxelMap<length>* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    return collideMapMap(map, coll_threshold, offset);
  }
  size_t collideWith(const ProbVoxelMap* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    return collideMapMap(map, coll_threshold, offset);
  }
  // GV/voxelmap/BitVoxelMap.hpp:91-135: swept-volume check with a bit window; colliding_meanings is OR-ed into
  uint32_t collisionCheckBitvector(const BitVoxelMap<length>* other, const SVCollider& collider,
                                   BitVector<length>& colliding_meanings, const uint16_t sv_offset = 0)
  {
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    uint64_t n = 0;
    if (!gvxCheck(gvx_collide_map_map_bitcheck(m_handle, other->getHandle(), uint32_t(uint8_t(collider.getWindowSize())),
                                               sv_offset, 1, colliding_meanings.bytes(), &n), "collisionCheckBitvector"))
      return UINT_MAX;
    return uint32_t(n);
  }
  // GV/voxelmap/BitVoxelMap.hpp:139-188 with a ProbVoxelMap: the SVCollider's threshold, no window
  uint32_t collisionCheckBitvector(const ProbVoxelMap* other, const SVCollider& collider,
                                   BitVector<length>& colliding_meanings, const uint16_t sv_offset = 0)
  {
    (void)sv_offset;
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    uint64_t n = 0;
    if (!gvxCheck(gvx_collide_map_map_types(m_handle, other->getHandle(), collider.getThreshold(), 1,
                                            colliding_meanings.bytes(), &n), "collisionCheckBitvector"))
      return UINT_MAX;
    return uint32_t(n);
  }
  // GV/voxelmap/BitVoxelMap.hpp:205-210: SVCollider(coll_threshold), window 0
  size_t collideWithTypes(const BitVoxelMap<length>* map, BitVoxel<length>& types_in_collision, float coll_threshold = 1.0,
                          const Vector3i& offset = Vector3i())
  {
    (void)offset;
    return collisionCheckBitvector(map, SVCollider(coll_threshold), types_in_collision.bitVector());
  }
  // GV/voxelmap/BitVoxelMap.hpp:234-246
  void shiftLeftSweptVolumeIDs(uint8_t shift_size)
  {
    if (shift_size > 56)
    {
      LOGGING_ERROR_C(VoxelmapLog, BitVoxelMap, "Maximum shift size is 56! Higher shift number requested. Not performing shift operation." << endl);
      return;
    }
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_map_shift_sv_ids(m_handle, shift_size), "shiftLeftSweptVolumeIDs");
  }

  // GV/voxelmap/BitVoxelMap.hpp:212-217: types_in_collision is OR-ed into; the offset is ignored by the reference too
  size_t collideWithTypes(const ProbVoxelMap* map, BitVoxel<length>& types_in_collision, float coll_threshold = 1.0,
                          const Vector3i& offset = Vector3i())
  {
    (void)offset;
    detail::DualLock guard(this->m_mutex, map->m_mutex);
    uint64_t n = 0;
    if (!gvxCheck(gvx_collide_map_map_types(m_handle, map->getHandle(), coll_threshold, 1,
                                            types_in_collision.bitVector().bytes(), &n), "collideWithTypes"))
      return SSIZE_MAX;
    return size_t(n);
  }
};

inline size_t ProbVoxelMap::collideWith(const BitVoxelMap<BIT_VECTOR_LENGTH>* map, float coll_threshold, const Vector3i& offset)
{
  return collideMapMap(map, coll_threshold, offset);
}

typedef BitVoxelMap<BIT_VECTOR_LENGTH> BitVectorVoxelMap;  // GV/voxelmap/VoxelMap.h

}  // namespace voxelmap
}  // namespace gpu_voxels
#endif
