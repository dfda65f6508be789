// gpu_voxels/voxellist/VoxelList.h -- voxellist::BitVectorVoxelList of the reference's API over the C ABI
// (GV/voxellist/BitVoxelList.h:139-150, .hpp:70-265; GV/voxellist/TemplateVoxelList.h, .hpp:284-422, 731-774).
#ifndef GPU_VOXELS_B200_VOXELLIST_VOXELLIST_H
#define GPU_VOXELS_B200_VOXELLIST_VOXELLIST_H

#include <climits>
#include <cmath>
#include <iostream>

#include <gpu_voxels/GpuVoxelsMap.h>
#include <gpu_voxels/voxelmap/VoxelMap.h>

namespace gpu_voxels {
namespace voxellist {

template <std::size_t length>
class BitVoxelList : public GpuVoxelsMap
{
public:
  BitVoxelList(const Vector3ui ref_map_dim, const float voxel_side_length, const MapType map_type = MT_BITVECTOR_VOXELLIST)
    : m_ref_map_dim(ref_map_dim), m_voxel_side_length(voxel_side_length), m_handle(0)
  {
    gvxEnsureInit();
    this->m_map_type = map_type;
    gvxCheck(gvx_list_create(&m_handle, ref_map_dim.x, ref_map_dim.y, ref_map_dim.z, voxel_side_length), "voxel list construction");
  }
  virtual ~BitVoxelList() { if (m_handle) gvx_list_destroy(m_handle); }

  virtual void insertPointCloud(const std::vector<Vector3f>& points, const BitVoxelMeaning voxel_meaning)
  {
    lock_guard guard(this->m_mutex);
    if (points.empty()) return;
    gvxCheck(gvx_list_insert_points(m_handle, reinterpret_cast<const float*>(&points[0]), points.size(), GVX_MEM_HOST,
                                    int(voxel_meaning)), "insertPointCloud");
  }
  virtual void insertPointCloud(const PointCloud& pointcloud, const BitVoxelMeaning voxel_meaning)
  {
    lock_guard guard(this->m_mutex);
    if (!pointcloud.getPointCloudSize()) return;
    gvxCheck(gvx_list_insert_points(m_handle, reinterpret_cast<const float*>(pointcloud.getConstDevicePointer()),
                                    pointcloud.getPointCloudSize(), GVX_MEM_DEVICE, int(voxel_meaning)), "insertPointCloud");
  }
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud, BitVoxelMeaning voxel_meaning)
  {
    std::vector<BitVoxelMeaning> one(1, voxel_meaning);
    insertMeta(meta_point_cloud, one);
  }
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud,
                                    const std::vector<BitVoxelMeaning>& voxel_meanings)
  {
    if (voxel_meanings.size() != meta_point_cloud.getNumberOfPointclouds())
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "Number of voxel meanings differs from number of sub-clouds" << endl);
      return;
    }
    insertMeta(meta_point_cloud, voxel_meanings);
  }
  // TemplateVoxelList::insertPointCloud(const Vector3f* points_d, uint32_t size, ...) (GV/voxellist/TemplateVoxelList.hpp:311-340)
  virtual void insertPointCloud(const Vector3f* points_d, uint32_t size, const BitVoxelMeaning voxel_meaning)
  {
    lock_guard guard(this->m_mutex);
    if (!size) return;
    gvxCheck(gvx_list_insert_points(m_handle, reinterpret_cast<const float*>(points_d), size, GVX_MEM_DEVICE,
                                    int(voxel_meaning)), "insertPointCloud");
  }
  // GV/voxellist/BitVoxelList.hpp:57-62, AbstractVoxelList.cu:26-36
  virtual bool insertRobotConfiguration(const MetaPointCloud*, bool)
  {
    LOGGING_ERROR_C(VoxellistLog, BitVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_YET_SUPPORTED << endl);
    return false;
  }
  virtual bool needsRebuild() const
  {
    LOGGING_ERROR_C(VoxellistLog, AbstractVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
    return false;
  }
  virtual bool rebuild()
  {
    LOGGING_ERROR_C(VoxellistLog, AbstractVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
    return false;
  }
  virtual MapType getTemplateType() { return this->m_map_type; }
  // the three device arrays of the list (GV/voxellist/TemplateVoxelList.h:95-116); they move when the list grows
  BitVoxel<length>* getDeviceDataPtr() { void* p = NULL; gvx_list_device_ptrs(m_handle, NULL, NULL, &p, NULL); return static_cast<BitVoxel<length>*>(p); }
  uint32_t* getDeviceIdPtr() { void* p = NULL; gvx_list_device_ptrs(m_handle, &p, NULL, NULL, NULL); return static_cast<uint32_t*>(p); }
  Vector3ui* getDeviceCoordPtr() { void* p = NULL; gvx_list_device_ptrs(m_handle, NULL, &p, NULL, NULL); return static_cast<Vector3ui*>(p); }
  virtual void* getVoidDeviceDataPtr() { return getDeviceDataPtr(); }
  // entries are unique and sorted after every insert / merge here, and storage is sized exactly: nothing to do
  void make_unique() {}
  virtual void shrinkToFit() {}
  // TemplateVoxelList::screendump (GV/voxellist/TemplateVoxelList.hpp:776-802)
  virtual void screendump(bool with_voxel_content = true)
  {
    LOGGING_INFO_C(VoxellistLog, TemplateVoxelList, "Dumping VoxelList with " << getDimensions().x << " entries to screen: " << endl);
    lock_guard guard(this->m_mutex);
    std::vector<uint32_t> ids; std::vector<Vector3ui> coords; std::vector<BitVectorVoxel> voxels;
    copyToHost(ids, coords, voxels);
    for (std::size_t i = 0; i < ids.size(); ++i)
    {
      std::cout << "[" << i << "] ID = " << ids[i] << " Coords: " << coords[i];
      if (with_voxel_content) std::cout << "Voxel: " << voxels[i];
      std::cout << std::endl;
    }
    LOGGING_INFO_C(VoxellistLog, TemplateVoxelList, "Dumped " << ids.size() << " Voxels." << endl);
  }
  // TemplateVoxelList::collisionCheckWithCollider (GV/voxellist/TemplateVoxelList.hpp:690-774) with the default
  // collider: what collideWith forwards to. The list's voxels are BitVoxels (side 1 of the collider), so of a
  // DefaultCollider(t1, t2) only t2 -- the test of the map's ProbabilisticVoxel -- matters
  // (DefaultCollider::collide(BitVoxel, ProbabilisticVoxel), GV/voxel/DefaultCollider.hpp:66-72)
  size_t collisionCheckWithCollider(const voxelmap::AbstractVoxelMap* other, DefaultCollider collider = DefaultCollider(),
                                    const Vector3i& offset = Vector3i())
  {
    float thr = 0.f;
    if (!DefaultCollider(collider.getThreshold2(), collider.getThreshold2()).asFloatThreshold(&thr))
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "DefaultCollider threshold without a float form: " << GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
      return 0;
    }
    return collide(0, other, thr, offset, NULL, NULL);
  }
  // list x list: both sides are BitVoxels, the thresholds play no role (DefaultCollider::collide(BitVoxel, BitVoxel))
  size_t collisionCheckWithCollider(const BitVoxelList<length>* other, DefaultCollider = DefaultCollider(),
                                    const Vector3i& offset = Vector3i())
  {
    return collideWith(other, 1.0f, offset);
  }
  virtual void clearMap()
  {
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_list_clear(m_handle), "clearMap");
  }
  virtual void clearBitVoxelMeaning(BitVoxelMeaning)
  {
    LOGGING_ERROR_C(VoxellistLog, BitVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
  }
  // the reference reports (number of entries, 1, 1) (GV/voxellist/TemplateVoxelList.hpp:418-422)
  virtual Vector3ui getDimensions() const { return Vector3ui(uint32_t(size()), 1, 1); }
  virtual Vector3f getMetricDimensions() const
  {
    return Vector3f(float(m_ref_map_dim.x), float(m_ref_map_dim.y), float(m_ref_map_dim.z)) * m_voxel_side_length;
  }
  virtual std::size_t getMemoryUsage() const { return size() * (4 + 12 + 32); }
  virtual gvx_handle_t getHandle() const { return m_handle; }
  float getVoxelSideLength() const { return m_voxel_side_length; }
  std::size_t size() const { uint64_t n = 0; gvx_list_size(m_handle, &n); return std::size_t(n); }
  // copyCoordsToHost (GV/voxellist/TemplateVoxelList.hpp:444-471)
  void copyCoordsToHost(std::vector<Vector3ui>& host_vec)
  {
    host_vec.resize(size());
    if (!host_vec.empty()) gvxCheck(gvx_list_download(m_handle, NULL, &host_vec[0].x, NULL), "copyCoordsToHost");
  }

  // ids / coords / voxels of all entries (what extractCubes exposes in the reference, GV/voxellist/TemplateVoxelList.hpp)
  void copyToHost(std::vector<uint32_t>& ids, std::vector<Vector3ui>& coords, std::vector<BitVectorVoxel>& voxels)
  {
    const std::size_t n = size();
    ids.resize(n); coords.resize(n); voxels.resize(n);
    if (n) gvxCheck(gvx_list_download(m_handle, &ids[0], &coords[0].x, voxels[0].bitVector().bytes()), "copyToHost");
  }

  // GV/voxellist/BitVoxelList.hpp:70-84
  size_t collideWith(const voxelmap::ProbVoxelMap* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    return collide(0, map, coll_threshold, offset, NULL, NULL);
  }
  size_t collideWith(const voxelmap::BitVectorVoxelMap* map, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    return collide(0, map, coll_threshold, offset, NULL, NULL);
  }
  // GV/voxellist/BitVoxelList.hpp:127-218
  size_t collideWithTypes(const voxelmap::ProbVoxelMap* map, BitVectorVoxel& types_in_collision, float coll_threshold = 1.0,
                          const Vector3i& offset = Vector3i())
  {
    return collide(1, map, coll_threshold, offset, NULL, types_in_collision.bitVector().bytes());
  }
  size_t collideWithTypes(const voxelmap::BitVectorVoxelMap* map, BitVectorVoxel& types_in_collision, float coll_threshold = 1.0,
                          const Vector3i& offset = Vector3i())
  {
    return collide(1, map, coll_threshold, offset, NULL, types_in_collision.bitVector().bytes());
  }
  // GV/voxellist/BitVoxelList.hpp:220-265
  template <class MapT>
  size_t collideWithTypeMask(const MapT* map, const BitVectorVoxel& types_to_check, float coll_threshold = 1.0,
                             const Vector3i& offset = Vector3i())
  {
    return collide(2, map, coll_threshold, offset, types_to_check.bitVector().bytes(), NULL);
  }
  // list x map per-meaning counts with the semantics of collideCountingPerMeaning (GV/voxellist/BitVoxelList.hpp:300-344)
  template <class MapT>
  size_t collideCountingPerMeaning(const MapT* map, std::vector<size_t>& collisions_per_meaning, float coll_threshold = 1.0,
                                   const Vector3i& offset = Vector3i())
  {
    detail::DualLock guard(this->m_mutex, map->m_mutex);
    const int32_t off[3] = {offset.x, offset.y, offset.z};
    uint64_t counts[256], total = 0;
    const int rc = gvx_collide_list_map_per_meaning(m_handle, map->getHandle(), coll_threshold, off, 1, counts, &total);
    if (rc != GVX_OK) { gvxCheck(rc, "collideCountingPerMeaning"); return SSIZE_MAX; }
    collisions_per_meaning.assign(counts, counts + 256);
    return size_t(total);
  }

  // ---- list x list (GV/voxellist/BitVoxelList.hpp:86-125, 267-344; offsets between lists are refused there) ----
  size_t collideWith(const BitVoxelList<length>* other, float coll_threshold = 1.0, const Vector3i& offset = Vector3i())
  {
    (void)coll_threshold;
    uint64_t n = 0;
    const int rc = gvx_collide_list_list(m_handle, other->m_handle, &n);
    return listList(other, offset, rc, n);
  }
  size_t collideWithTypes(const BitVoxelList<length>* other, BitVectorVoxel& types_in_collision, float coll_threshold = 1.0,
                          const Vector3i& offset = Vector3i())
  {
    (void)coll_threshold;
    uint64_t n = 0;
    const int rc = gvx_collide_list_list_types(m_handle, other->m_handle, types_in_collision.bitVector().bytes(), &n);
    return listList(other, offset, rc, n);
  }
  size_t collideWithBitcheck(const BitVoxelList<length>* other, const uint8_t margin = 0, const Vector3i& offset = Vector3i())
  {
    uint64_t n = 0;
    const int rc = gvx_collide_list_list_bitcheck(m_handle, other->m_handle, margin, &n);
    return listList(other, offset, rc, n);
  }
  size_t collideCountingPerMeaning(const GpuVoxelsMapSharedPtr other, std::vector<size_t>& collisions_per_meaning,
                                   const Vector3i& offset = Vector3i())
  {
    if (other->getMapType() != MT_BITVECTOR_VOXELLIST)
    {
      LOGGING_ERROR_C(VoxellistLog, BitVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED << endl);
      return SSIZE_MAX;
    }
    uint64_t counts[256], total = 0;
    const int rc = gvx_collide_list_list_per_meaning(m_handle, other->getHandle(), counts, &total);
    const size_t r = listList(other.get(), offset, rc, total);
    if (r != SSIZE_MAX && offset == Vector3i())
    {
      if (collisions_per_meaning.size() < 256) collisions_per_meaning.resize(256, 0);
      for (int k = 0; k < 256; ++k) collisions_per_meaning[k] += size_t(counts[k]);  // the reference increments
    }
    return r;
  }
  // GV/voxellist/TemplateVoxelList.hpp:609-641
  bool subtract(const BitVoxelList<length>* other, const Vector3i& voxel_offset = Vector3i())
  {
    if (!(voxel_offset == Vector3i()))
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "Offset for VoxelList collision not supported! Not performing collision check." << endl);
      return true;  // the reference removes nothing and still returns true
    }
    detail::DualLock guard(this->m_mutex, other->m_mutex);
    return gvxCheck(gvx_list_subtract(m_handle, other->m_handle), "subtract");
  }
  bool subtract(const BitVoxelList<length>* other, const Vector3f& metric_offset)
  {
    return subtract(other, Vector3i(int32_t(std::floor(metric_offset.x / m_voxel_side_length)),
                                    int32_t(std::floor(metric_offset.y / m_voxel_side_length)),
                                    int32_t(std::floor(metric_offset.z / m_voxel_side_length))));
  }
  // GV/voxellist/TemplateVoxelList.hpp:548-608: only another BitVectorVoxelList can be merged; its entries are
  // appended (shifted by voxel_offset, voxels replaced by *new_meaning if given) and the list is made unique again
  virtual bool merge(const GpuVoxelsMapSharedPtr other, const Vector3i& voxel_offset = Vector3i(),
                     const BitVoxelMeaning* new_meaning = NULL)
  {
    if (!other || other->getMapType() != MT_BITVECTOR_VOXELLIST)
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, GPU_VOXELS_MAP_OPERATION_NOT_YET_SUPPORTED << endl);
      return false;
    }
    const BitVoxelList<length>* o = static_cast<const BitVoxelList<length>*>(other.get());
    detail::DualLock guard(this->m_mutex, o->m_mutex);
    const int32_t off[3] = {voxel_offset.x, voxel_offset.y, voxel_offset.z};
    return gvxCheck(gvx_list_merge(m_handle, o->m_handle, off, new_meaning ? int(*new_meaning) : -1), "merge");
  }
  virtual bool merge(const GpuVoxelsMapSharedPtr other, const Vector3f& metric_offset, const BitVoxelMeaning* new_meaning = NULL)
  {
    return merge(other, Vector3i(int32_t(std::floor(metric_offset.x / m_voxel_side_length)),
                                 int32_t(std::floor(metric_offset.y / m_voxel_side_length)),
                                 int32_t(std::floor(metric_offset.z / m_voxel_side_length))), new_meaning);
  }
  // GV/voxellist/TemplateVoxelList.hpp:806-825
  bool equals(const BitVoxelList<length>& other) const
  {
    int eq = 0;
    gvxCheck(gvx_list_equals(m_handle, other.m_handle, &eq), "equals");
    return eq != 0;
  }
  // GV/voxellist/BitVoxelList.hpp:560-582
  void shiftLeftSweptVolumeIDs(uint8_t shift_size)
  {
    if (shift_size > 56)
    {
      LOGGING_ERROR_C(VoxellistLog, BitVoxelList, "Maximum shift size is 56! Higher shift number requested. Not performing shift operation." << endl);
      return;
    }
    lock_guard guard(this->m_mutex);
    gvxCheck(gvx_list_shift_sv_ids(m_handle, shift_size), "shiftLeftSweptVolumeIDs");
  }
  // GV/voxellist/TemplateVoxelList.hpp:444-526
  virtual bool writeToDisk(const std::string path)
  {
    lock_guard guard(this->m_mutex);
    return gvxCheck(gvx_list_write_file(m_handle, path.c_str()), "writeToDisk");
  }
  virtual bool readFromDisk(const std::string path)
  {
    lock_guard guard(this->m_mutex);
    return gvxCheck(gvx_list_read_file(m_handle, path.c_str()), "readFromDisk");
  }

private:
  size_t listList(const GpuVoxelsMap* other, const Vector3i& offset, int rc, uint64_t n)
  {
    (void)other;
    if (!(offset == Vector3i()))
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "Offset for VoxelList collision not supported! Not performing collision check." << endl);
      return 0;
    }
    if (rc == GVX_ERR_DIM_MISMATCH)
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "The Voxellist reference map dimensions differ!" << endl);
      return SSIZE_MAX;
    }
    if (!gvxCheck(rc, "list x list")) return SSIZE_MAX;
    return size_t(n);
  }
  size_t collide(int mode, const GpuVoxelsMap* map, float thr, const Vector3i& offset, const uint8_t* mask_in, uint8_t* mask_out)
  {
    detail::DualLock guard(this->m_mutex, map->m_mutex);
    const int32_t off[3] = {offset.x, offset.y, offset.z};
    uint64_t n = 0;
    int rc;
    if (mode == 0) rc = gvx_collide_list_map(m_handle, map->getHandle(), thr, off, 1, &n);
    else if (mode == 1) rc = gvx_collide_list_map_types(m_handle, map->getHandle(), thr, off, 1, mask_out, &n);
    else rc = gvx_collide_list_map_mask(m_handle, map->getHandle(), thr, off, 1, mask_in, &n);
    if (rc == GVX_ERR_DIM_MISMATCH)  // GV/voxellist/TemplateVoxelList.hpp:737-742
    {
      LOGGING_ERROR_C(VoxellistLog, TemplateVoxelList, "The Voxellist reference map dimension differ from the map dimension!" << endl);
      return SSIZE_MAX;
    }
    if (!gvxCheck(rc, "collideWith")) return SSIZE_MAX;
    return size_t(n);
  }
  void insertMeta(const MetaPointCloud& mpc, const std::vector<BitVoxelMeaning>& meanings)
  {
    lock_guard guard(this->m_mutex);
    if (mpc.getNumberOfPointclouds() == 0 || mpc.getAccumulatedPointcloudSize() == 0) return;
    const std::vector<int32_t> mg = detail::meaningsToInts(meanings);
    gvxCheck(gvx_list_insert_meta(m_handle, reinterpret_cast<const float*>(mpc.getDeviceBasePointer()), GVX_MEM_DEVICE,
                                  &mpc.getPointcloudSizes()[0], mpc.getNumberOfPointclouds(), &mg[0], uint32_t(mg.size())),
             "insertMetaPointCloud");
  }

  Vector3ui m_ref_map_dim;
  float m_voxel_side_length;
  gvx_handle_t m_handle;
};

typedef BitVoxelList<BIT_VECTOR_LENGTH> BitVectorVoxelList;  // GV/voxellist/VoxelList.h

}  // namespace voxellist
}  // namespace gpu_voxels
#endif
