// gpu_voxels/GpuVoxelsMap.h -- abstract base of all maps, as in the reference (GV/GpuVoxelsMap.h:50-276).
// std::shared_ptr / std::recursive_timed_mutex stand in for the boost types of the same role (Boost is not a
// dependency of this build); names of the typedefs are the reference's.
#ifndef GPU_VOXELS_B200_GPUVOXELSMAP_H
#define GPU_VOXELS_B200_GPUVOXELSMAP_H

#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <gpu_voxels/helpers/BoostCompat.h>
#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/MetaPointCloud.h>
#include <gpu_voxels/helpers/PointCloud.h>
#include <gpu_voxels/helpers/common_defines.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>
#include <gpu_voxels/logging/logging_gpu_voxels.h>
#include <gpu_voxels/voxel/BitVoxel.h>

namespace gpu_voxels {

class GpuVoxelsMap;
typedef std::shared_ptr<GpuVoxelsMap> GpuVoxelsMapSharedPtr;
typedef std::unique_lock<std::recursive_timed_mutex> lock_guard;

class GpuVoxelsMap
{
public:
  GpuVoxelsMap() : m_map_type(MT_BITVECTOR_VOXELMAP) {}
  virtual ~GpuVoxelsMap() {}

  template <class T> const T* as() const { return static_cast<const T*>(this); }
  template <class T> T* as() { return static_cast<T*>(this); }
  template <class T> bool is() { return dynamic_cast<T*>(this) != NULL; }
  template <class T> bool is() const { return dynamic_cast<const T*>(this) != NULL; }

  MapType getMapType() const { return m_map_type; }

  virtual void insertPointCloud(const std::vector<Vector3f>& point_cloud, const BitVoxelMeaning voxel_meaning) = 0;
  virtual void insertPointCloud(const PointCloud& pointcloud, const BitVoxelMeaning voxel_meaning) = 0;
  // GV/GpuVoxelsMap.h:117: no map of the path implements it (each logs "not supported" and returns false)
  virtual bool insertRobotConfiguration(const MetaPointCloud* robot_links, bool with_self_collision_test) = 0;
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud, BitVoxelMeaning voxel_meaning) = 0;
  virtual void insertMetaPointCloud(const MetaPointCloud& meta_point_cloud,
                                    const std::vector<BitVoxelMeaning>& voxel_meanings) = 0;
  // GV/GpuVoxelsMap.cpp:36-49: load (.xyz / .pcd / .binvox), then insertPointCloud; false if the file cannot be read
  bool insertPointCloudFromFile(const std::string path, const bool use_model_path, const BitVoxelMeaning voxel_meaning,
                                const bool shift_to_zero = false, const Vector3f& offset_XYZ = Vector3f(),
                                const float scaling = 1.0)
  {
    std::vector<Vector3f> points;
    if (!file_handling::PointcloudFileHandler::Instance()->loadPointCloud(path, use_model_path, points, shift_to_zero,
                                                                         offset_XYZ, scaling))
      return false;
    insertPointCloud(points, voxel_meaning);
    return true;
  }
  // GV/GpuVoxelsMap.h:186-204: copy the other map's data into this one; lists merge lists, dense maps refuse
  virtual bool merge(const GpuVoxelsMapSharedPtr other, const Vector3i& voxel_offset = Vector3i(),
                     const BitVoxelMeaning* new_meaning = NULL) = 0;
  virtual bool merge(const GpuVoxelsMapSharedPtr other, const Vector3f& metric_offset,
                     const BitVoxelMeaning* new_meaning = NULL) = 0;
  // the reference's binary dumps (GV/GpuVoxelsMap.h:206-216)
  virtual bool writeToDisk(const std::string path) = 0;
  virtual bool readFromDisk(const std::string path) = 0;
  void generateVisualizerData() {}  // GV/GpuVoxelsMap.cpp:50-52: empty there too; the visualizer is out of scope
  virtual void clearMap() = 0;
  virtual void clearBitVoxelMeaning(BitVoxelMeaning voxel_meaning) = 0;
  // GV/GpuVoxelsMap.h:229-247, .cpp:54-63: only octrees fragment; dense maps and lists answer "not supported" / false
  virtual bool needsRebuild() const = 0;
  virtual bool rebuild() = 0;
  virtual bool rebuildIfNeeded()
  {
    if (needsRebuild())
    {
      rebuild();
      return true;
    }
    return false;
  }
  virtual std::size_t getMemoryUsage() const = 0;
  virtual Vector3ui getDimensions() const = 0;
  virtual Vector3f getMetricDimensions() const = 0;

  // the C-ABI handle behind this map (not in the reference)
  virtual gvx_handle_t getHandle() const = 0;

  // public so that callers can lock a map around several operations, as in the reference (GV/GpuVoxelsMap.h:267-269)
  mutable std::recursive_timed_mutex m_mutex;

protected:
  MapType m_map_type;

private:
  GpuVoxelsMap(const GpuVoxelsMap&);
  const GpuVoxelsMap& operator=(const GpuVoxelsMap&);
};

namespace detail {
inline std::vector<int32_t> meaningsToInts(const std::vector<BitVoxelMeaning>& v)
{
  std::vector<int32_t> out(v.size());
  for (size_t i = 0; i < v.size(); ++i) out[i] = int32_t(v[i]);
  return out;
}
// lock two maps without deadlock, like boost::lock(a, b) in the reference's binary operations
struct DualLock
{
  DualLock(std::recursive_timed_mutex& a, std::recursive_timed_mutex& b) : m_a(a), m_b(b), m_same(&a == &b)
  {
    if (m_same) m_a.lock(); else std::lock(m_a, m_b);
  }
  ~DualLock() { m_a.unlock(); if (!m_same) m_b.unlock(); }
  std::recursive_timed_mutex& m_a; std::recursive_timed_mutex& m_b; bool m_same;
};
// One collision lane per WORKER thread (include/gvx.h "collision lanes"): the boolean collisionCheck of a thread other
// than the first one to ask runs on a stream and result block of its own, so that planner threads checking disjoint
// map pairs overlap on the GPU -- what the reference's per-map m_mutex allows (GV/voxelmap/TemplateVoxelMap.hpp:329-414).
// The first thread keeps the serial path; 0 = no lane (first thread, or lane creation failed).
struct ThreadLane
{
  gvx_handle_t handle;
  bool tried;
  ThreadLane() : handle(0), tried(false) {}
  ~ThreadLane() { if (handle) gvx_lane_destroy(handle); }
};
inline gvx_handle_t threadLane()
{
  static const std::thread::id first = std::this_thread::get_id();
  if (std::this_thread::get_id() == first) return 0;
  thread_local ThreadLane tl;
  if (!tl.tried)
  {
    tl.tried = true;
    if (gvx_lane_create(&tl.handle) != GVX_OK) tl.handle = 0;
  }
  return tl.handle;
}
}  // namespace detail

}  // namespace gpu_voxels
#endif
