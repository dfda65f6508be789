// gpu_voxels/helpers/MetaPointCloud.h -- MetaPointCloud of the reference's API (GV/helpers/MetaPointCloud.h:41-241,
// GV/helpers/MetaPointCloud.cu:24-140, 417-447, 596-649): several sub-clouds in ONE contiguous array plus their sizes.
// Host copy + one device buffer; the per-sub-cloud transform runs in libgvx.so.
#ifndef GPU_VOXELS_B200_HELPERS_METAPOINTCLOUD_H
#define GPU_VOXELS_B200_HELPERS_METAPOINTCLOUD_H

#include <cstring>
#include <iostream>
#include <map>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/PointCloud.h>
#include <gpu_voxels/helpers/PointcloudFileHandler.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>
#include <gpu_voxels/logging/logging_gpu_voxels.h>

namespace gpu_voxels {

class MetaPointCloud
{
public:
  MetaPointCloud() : m_buf(0), m_dev(NULL), m_cap(0) { gvxEnsureInit(); }
  // sub-clouds from .xyz / .pcd / .binvox files (GV/helpers/MetaPointCloud.cu:115-171)
  explicit MetaPointCloud(const std::vector<std::string>& point_cloud_files, bool use_model_path)
    : m_buf(0), m_dev(NULL), m_cap(0)
  {
    gvxEnsureInit();
    addClouds(point_cloud_files, use_model_path);
  }
  explicit MetaPointCloud(const std::vector<std::string>& point_cloud_files,
                          const std::vector<std::string>& point_cloud_names, bool use_model_path)
    : m_buf(0), m_dev(NULL), m_cap(0)
  {
    gvxEnsureInit();
    addClouds(point_cloud_files, use_model_path);
    if (point_cloud_files.size() == point_cloud_names.size())
    {
      if (m_sizes.size() == point_cloud_names.size())
        for (size_t i = 0; i < point_cloud_names.size(); ++i) m_names[uint16_t(i)] = point_cloud_names[i];
    }
    else
      std::cerr << "<Error MetaPointCloud> Number of names differs to number of pointcloud files!" << std::endl;
  }
  // all files or none: the first unreadable file aborts and leaves the MetaPointCloud as it was
  void addClouds(const std::vector<std::string>& point_cloud_files, bool use_model_path)
  {
    std::vector<std::vector<Vector3f> > clouds(point_cloud_files.size());
    for (size_t i = 0; i < point_cloud_files.size(); ++i)
      if (!file_handling::PointcloudFileHandler::Instance()->loadPointCloud(point_cloud_files[i], use_model_path, clouds[i]))
      {
        std::cerr << "<Error MetaPointCloud> Could not read file " << point_cloud_files[i] << std::endl;
        return;
      }
    for (size_t i = 0; i < clouds.size(); ++i) addCloud(clouds[i], false);
    syncToDevice();
  }
  explicit MetaPointCloud(const std::vector<uint32_t>& point_cloud_sizes) : m_buf(0), m_dev(NULL), m_cap(0)
  {
    gvxEnsureInit();
    for (size_t i = 0; i < point_cloud_sizes.size(); ++i) addCloud(point_cloud_sizes[i]);
  }
  explicit MetaPointCloud(const std::vector<std::vector<Vector3f> >& point_clouds) : m_buf(0), m_dev(NULL), m_cap(0)
  {
    gvxEnsureInit();
    for (size_t i = 0; i < point_clouds.size(); ++i) addCloud(point_clouds[i], false);
    syncToDevice();
  }
  MetaPointCloud(const MetaPointCloud& other) : m_buf(0), m_dev(NULL), m_cap(0) { *this = other; }
  MetaPointCloud& operator=(const MetaPointCloud& other)
  {
    if (this == &other) return *this;
    m_sizes = other.m_sizes; m_points = other.m_points; m_names = other.m_names;
    syncToDevice();
    return *this;
  }
  ~MetaPointCloud() { if (m_buf) gvx_buffer_destroy(m_buf); }

  // GV/helpers/MetaPointCloud.cu:226-287: same accumulated size, same NUMBER of sub-clouds and the same bytes of the
  // accumulated array (the individual sub-cloud sizes are not compared there; NaN bit patterns compare equal)
  bool operator==(const MetaPointCloud& other) const
  {
    if (this == &other) return true;
    if (m_points.size() != other.m_points.size() || m_sizes.size() != other.m_sizes.size()) return false;
    return m_points.empty() || std::memcmp(&m_points[0], &other.m_points[0], m_points.size() * sizeof(Vector3f)) == 0;
  }

  void addCloud(uint32_t cloud_size)
  {
    m_sizes.push_back(cloud_size);
    m_points.resize(m_points.size() + cloud_size);
  }
  void addCloud(const std::vector<Vector3f>& cloud, bool sync = false, const std::string& name = "")
  {
    addCloud(cloud.empty() ? NULL : &cloud[0], uint32_t(cloud.size()), sync, name);
  }
  void addCloud(const Vector3f* points, uint32_t pointcloud_size, bool sync = false, const std::string& name = "")
  {
    m_sizes.push_back(pointcloud_size);
    m_points.insert(m_points.end(), points, points + pointcloud_size);
    if (!name.empty()) m_names[uint16_t(m_sizes.size() - 1)] = name;
    if (sync) syncToDevice();
  }
  void addCloud(const PointCloud& cloud, bool sync = false, const std::string& name = "")
  {
    addCloud(cloud.getPoints(), uint32_t(cloud.getPointCloudSize()), sync, name);
  }
  std::string getCloudName(uint16_t i) const
  {
    std::map<uint16_t, std::string>::const_iterator it = m_names.find(i);
    return it == m_names.end() ? std::string() : it->second;
  }
  const std::map<uint16_t, std::string> getCloudNames() const { return m_names; }
  int16_t getCloudNumber(const std::string& name) const
  {
    for (std::map<uint16_t, std::string>::const_iterator it = m_names.begin(); it != m_names.end(); ++it)
      if (it->second == name) return int16_t(it->first);
    return -1;
  }

  // sizes, names and bounding boxes of the sub-clouds on stdout (GV/helpers/MetaPointCloud.cu:543-590)
  void debugPointCloud() const
  {
    std::cout << "================== hostMetaPointCloud DBG ==================" << std::endl;
    std::cout << "hostDebugMetaPointCloud DBG: NumClouds: " << m_sizes.size() << std::endl;
    size_t first = 0;
    for (size_t c = 0; c < m_sizes.size(); first += m_sizes[c], ++c)
    {
      std::cout << "hostDebugMetaPointCloud DBG: '" << getCloudName(uint16_t(c)) << "' CloudSize[" << c << "]: "
                << m_sizes[c] << std::endl;
      if (m_sizes[c] == 0) continue;
      Vector3f lo = m_points[first], hi = m_points[first];
      for (size_t j = first + 1; j < first + m_sizes[c]; ++j)
      {
        const Vector3f& p = m_points[j];
        lo.x = p.x < lo.x ? p.x : lo.x; lo.y = p.y < lo.y ? p.y : lo.y; lo.z = p.z < lo.z ? p.z : lo.z;
        hi.x = hi.x < p.x ? p.x : hi.x; hi.y = hi.y < p.y ? p.y : hi.y; hi.z = hi.z < p.z ? p.z : hi.z;
      }
      std::cout << "hostDebugMetaPointCloud DBG: CloudSize[" << c << "] bounds: Min[" << lo.x << ", " << lo.y << ", "
                << lo.z << "], Max[" << hi.x << ", " << hi.y << ", " << hi.z << "]" << std::endl;
    }
    std::cout << "================== END hostMetaPointCloud DBG ==================" << std::endl;
  }

  void syncToDevice()
  {
    reserveDevice(m_points.size());
    if (!m_points.empty())
      gvxCheck(gvx_buffer_upload(m_buf, 0, &m_points[0], m_points.size() * sizeof(Vector3f)), "MetaPointCloud::syncToDevice");
  }
  void syncToDevice(uint16_t cloud)
  {
    if (m_points.size() > m_cap || !m_buf) { syncToDevice(); return; }
    const size_t off = offsetOf(cloud);
    if (m_sizes[cloud])
      gvxCheck(gvx_buffer_upload(m_buf, off * sizeof(Vector3f), &m_points[off], m_sizes[cloud] * sizeof(Vector3f)),
               "MetaPointCloud::syncToDevice");
  }
  void syncToHost()
  {
    if (!m_points.empty() && m_buf)
      gvxCheck(gvx_buffer_download(m_buf, 0, &m_points[0], m_points.size() * sizeof(Vector3f)), "MetaPointCloud::syncToHost");
  }

  // same-size update keeps the storage; a different size re-packs the contiguous array (the reference's slow path)
  void updatePointCloud(uint16_t cloud, const Vector3f* pointcloud, uint32_t pointcloud_size, bool sync = false)
  {
    if (cloud >= m_sizes.size()) { LOGGING_ERROR(Gpu_voxels_helpers, "Cloud " << cloud << " does not exist" << endl); return; }
    const size_t off = offsetOf(cloud);
    if (pointcloud_size != m_sizes[cloud])
    {
      m_points.erase(m_points.begin() + off, m_points.begin() + off + m_sizes[cloud]);
      m_points.insert(m_points.begin() + off, pointcloud, pointcloud + pointcloud_size);
      m_sizes[cloud] = pointcloud_size;
      if (sync) syncToDevice();
      return;
    }
    std::copy(pointcloud, pointcloud + pointcloud_size, m_points.begin() + off);
    if (sync) syncToDevice(cloud);
  }
  void updatePointCloud(uint16_t cloud, const std::vector<Vector3f>& pointcloud, bool sync = false)
  {
    updatePointCloud(cloud, pointcloud.empty() ? NULL : &pointcloud[0], uint32_t(pointcloud.size()), sync);
  }
  void updatePointCloud(uint16_t cloud, const PointCloud& pointcloud, bool sync = false)
  {
    updatePointCloud(cloud, pointcloud.getPoints(), uint32_t(pointcloud.getPointCloudSize()), sync);
  }
  void updatePointCloud(const std::string& cloud_name, const std::vector<Vector3f>& pointcloud, bool sync = false)
  {
    const int16_t c = getCloudNumber(cloud_name);
    if (c < 0) { LOGGING_ERROR(Gpu_voxels_helpers, "Cloud '" << cloud_name << "' does not exist" << endl); return; }
    updatePointCloud(uint16_t(c), pointcloud, sync);
  }

  uint16_t getNumberOfPointclouds() const { return uint16_t(m_sizes.size()); }
  uint32_t getPointcloudSize(uint16_t cloud = 0) const { return m_sizes[cloud]; }
  uint32_t getAccumulatedPointcloudSize() const { return uint32_t(m_points.size()); }
  const std::vector<uint32_t>& getPointcloudSizes() const { return m_sizes; }
  const Vector3f* getPointCloud(uint16_t cloud) const { return m_points.empty() ? NULL : &m_points[offsetOf(cloud)]; }
  // the contiguous device array of all sub-clouds (the reference hands out a MetaPointCloudStruct*; kernels here take
  // the base pointer + sizes by value instead of chasing three dependent device pointers)
  const Vector3f* getDeviceBasePointer() const
  {
    if (!m_buf || m_cap < m_points.size()) const_cast<MetaPointCloud*>(this)->syncToDevice();  // never synced yet
    return m_dev;
  }

  // MetaPointCloud::transform / transformSubCloud / transformSelf / transformSelfSubCloud
  // (GV/helpers/MetaPointCloud.cu:596-649): device data is transformed; the host copy follows on syncToHost()
  void transform(const Matrix4f* transformation, MetaPointCloud* transformed_cloud) const
  {
    if (!prepareTarget(transformed_cloud) || m_points.empty()) return;
    gvxCheck(gvx_transform_points((const float*)m_dev, (float*)transformed_cloud->m_dev, m_points.size(), GVX_MEM_DEVICE,
                                  transformation->data()), "MetaPointCloud::transform");
  }
  void transformSubCloud(uint8_t subcloud_to_transform, const Matrix4f* transformation, MetaPointCloud* transformed_cloud) const
  {
    if (subcloud_to_transform >= m_sizes.size() || !prepareTarget(transformed_cloud)) return;
    const size_t off = offsetOf(subcloud_to_transform);
    if (!m_sizes[subcloud_to_transform]) return;
    gvxCheck(gvx_transform_points((const float*)(m_dev + off), (float*)(transformed_cloud->m_dev + off),
                                  m_sizes[subcloud_to_transform], GVX_MEM_DEVICE, transformation->data()),
             "MetaPointCloud::transformSubCloud");
  }
  void transformSelf(const Matrix4f* transformation) { transform(transformation, this); }
  void transformSelfSubCloud(uint8_t subcloud_to_transform, const Matrix4f* transformation)
  {
    transformSubCloud(subcloud_to_transform, transformation, this);
  }

private:
  size_t offsetOf(uint16_t cloud) const
  {
    size_t off = 0;
    for (uint16_t c = 0; c < cloud; ++c) off += m_sizes[c];
    return off;
  }
  bool prepareTarget(MetaPointCloud* t) const
  {
    if (t == this) return m_buf != 0 || m_points.empty();
    if (t->m_sizes != m_sizes)
    {
      LOGGING_ERROR(Gpu_voxels_helpers, "Size of target pointcloud does not match local pointcloud. Not transforming!" << endl);
      return false;
    }
    if (!t->m_buf) t->syncToDevice();
    return true;
  }
  void reserveDevice(size_t n)
  {
    if (n <= m_cap && m_buf) return;
    if (m_buf) gvx_buffer_destroy(m_buf);
    m_buf = 0; m_dev = NULL; m_cap = 0;
    if (!gvxCheck(gvx_buffer_create(&m_buf, n * sizeof(Vector3f)), "MetaPointCloud allocation")) return;
    void* p = NULL;
    gvx_buffer_device_ptr(m_buf, &p, NULL);
    m_dev = static_cast<Vector3f*>(p);
    m_cap = n;
  }

  std::vector<uint32_t> m_sizes;
  std::vector<Vector3f> m_points;
  std::map<uint16_t, std::string> m_names;
  gvx_handle_t m_buf;
  Vector3f* m_dev;
  size_t m_cap;
};

}  // namespace gpu_voxels
#endif
