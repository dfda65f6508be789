// gpu_voxels/helpers/PointCloud.h -- PointCloud of the reference's API (GV/helpers/PointCloud.h:40-155,
// GV/helpers/PointCloud.cu:32-346): a host copy plus a device copy of n packed Vector3f, kept in sync.
// Device storage is a gvx buffer; transforms run in libgvx.so (kernelTransformCloud's replacement).
#ifndef GPU_VOXELS_B200_HELPERS_POINTCLOUD_H
#define GPU_VOXELS_B200_HELPERS_POINTCLOUD_H

#include <cstring>
#include <iostream>
#include <string>
#include <vector>

#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/PointcloudFileHandler.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>

namespace gpu_voxels {

class PointCloud
{
public:
  PointCloud() : m_buf(0), m_dev(NULL), m_cap(0) { gvxEnsureInit(); }
  explicit PointCloud(const std::vector<Vector3f>& points) : m_buf(0), m_dev(NULL), m_cap(0) { gvxEnsureInit(); update(points); }
  explicit PointCloud(const Vector3f* points, uint32_t size) : m_buf(0), m_dev(NULL), m_cap(0) { gvxEnsureInit(); update(points, size); }
  // from a .xyz / .pcd / .binvox file (GV/helpers/PointCloud.cu:84-105); an unreadable file leaves the cloud empty
  explicit PointCloud(const std::string& path_to_file, bool use_model_path = true) : m_buf(0), m_dev(NULL), m_cap(0)
  {
    gvxEnsureInit();
    std::vector<Vector3f> points;
    if (!file_handling::PointcloudFileHandler::Instance()->loadPointCloud(path_to_file, use_model_path, points))
    {
      std::cerr << "<Error PointCloud> Could not read file " << path_to_file << std::endl;
      return;
    }
    update(points);
  }
  PointCloud(const PointCloud& other) : m_buf(0), m_dev(NULL), m_cap(0) { update(&other); }
  PointCloud& operator=(const PointCloud& other) { if (this != &other) update(&other); return *this; }
  ~PointCloud() { if (m_buf) gvx_buffer_destroy(m_buf); }

  // equal = same number of points and the same BYTES (GV/helpers/PointCloud.cu:147-210 compares the device arrays with
  // kernelCompareMem): NaN coordinates with equal bit patterns are equal, +0 and -0 are not
  bool operator==(const PointCloud& other) const
  {
    if (this == &other) return true;
    if (m_host.size() != other.m_host.size()) return false;
    return m_host.empty() || std::memcmp(&m_host[0], &other.m_host[0], m_host.size() * sizeof(Vector3f)) == 0;
  }

  void add(const PointCloud* cloud) { add(cloud->m_host); }
  void add(const std::vector<Vector3f>& points) { add(points.empty() ? NULL : &points[0], uint32_t(points.size())); }
  void add(const Vector3f* points, uint32_t size)
  {
    std::vector<Vector3f> all(m_host);
    all.insert(all.end(), points, points + size);
    update(all);
  }
  void update(const PointCloud* cloud) { update(cloud->m_host); }
  void update(const std::vector<Vector3f>& points) { update(points.empty() ? NULL : &points[0], uint32_t(points.size())); }
  void update(const Vector3f* points, uint32_t size)
  {
    m_host.assign(points, points + size);
    toDevice();
  }

  // PointCloud::transformSelf / transform (GV/helpers/PointCloud.cu:258-287)
  void transformSelf(const Matrix4f* transform) { this->transform(transform, this); }
  void transform(const Matrix4f* transform, PointCloud* transformed_cloud)
  {
    if (transformed_cloud != this) transformed_cloud->resize(uint32_t(m_host.size()));
    if (m_host.empty()) return;
    gvxCheck(gvx_transform_points((const float*)m_dev, (float*)transformed_cloud->m_dev, m_host.size(), GVX_MEM_DEVICE,
                                  transform->data()), "PointCloud::transform");
    transformed_cloud->toHost();
  }
  void scaleSelf(const Vector3f* scaling) { scale(scaling, this); }
  void scale(const Vector3f* scaling, PointCloud* scaled_cloud)
  {
    const Matrix4f s(scaling->x, 0, 0, 0, 0, scaling->y, 0, 0, 0, 0, scaling->z, 0, 0, 0, 0, 1);
    transform(&s, scaled_cloud);
  }

  Vector3f* getDevicePointer() { return m_dev; }
  const Vector3f* getConstDevicePointer() const { return m_dev; }
  size_t getPointCloudSize() const { return m_host.size(); }
  Vector3f* getPoints() const { return m_host.empty() ? NULL : const_cast<Vector3f*>(&m_host[0]); }
  const std::vector<Vector3f>& getPointVector() const { return m_host; }

  // GV/helpers/PointCloud.cu:349-358
  void print()
  {
    for (size_t i = 0; i < m_host.size(); i++)
      std::cout << "Point " << i << ": (" << m_host[i].x << ", " << m_host[i].y << ", " << m_host[i].z << ")" << std::endl;
    std::cout << std::endl << std::endl;
  }

  void resize(uint32_t new_number_of_points)
  {
    m_host.resize(new_number_of_points);
    reserveDevice(new_number_of_points);
  }

private:
  void reserveDevice(size_t n)
  {
    if (n <= m_cap && m_buf) return;
    if (m_buf) gvx_buffer_destroy(m_buf);
    m_buf = 0; m_dev = NULL; m_cap = 0;
    if (!gvxCheck(gvx_buffer_create(&m_buf, n * sizeof(Vector3f)), "PointCloud allocation")) return;
    void* p = NULL;
    gvx_buffer_device_ptr(m_buf, &p, NULL);
    m_dev = static_cast<Vector3f*>(p);
    m_cap = n;
  }
  void toDevice()
  {
    reserveDevice(m_host.size());
    if (!m_host.empty()) gvxCheck(gvx_buffer_upload(m_buf, 0, &m_host[0], m_host.size() * sizeof(Vector3f)), "PointCloud upload");
  }
  void toHost()
  {
    if (!m_host.empty()) gvxCheck(gvx_buffer_download(m_buf, 0, &m_host[0], m_host.size() * sizeof(Vector3f)), "PointCloud download");
  }

  std::vector<Vector3f> m_host;
  gvx_handle_t m_buf;
  Vector3f* m_dev;
  size_t m_cap;
};

}  // namespace gpu_voxels
#endif
