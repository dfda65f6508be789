// gpu_voxels/helpers/PointcloudFileHandler.h -- file_handling::PointcloudFileHandler of the reference's API
// (GV/helpers/PointcloudFileHandler.h:40-100, .cpp:64-137): loads .xyz / .pcd / .binvox files into a vector of points.
// The parsing is host code in libgvx.so (gvx_pointcloud_file_load); no GPU is touched.
#ifndef GPU_VOXELS_B200_HELPERS_POINTCLOUDFILEHANDLER_H
#define GPU_VOXELS_B200_HELPERS_POINTCLOUDFILEHANDLER_H

#include <string>
#include <vector>

#include <gpu_voxels/helpers/GvxCall.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>

namespace gpu_voxels {
namespace file_handling {

class PointcloudFileHandler
{
public:
  static PointcloudFileHandler* Instance()
  {
    static PointcloudFileHandler instance;
    return &instance;
  }

  // Appends the file's points to `points`. As in the reference, shift_to_zero, scaling and offset_XYZ then apply to
  // EVERYTHING in `points`, including what it held before the call.
  bool loadPointCloud(const std::string _path, const bool use_model_path, std::vector<Vector3f>& points,
                      const bool shift_to_zero = false, const Vector3f& offset_XYZ = Vector3f(),
                      const float scaling = 1.0)
  {
    float* xyz = NULL;
    uint64_t n = 0;
    // the plain file content first; the post-processing below must see the caller's earlier points too
    if (!gvxCheck(gvx_pointcloud_file_load(_path.c_str(), use_model_path ? 1 : 0, 0, NULL, 1.0f, &xyz, &n),
                  "loadPointCloud"))
      return false;
    const Vector3f* loaded = reinterpret_cast<const Vector3f*>(xyz);
    points.insert(points.end(), loaded, loaded + n);
    gvx_pointcloud_file_free(xyz);
    if (shift_to_zero && !points.empty())
    {
      Vector3f lo = points[0];
      for (size_t i = 1; i < points.size(); ++i)
      {
        if (points[i].x < lo.x) lo.x = points[i].x;
        if (points[i].y < lo.y) lo.y = points[i].y;
        if (points[i].z < lo.z) lo.z = points[i].z;
      }
      for (size_t i = 0; i < points.size(); ++i)
      {
        points[i].x -= lo.x;
        points[i].y -= lo.y;
        points[i].z -= lo.z;
      }
    }
    for (size_t i = 0; i < points.size(); ++i)
    {
      // one rounding per operation, as the reference's host build (x86-64 without FMA) computes it
      volatile float sx = scaling * points[i].x, sy = scaling * points[i].y, sz = scaling * points[i].z;
      points[i].x = sx + offset_XYZ.x;
      points[i].y = sy + offset_XYZ.y;
      points[i].z = sz + offset_XYZ.z;
    }
    return true;
  }

private:
  PointcloudFileHandler() {}
  PointcloudFileHandler(const PointcloudFileHandler&);
  PointcloudFileHandler& operator=(const PointcloudFileHandler&);
};

}  // namespace file_handling
}  // namespace gpu_voxels
#endif
