// gpu_voxels/helpers/GeometryGeneration.h -- the host-side point generators the facade and the reference's tests use
// (GV/helpers/GeometryGeneration.cpp:31-221), pinned against the reference's own translation unit by
// tests/golden/geometry.npz (tests/test_host_helpers.py). The exact point set depends on fp32 accumulation
// (`x += delta` in float), so the loops are stated with the same types.
#ifndef GPU_VOXELS_B200_HELPERS_GEOMETRYGENERATION_H
#define GPU_VOXELS_B200_HELPERS_GEOMETRYGENERATION_H

#include <cmath>
#include <vector>

#include <gpu_voxels/helpers/PointCloud.h>
#include <gpu_voxels/helpers/cuda_datatypes.h>

namespace gpu_voxels {
namespace geometry_generation {

inline std::vector<Vector3f> createBoxOfPoints(Vector3f min, Vector3f max, float delta)
{
  std::vector<Vector3f> cloud;
  for (float x = min.x; x <= max.x; x += delta)
    for (float y = min.y; y <= max.y; y += delta)
      for (float z = min.z; z <= max.z; z += delta)
        cloud.push_back(Vector3f(x, y, z));
  return cloud;
}

// GV/helpers/GeometryGeneration.cpp:107-129: the grid points of the bounding box that lie within `radius` of the centre
inline std::vector<Vector3f> createSphereOfPoints(Vector3f center, float radius, float delta)
{
  std::vector<Vector3f> cloud;
  const Vector3f bbox_min(center - Vector3f(radius));
  const Vector3f bbox_max(center + Vector3f(radius));
  for (float x = bbox_min.x; x <= bbox_max.x; x += delta)
    for (float y = bbox_min.y; y <= bbox_max.y; y += delta)
      for (float z = bbox_min.z; z <= bbox_max.z; z += delta)
      {
        const Vector3f point(x, y, z);
        if ((center - point).length() <= radius) cloud.push_back(point);
      }
  return cloud;
}

// GV/helpers/GeometryGeneration.cpp:131-156: axis along z; the half length is formed in double and narrowed to float
inline std::vector<Vector3f> createCylinderOfPoints(Vector3f center, float radius, float length_along_z, float delta)
{
  std::vector<Vector3f> cloud;
  const float half = float(double(length_along_z) / 2.0);
  const Vector3f bbox_min(center - Vector3f(radius, radius, half));
  const Vector3f bbox_max(center + Vector3f(radius, radius, half));
  for (float x = bbox_min.x; x <= bbox_max.x; x += delta)
    for (float y = bbox_min.y; y <= bbox_max.y; y += delta)
      for (float z = bbox_min.z; z <= bbox_max.z; z += delta)
      {
        const float dx = center.x - x, dy = center.y - y;
        if (std::sqrt(dx * dx + dy * dy) <= radius) cloud.push_back(Vector3f(x, y, z));
      }
  return cloud;
}

// GV/helpers/GeometryGeneration.cpp:31-87: the cloud of a box around the origin (params.dim = HALF side lengths), then
// rotated by roll / pitch / yaw and moved to params.center on the device (PointCloud::transformSelf)
namespace detail {
inline Matrix4f orientedBoxTransform(const OrientedBoxParams& params)
{
  return Matrix4f::createFromRotationAndTranslation(Matrix3f::createFromRPY(params.rot), params.center);
}
inline std::vector<Vector3f> orientedBoxCloud(const OrientedBoxParams& params, float spacing)
{
  std::vector<Vector3f> cloud;
  for (float x = -params.dim.x; x <= params.dim.x; x += spacing)
    for (float y = -params.dim.y; y <= params.dim.y; y += spacing)
      for (float z = -params.dim.z; z <= params.dim.z; z += spacing)
        cloud.push_back(Vector3f(x, y, z));
  return cloud;
}
inline std::vector<Vector3f> orientedBoxEdgesCloud(const OrientedBoxParams& params, float spacing)
{
  std::vector<Vector3f> cloud;
  const Vector3f d = params.dim;
  for (float x = -d.x; x <= d.x; x += spacing)
  {
    cloud.push_back(Vector3f(x, d.y, d.z));
    cloud.push_back(Vector3f(x, d.y, -d.z));
    cloud.push_back(Vector3f(x, -d.y, d.z));
    cloud.push_back(Vector3f(x, -d.y, -d.z));
  }
  for (float y = -d.y; y <= d.y; y += spacing)
  {
    cloud.push_back(Vector3f(d.x, y, d.z));
    cloud.push_back(Vector3f(d.x, y, -d.z));
    cloud.push_back(Vector3f(-d.x, y, d.z));
    cloud.push_back(Vector3f(-d.x, y, -d.z));
  }
  for (float z = -d.z; z <= d.z; z += spacing)
  {
    cloud.push_back(Vector3f(d.x, d.y, z));
    cloud.push_back(Vector3f(d.x, -d.y, z));
    cloud.push_back(Vector3f(-d.x, d.y, z));
    cloud.push_back(Vector3f(-d.x, -d.y, z));
  }
  return cloud;
}
}  // namespace detail

inline void createOrientedBoxEdges(const OrientedBoxParams& params, float spacing, PointCloud& ret)
{
  ret.update(detail::orientedBoxEdgesCloud(params, spacing));
  const Matrix4f transformation = detail::orientedBoxTransform(params);
  ret.transformSelf(&transformation);
}
inline void createOrientedBox(const OrientedBoxParams& params, float spacing, PointCloud& ret)
{
  ret.update(detail::orientedBoxCloud(params, spacing));
  const Matrix4f transformation = detail::orientedBoxTransform(params);
  ret.transformSelf(&transformation);
}

// centre of every second voxel per axis, starting at voxel 0 (black) / voxel 1 (white): coordinate =
// float(index * side) + side / 2.0 evaluated in double, stored as float
inline float gvxVoxelCentre(uint32_t index, float side) { return float(double(float(index) * side) + double(side) / 2.0); }

inline void createEquidistantPointsInBox(const size_t max_nr_points, const Vector3ui max_coords, const float side_length,
                                         std::vector<Vector3f>& points)
{
  const size_t start = points.size();  // appends, like the reference
  for (uint32_t i = 0; i < (max_coords.x - 1) / 2; i++)
    for (uint32_t j = 0; j < (max_coords.y - 1) / 2; j++)
      for (uint32_t k = 0; k < (max_coords.z - 1) / 2; k++)
      {
        if (points.size() - start >= max_nr_points) return;
        points.push_back(Vector3f(gvxVoxelCentre(i * 2, side_length), gvxVoxelCentre(j * 2, side_length),
                                  gvxVoxelCentre(k * 2, side_length)));
      }
}

inline void createNonOverlapping3dCheckerboard(const size_t max_nr_points, const Vector3ui max_coords,
                                               const float side_length, std::vector<Vector3f>& black_points,
                                               std::vector<Vector3f>& white_points)
{
  const size_t start = black_points.size();
  for (uint32_t i = 0; i < (max_coords.x - 1) / 2; i++)
    for (uint32_t j = 0; j < (max_coords.y - 1) / 2; j++)
      for (uint32_t k = 0; k < (max_coords.z - 1) / 2; k++)
      {
        if (black_points.size() - start >= max_nr_points) return;
        black_points.push_back(Vector3f(gvxVoxelCentre(i * 2, side_length), gvxVoxelCentre(j * 2, side_length),
                                        gvxVoxelCentre(k * 2, side_length)));
        white_points.push_back(Vector3f(gvxVoxelCentre(i * 2 + 1, side_length), gvxVoxelCentre(j * 2 + 1, side_length),
                                        gvxVoxelCentre(k * 2 + 1, side_length)));
      }
}

}  // namespace geometry_generation
}  // namespace gpu_voxels
#endif
