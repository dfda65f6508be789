// gpu_voxels/helpers/GeometryGeneration.h -- the host-side point generators the facade and the reference's tests use
// (GV/helpers/GeometryGeneration.cpp:90-105, 159-221). The exact point set depends on fp32 accumulation
// (`x += delta` in float), so the loops are stated with the same types.
#ifndef GPU_VOXELS_B200_HELPERS_GEOMETRYGENERATION_H
#define GPU_VOXELS_B200_HELPERS_GEOMETRYGENERATION_H

#include <vector>

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
