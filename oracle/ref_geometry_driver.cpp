// oracle/ref_geometry_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The reference's synthetic-geometry generators (GV/helpers/GeometryGeneration.cpp: boxes, spheres, cylinders, oriented
// boxes, equidistant points, checkerboards) and its rotation-matrix constructors (GV/helpers/cuda_matrices.h) are host
// C++. The translation unit is compiled here UNMODIFIED, by include path from /root/reference, behind a small C
// interface, so that tests/golden/make_golden_geometry.py can record what the reference itself produces and
// tests/test_host_helpers.py can compare the facade's generators (include/gpu_voxels/helpers/GeometryGeneration.h,
// cuda_datatypes.h) with it bit for bit.
//
// createOrientedBox[Edges] hand their cloud to a gpu_voxels::PointCloud (device storage, GV/helpers/PointCloud.cu, not
// buildable without the library). The four PointCloud members the generators touch are defined HERE as recorders:
// update() keeps the host cloud, transformSelf() keeps the matrix -- which is exactly what there is to pin on the host;
// the transform itself is kernelTransformCloud, pinned on the GPU by tests/test_gpu_vs_ref_cuda.py.
// Built only where /root/reference exists, into oracle/_ref/libgvref_geom.so.
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <vector>
#include <gpu_voxels/helpers/GeometryGeneration.cpp>

using gpu_voxels::Matrix3f;
using gpu_voxels::Matrix4f;
using gpu_voxels::Vector3f;
using gpu_voxels::Vector3ui;

static std::vector<Vector3f> g_cloud, g_cloud2;
static Matrix4f g_matrix;

namespace gpu_voxels {
PointCloud::PointCloud() { std::memset((void*)this, 0, sizeof(*this)); }
PointCloud::~PointCloud() {}
void PointCloud::update(const std::vector<Vector3f>& points) { g_cloud = points; }
void PointCloud::transformSelf(const Matrix4f* transform) { g_matrix = *transform; }
}  // namespace gpu_voxels

static Vector3f v3(const float* p) { return Vector3f(p[0], p[1], p[2]); }

extern "C" {

int64_t gvrefgeo_box(const float* min3, const float* max3, float delta) {
  g_cloud = gpu_voxels::geometry_generation::createBoxOfPoints(v3(min3), v3(max3), delta);
  return (int64_t)g_cloud.size();
}
int64_t gvrefgeo_sphere(const float* center3, float radius, float delta) {
  g_cloud = gpu_voxels::geometry_generation::createSphereOfPoints(v3(center3), radius, delta);
  return (int64_t)g_cloud.size();
}
int64_t gvrefgeo_cylinder(const float* center3, float radius, float length_along_z, float delta) {
  g_cloud = gpu_voxels::geometry_generation::createCylinderOfPoints(v3(center3), radius, length_along_z, delta);
  return (int64_t)g_cloud.size();
}
// cloud as handed to PointCloud::update (before the transform) + the Matrix4f handed to transformSelf
int64_t gvrefgeo_oriented_box(const float* dim3, const float* center3, const float* rot3, float spacing, int edges,
                              float* matrix16) {
  gpu_voxels::OrientedBoxParams p;
  p.dim = v3(dim3); p.center = v3(center3); p.rot = v3(rot3);
  gpu_voxels::PointCloud pc;
  g_cloud.clear();
  if (edges) gpu_voxels::geometry_generation::createOrientedBoxEdges(p, spacing, pc);
  else gpu_voxels::geometry_generation::createOrientedBox(p, spacing, pc);
  std::memcpy(matrix16, &g_matrix, 64);
  return (int64_t)g_cloud.size();
}
int64_t gvrefgeo_equidistant(uint64_t max_nr_points, const uint32_t* max_coords3, float side) {
  g_cloud.clear();
  gpu_voxels::geometry_generation::createEquidistantPointsInBox(
      (size_t)max_nr_points, Vector3ui(max_coords3[0], max_coords3[1], max_coords3[2]), side, g_cloud);
  return (int64_t)g_cloud.size();
}
// returns the black count; the white cloud is fetched with which = 1
int64_t gvrefgeo_checkerboard(uint64_t max_nr_points, const uint32_t* max_coords3, float side, int64_t* n_white) {
  g_cloud.clear(); g_cloud2.clear();
  gpu_voxels::geometry_generation::createNonOverlapping3dCheckerboard(
      (size_t)max_nr_points, Vector3ui(max_coords3[0], max_coords3[1], max_coords3[2]), side, g_cloud, g_cloud2);
  *n_white = (int64_t)g_cloud2.size();
  return (int64_t)g_cloud.size();
}
void gvrefgeo_fetch(int which, float* xyz_out) {
  const std::vector<Vector3f>& c = which ? g_cloud2 : g_cloud;
  if (!c.empty()) std::memcpy(xyz_out, c.data(), c.size() * sizeof(Vector3f));
}

// Matrix3f::createFromRPY(Vector3f) and Matrix4f::createFromRotationAndTranslation (GV/helpers/cuda_matrices.h)
void gvrefgeo_rpy(float roll, float pitch, float yaw, float* m9) {
  const Matrix3f m = Matrix3f::createFromRPY(Vector3f(roll, pitch, yaw));
  std::memcpy(m9, &m, 36);
}
void gvrefgeo_rot_trans(const float* m9, const float* t3, float* m16) {
  Matrix3f r;
  std::memcpy((void*)&r, m9, 36);
  const Matrix4f m = Matrix4f::createFromRotationAndTranslation(r, v3(t3));
  std::memcpy(m16, &m, 64);
}

}  // extern "C"
