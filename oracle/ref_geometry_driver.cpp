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
#include <sstream>
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

// ---- value API of the vector / matrix types (GV/helpers/cuda_vectors.h:247-331, cuda_matrices.h:285-660) ----------
static Matrix3f m3(const float* p) { Matrix3f m; std::memcpy((void*)&m, p, 36); return m; }
static Matrix4f m4(const float* p) { Matrix4f m; std::memcpy((void*)&m, p, 64); return m; }
static void put3(const Vector3f& v, float* o) { o[0] = v.x; o[1] = v.y; o[2] = v.z; }

void gvrefgeo_to_rpy(const float* m9, unsigned solution, float* out3) { put3(m3(m9).toRPY(solution), out3); }
void gvrefgeo_orientation_diff(const float* a9, const float* b9, float* out3) { put3(m3(a9).orientationMatrixDiff(m3(b9)), out3); }
int gvrefgeo_inverse3(const float* m9, float* out9) {
  Matrix3f inv;
  const bool ok = m3(m9).invertMatrix(inv);
  std::memcpy(out9, &inv, 36);
  return ok ? 1 : 0;
}
int gvrefgeo_inverse4(const float* m16, float* out16) {
  Matrix4f inv;
  const bool ok = m4(m16).invertMatrix(inv);
  std::memcpy(out16, &inv, 64);
  return ok ? 1 : 0;
}
// out: length, dot, angleBetween, cross[3], a.normalized[3], a.abs[3], a normalised in place [3], M4*a [3], M3*a [3],
// then flags as floats: compVec(a,b), apprx_equal(a,b,eps)
void gvrefgeo_vector_ops(const float* a3, const float* b3, const float* m16, double eps, float* out20) {
  Vector3f a = v3(a3), b = v3(b3);
  out20[0] = a.length();
  out20[1] = a.dot(b);
  out20[2] = a.angleBetween(b);
  put3(a.cross(b), out20 + 3);
  put3(a.normalized(), out20 + 6);
  put3(a.abs(), out20 + 9);
  Vector3f n = a;
  n.normalize();
  put3(n, out20 + 12);
  put3(m4(m16) * a, out20 + 15);
  out20[18] = Vector3f::compVec(a, b) ? 1.f : 0.f;
  out20[19] = a.apprx_equal(b, eps) ? 1.f : 0.f;
}
void gvrefgeo_matrix_ops(const float* a16, const float* b16, double eps, float* transpose16, float* minus16, float* prod16,
                         float* rot_apply3, const float* v, int* approx) {
  const Matrix4f a = m4(a16), b = m4(b16);
  const Matrix4f t = a.transpose(), d = a - b, pr = a * b;
  std::memcpy(transpose16, &t, 64);
  std::memcpy(minus16, &d, 64);
  std::memcpy(prod16, &pr, 64);
  put3(a.getRotation() * v3(v), rot_apply3);
  *approx = a.apprx_equal(b, eps) ? 1 : 0;
}
// operator<< of Vector3f / Matrix3f / Matrix4f into buf (NUL-terminated); returns the length
int gvrefgeo_format(int what, const float* p, char* buf, int cap) {
  std::stringstream ss;
  if (what == 0) ss << v3(p);
  else if (what == 1) ss << m3(p);
  else ss << m4(p);
  const std::string s = ss.str();
  const int n = (int)s.size() < cap - 1 ? (int)s.size() : cap - 1;
  std::memcpy(buf, s.data(), n);
  buf[n] = 0;
  return n;
}

}  // extern "C"
