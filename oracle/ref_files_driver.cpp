// oracle/ref_files_driver.cpp -- TEST INFRASTRUCTURE, not product code.
//
// The reference's point-cloud file readers are plain host C++ (GV/helpers/PointcloudFileHandler.cpp,
// BinvoxFileReader.cpp, XyzFileReader.cpp); their translation units are compiled here UNMODIFIED, by include
// path from /root/reference, behind a two-function C interface, so that tests/golden/make_golden.py can
// record what the reference itself returns for a set of files and tests can compare the product's loader
// (gvx_pointcloud_file_*) with it. The PCD reader is not part of it: it is a thin wrapper around PCL
// (pcl::io::loadPCDFile), which is not installed and not vendored in the reference.
// GV/helpers/MathHelpers.cpp (computeLinearLoad, interpolateLinear) rides along: host code of the same kind.
// Built only where /root/reference exists, into oracle/_ref/libgvref_files.so.
#include <cstdint>
#include <cstring>
#include <map>   // MathHelpers.h relies on a transitive include of <map>
#include <string>
#include <vector>
#include <gpu_voxels/helpers/PointcloudFileHandler.cpp>
#include <gpu_voxels/helpers/BinvoxFileReader.cpp>
#include <gpu_voxels/helpers/XyzFileReader.cpp>
#include <gpu_voxels/helpers/MathHelpers.cpp>

using gpu_voxels::Vector3f;
static std::vector<Vector3f> g_points;

extern "C" {

// loadPointCloud(path, use_model_path, points, shift_to_zero, offset_XYZ, scaling); returns -1 on failure, else the
// number of points now held for gvrefio_fetch
int64_t gvrefio_load(const char* path, int use_model_path, int shift_to_zero, const float* offset3, float scaling) {
  g_points.clear();
  bool ok = gpu_voxels::file_handling::PointcloudFileHandler::Instance()->loadPointCloud(
      std::string(path), use_model_path != 0, g_points, shift_to_zero != 0,
      Vector3f(offset3[0], offset3[1], offset3[2]), scaling);
  return ok ? (int64_t)g_points.size() : -1;
}

void gvrefio_fetch(float* xyz_out) {
  if (!g_points.empty()) memcpy(xyz_out, g_points.data(), g_points.size() * sizeof(Vector3f));
}

// computeLinearLoad / interpolateLinear (GV/helpers/MathHelpers.cpp:29-97), the reference's own translation unit
void gvrefio_linear_load(uint32_t n, uint32_t* blocks, uint32_t* threads) { gpu_voxels::computeLinearLoad(n, blocks, threads); }
float gvrefio_lerp_f(float a, float b, float r) { return gpu_voxels::interpolateLinear(a, b, r); }
double gvrefio_lerp_d(double a, double b, double r) { return gpu_voxels::interpolateLinear(a, b, r); }

}  // extern "C"
