// gvx_api_files.cu -- point-cloud file loaders of the reference (host code; include/gvx.h "point-cloud files").
//
//   file_handling::PointcloudFileHandler::loadPointCloud   GV/helpers/PointcloudFileHandler.cpp:64-137
//   XyzFileReader::readPointCloud                          GV/helpers/XyzFileReader.cpp:28-59
//   BinvoxFileReader::readPointCloud                       GV/helpers/BinvoxFileReader.cpp:38-166
//   PcdFileReader::readPointCloud                          GV/helpers/PcdFileReader.cpp:31-54 (a wrapper around
//       pcl::io::loadPCDFile<PointXYZ>; PCL is a third-party dependency that is neither vendored in the reference nor
//       version-pinned by its CMakeLists (FIND_PACKAGE(PCL)); the PCD v0.7 format is restated from its published
//       description: ascii / binary / binary_compressed (LZF, fields stored one after the other))
//   GpuVoxelsMap::insertPointCloudFromFile                 GV/GpuVoxelsMap.cpp:36-49
//
// Parsing needs no GPU: gvx_pointcloud_file_load works before gvx_init. Numbers are read with the C++ stream
// extractors the reference uses (so "1e50", "nan", "0x10", a trailing comma ... fail or succeed exactly as there),
// arithmetic is single fp32 multiplies and adds in the reference's order (this file is compiled without FMA contraction).
#include <algorithm>
#include <cerrno>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <sstream>
#include <stdexcept>
#include <string>
#include <vector>
#include "gvx_context.h"

using namespace gvx;

namespace {

struct P3 { float x, y, z; };
typedef std::vector<P3> Cloud;

bool slurp(const std::string& path, std::string* out) {
  FILE* f = fopen(path.c_str(), "rb");
  if (!f) return false;
  char buf[1 << 16];
  size_t got;
  out->clear();
  while ((got = fread(buf, 1, sizeof buf, f)) > 0) out->append(buf, got);
  const bool ok = !ferror(f);
  fclose(f);
  return ok;
}

// ---- .xyz: every line holds any number of whitespace-separated "x y z" triples; a line is abandoned at the first
// token that is not a number, an incomplete triple at the end of a line is dropped (XyzFileReader.cpp:42-54)
int read_xyz(const std::string& path, Cloud* pts) {
  std::string text;
  if (!slurp(path, &text)) return fail(GVX_ERR_IO, "Could not open file %s !", path.c_str());
  size_t pos = 0;
  while (pos < text.size()) {
    size_t eol = text.find('\n', pos);
    if (eol == std::string::npos) eol = text.size();
    std::istringstream line(text.substr(pos, eol - pos));
    P3 p;
    while (line >> p.x && line >> p.y && line >> p.z) pts->push_back(p);
    pos = eol + 1;
  }
  return GVX_OK;
}

// ---- .binvox (run-length encoded occupancy grid, x most significant, then z, then y).
// The reference keeps the header values in function-local statics: a file without "translate" / "scale" / a full "dim"
// line inherits them from the file read before (zero in a fresh process), and "scale" is divided by the width known
// when the keyword is met. Kept, because it decides the coordinates of such files (BinvoxFileReader.cpp:44-49,75-86).
struct BinvoxSticky { int depth, height, width; float tx, ty, tz, scale; };
BinvoxSticky g_binvox = {0, 0, 0, 0.f, 0.f, 0.f, 0.f};
std::mutex g_binvox_mu;

int read_binvox(const std::string& path, Cloud* pts) {
  std::string raw;
  if (!slurp(path, &raw)) return fail(GVX_ERR_IO, "Binvox: Could not open file %s !", path.c_str());
  std::lock_guard<std::mutex> lk(g_binvox_mu);
  BinvoxSticky& h = g_binvox;
  std::istringstream in(raw);
  std::string word;
  in >> word;
  if (word != "#binvox") return fail(GVX_ERR_INVALID, "Binvox: First line reads [%s] instead of [#binvox]", word.c_str());
  int version = 0;
  in >> version;
  h.depth = -1;
  bool have_data = false;
  while (in.good() && !have_data) {
    in >> word;
    if (word == "data") have_data = true;
    else if (word == "dim") in >> h.depth >> h.height >> h.width;
    else if (word == "translate") in >> h.tx >> h.ty >> h.tz;
    else if (word == "scale") { in >> h.scale; h.scale = h.scale / h.width; }
    else {  // unknown keyword: the rest of its line is skipped
      int c;
      do c = in.get(); while (in.good() && c != '\n');
    }
  }
  if (!have_data) return fail(GVX_ERR_INVALID, "Binvox: Error reading header");
  if (h.depth == -1) return fail(GVX_ERR_INVALID, "Binvox: Missing dimensions in header");
  const long long cells = (long long)h.width * h.height * h.depth;
  if (cells < 0 || cells > 0x7fffffffLL)  // the reference computes this in int and allocates it
    return fail(GVX_ERR_INVALID, "Binvox: dimensions %d x %d x %d not representable", h.depth, h.height, h.width);
  // cells the file does not cover stay empty (the reference leaves them uninitialised): only the part of the grid the
  // run-length data can reach (255 cells per byte pair) is materialised, whatever the header claims
  size_t at = in.good() ? (size_t)in.tellg() + 1 : raw.size();  // one byte after "data": the line feed
  const long long reachable = at < raw.size() ? (long long)((raw.size() - at) / 2) * 255 : 0;
  const size_t n_grid = (size_t)std::min(cells, reachable);
  std::vector<unsigned char> grid(n_grid, 0);
  long long filled = 0;
  while (filled < cells && at + 1 < raw.size()) {
    const unsigned char value = (unsigned char)raw[at], count = (unsigned char)raw[at + 1];
    at += 2;
    if (filled + count > cells) return fail(GVX_ERR_INVALID, "Binvox: run-length data exceeds the grid");
    memset(grid.data() + filled, value, count);
    filled += count;
  }
  size_t cell = 0;
  for (int x = 0; x < h.depth; ++x)
    for (int z = 0; z < h.height; ++z)
      for (int y = 0; y < h.width; ++y, ++cell)
        if (cell < n_grid && grid[cell] == 1) {  // exactly 1: other non-zero run values are not voxels there either
          P3 p;
          p.x = h.scale * x; p.x = p.x + h.tx;
          p.y = h.scale * y; p.y = p.y + h.ty;
          p.z = h.scale * z; p.z = p.z + h.tz;
          pts->push_back(p);
        }
  return GVX_OK;
}

// ---- .pcd (PCD v0.7): fields x, y, z must be float32 with COUNT 1, as pcl::PointXYZ maps them by name and type
bool lzf_expand(const unsigned char* in, size_t n_in, unsigned char* out, size_t n_out) {
  size_t ip = 0, op = 0;
  while (ip < n_in) {
    unsigned ctrl = in[ip++];
    if (ctrl < 32) {  // literal run of ctrl + 1 bytes
      const size_t len = ctrl + 1;
      if (ip + len > n_in || op + len > n_out) return false;
      memcpy(out + op, in + ip, len);
      ip += len; op += len;
    } else {  // back reference: 3 bits of length (7 = one more length byte follows), 13 bits of distance
      size_t len = ctrl >> 5;
      if (len == 7) { if (ip >= n_in) return false; len += in[ip++]; }
      if (ip >= n_in) return false;
      const size_t dist = ((size_t)(ctrl & 0x1f) << 8) + in[ip++] + 1;
      len += 2;
      if (dist > op || op + len > n_out) return false;
      for (size_t k = 0; k < len; ++k, ++op) out[op] = out[op - dist];  // may overlap: byte by byte
    }
  }
  return op == n_out;
}

struct PcdField { std::string name; int size; char type; int count; size_t offset; };

int read_pcd(const std::string& path, Cloud* pts) {
  std::string raw;
  if (!slurp(path, &raw)) return fail(GVX_ERR_IO, "Could not open file %s !", path.c_str());
  std::vector<PcdField> fields;
  long long width = -1, height = 1, points = -1;
  std::string mode;
  size_t pos = 0;
  bool have_sizes = false, have_types = false;
  while (pos < raw.size() && mode.empty()) {
    size_t eol = raw.find('\n', pos);
    if (eol == std::string::npos) eol = raw.size();
    std::istringstream line(raw.substr(pos, eol - pos));
    pos = eol + 1;
    std::string key, tok;
    if (!(line >> key) || key[0] == '#') continue;
    if (key == "FIELDS" || key == "COLUMNS") {
      while (line >> tok) fields.push_back(PcdField{tok, 4, 'F', 1, 0});
    } else if (key == "SIZE") {
      for (size_t i = 0; i < fields.size(); ++i) if (!(line >> fields[i].size)) return fail(GVX_ERR_INVALID, "PCD: SIZE does not match FIELDS in %s", path.c_str());
      have_sizes = true;
    } else if (key == "TYPE") {
      for (size_t i = 0; i < fields.size(); ++i) if (!(line >> fields[i].type)) return fail(GVX_ERR_INVALID, "PCD: TYPE does not match FIELDS in %s", path.c_str());
      have_types = true;
    } else if (key == "COUNT") {
      for (size_t i = 0; i < fields.size(); ++i) if (!(line >> fields[i].count)) return fail(GVX_ERR_INVALID, "PCD: COUNT does not match FIELDS in %s", path.c_str());
    } else if (key == "WIDTH") line >> width;
    else if (key == "HEIGHT") line >> height;
    else if (key == "POINTS") line >> points;
    else if (key == "DATA") { if (!(line >> mode)) mode = "?"; }
    // VERSION, VIEWPOINT and unknown keys carry nothing the point list needs
  }
  if (fields.empty() || !have_sizes || !have_types || mode.empty())
    return fail(GVX_ERR_INVALID, "PCD: incomplete header in %s", path.c_str());
  if (points < 0) points = width * height;
  if (points < 0 || points > 0x7fffffffLL) return fail(GVX_ERR_INVALID, "PCD: bad point count in %s", path.c_str());
  size_t stride = 0;
  int fx = -1, fy = -1, fz = -1;
  for (size_t i = 0; i < fields.size(); ++i) {
    PcdField& f = fields[i];
    if (f.size <= 0 || f.size > 8 || f.count < 0) return fail(GVX_ERR_INVALID, "PCD: bad field layout in %s", path.c_str());
    f.offset = stride;
    stride += (size_t)f.size * f.count;
    if (f.name == "x") fx = (int)i; else if (f.name == "y") fy = (int)i; else if (f.name == "z") fz = (int)i;
  }
  const int want[3] = {fx, fy, fz};
  for (int a = 0; a < 3; ++a)
    if (want[a] < 0 || fields[want[a]].type != 'F' || fields[want[a]].size != 4 || fields[want[a]].count != 1)
      return fail(GVX_ERR_UNSUPPORTED, "PCD: %s needs float32 fields x, y and z", path.c_str());
  const size_t n = (size_t)points, first = pts->size();
  // the header's point count is checked against the bytes that are there BEFORE anything of that size is allocated
  const size_t left = pos < raw.size() ? raw.size() - pos : 0;
  if (stride > (1u << 20)) return fail(GVX_ERR_INVALID, "PCD: bad field layout in %s", path.c_str());
  if (mode == "ascii" && n > left / 2 + 1) return fail(GVX_ERR_IO, "PCD: %s holds fewer than the %zu points of its header", path.c_str(), n);
  if (mode == "binary" && n * stride > left) return fail(GVX_ERR_IO, "PCD: truncated point data in %s", path.c_str());
  if (mode == "binary_compressed" && n * stride > (left > 8 ? left - 8 : 0) * 90 + 64)  // LZF expands < 90 x
    return fail(GVX_ERR_IO, "PCD: bad compressed block in %s", path.c_str());
  pts->resize(first + n);
  P3* out = pts->data() + first;
  if (mode == "ascii") {
    size_t i = 0;
    while (i < n && pos < raw.size()) {
      size_t eol = raw.find('\n', pos);
      if (eol == std::string::npos) eol = raw.size();
      std::string text = raw.substr(pos, eol - pos);
      pos = eol + 1;
      const char* s = text.c_str();
      while (*s == ' ' || *s == '\t' || *s == '\r') ++s;
      if (!*s) continue;  // empty lines are skipped
      float v[3] = {0.f, 0.f, 0.f};
      for (size_t f = 0; f < fields.size(); ++f)
        for (int c = 0; c < fields[f].count; ++c) {
          char* end;
          const double d = strtod(s, &end);  // accepts "nan" / "inf", as PCL does for float fields
          if (end == s) { pts->resize(first); return fail(GVX_ERR_INVALID, "PCD: malformed point %zu in %s", i, path.c_str()); }
          s = end;
          for (int a = 0; a < 3; ++a) if ((int)f == want[a]) v[a] = (float)d;
        }
      out[i++] = P3{v[0], v[1], v[2]};
    }
    if (i != n) { pts->resize(first); return fail(GVX_ERR_IO, "PCD: %s holds %zu of %zu points", path.c_str(), i, n); }
  } else if (mode == "binary") {
    if (pos + n * stride > raw.size()) { pts->resize(first); return fail(GVX_ERR_IO, "PCD: truncated point data in %s", path.c_str()); }
    const char* base = raw.data() + pos;
    for (size_t i = 0; i < n; ++i) {
      memcpy(&out[i].x, base + i * stride + fields[fx].offset, 4);
      memcpy(&out[i].y, base + i * stride + fields[fy].offset, 4);
      memcpy(&out[i].z, base + i * stride + fields[fz].offset, 4);
    }
  } else if (mode == "binary_compressed") {
    uint32_t packed = 0, plain = 0;
    if (pos + 8 > raw.size()) { pts->resize(first); return fail(GVX_ERR_IO, "PCD: truncated point data in %s", path.c_str()); }
    memcpy(&packed, raw.data() + pos, 4);
    memcpy(&plain, raw.data() + pos + 4, 4);
    if (pos + 8 + packed > raw.size() || (size_t)plain != n * stride) { pts->resize(first); return fail(GVX_ERR_IO, "PCD: bad compressed block in %s", path.c_str()); }
    std::vector<unsigned char> soa(plain);
    if (!lzf_expand((const unsigned char*)raw.data() + pos + 8, packed, soa.data(), plain)) { pts->resize(first); return fail(GVX_ERR_IO, "PCD: corrupt LZF stream in %s", path.c_str()); }
    // decompressed layout: all values of field 0, then all of field 1, ...
    const unsigned char* col[3];
    for (int a = 0; a < 3; ++a) col[a] = soa.data() + fields[want[a]].offset * n;
    for (size_t i = 0; i < n; ++i) {
      memcpy(&out[i].x, col[0] + 4 * i, 4);
      memcpy(&out[i].y, col[1] + 4 * i, 4);
      memcpy(&out[i].z, col[2] + 4 * i, 4);
    }
  } else {
    pts->resize(first);
    return fail(GVX_ERR_INVALID, "PCD: unknown DATA mode '%s' in %s", mode.c_str(), path.c_str());
  }
  return GVX_OK;
}

// getGpuVoxelsPath(use_model_path) / path  (GV/helpers/common_defines.h:276-293; boost::filesystem::path::operator/)
std::string resolve(const char* path, int use_model_path) {
  std::string p(path);
  if (!use_model_path) return p;
  const char* root = getenv("GPU_VOXELS_MODEL_PATH");
  if (!root) {
    fprintf(stderr, "The environment variable 'GPU_VOXELS_MODEL_PATH' could not be read. Did you set it?\n");
    return p;
  }
  std::string r(root);
  if (r.empty() || p.empty()) return r + p;
  if (r[r.size() - 1] == '/' || p[0] == '/') return r + p;
  return r + "/" + p;
}

int load_cloud(const char* path_in, int use_model_path, int shift_to_zero, const float* offset, float scaling, Cloud* pts) {
  if (!path_in) return fail(GVX_ERR_INVALID, "null path");
  const std::string path = resolve(path_in, use_model_path);
  // the format is recognised by the FIRST of "xyz", "pcd", "binvox" found ANYWHERE in the full path (:84-122)
  int rc;
  try {
    if (path.find("xyz") != std::string::npos) rc = read_xyz(path, pts);
    else if (path.find("pcd") != std::string::npos) rc = read_pcd(path, pts);
    else if (path.find("binvox") != std::string::npos) rc = read_binvox(path, pts);
    else return fail(GVX_ERR_INVALID, "%s has no known file format.", path.c_str());
  } catch (const std::bad_alloc&) {  // no exception crosses the C ABI
    pts->clear();
    return fail(GVX_ERR_NOMEM, "out of host memory while reading %s", path.c_str());
  } catch (const std::exception& e) {
    pts->clear();
    return fail(GVX_ERR_INVALID, "%s: %s", path.c_str(), e.what());
  }
  if (rc) return rc;
  if (shift_to_zero && !pts->empty()) {  // (an empty cloud is undefined behaviour there: points[0])
    P3 lo = (*pts)[0];
    for (size_t i = 1; i < pts->size(); ++i) {
      lo.x = std::min(lo.x, (*pts)[i].x);
      lo.y = std::min(lo.y, (*pts)[i].y);
      lo.z = std::min(lo.z, (*pts)[i].z);
    }
    for (size_t i = 0; i < pts->size(); ++i) { (*pts)[i].x -= lo.x; (*pts)[i].y -= lo.y; (*pts)[i].z -= lo.z; }
  }
  const float ox = offset ? offset[0] : 0.f, oy = offset ? offset[1] : 0.f, oz = offset ? offset[2] : 0.f;
  for (size_t i = 0; i < pts->size(); ++i) {
    P3& p = (*pts)[i];
    p.x = scaling * p.x; p.x = p.x + ox;
    p.y = scaling * p.y; p.y = p.y + oy;
    p.z = scaling * p.z; p.z = p.z + oz;
  }
  return GVX_OK;
}

}  // namespace

extern "C" {

int gvx_pointcloud_file_load(const char* path, int use_model_path, int shift_to_zero, const float* offset_xyz,
                             float scaling, float** xyz_out, uint64_t* n_out) {
  if (!xyz_out || !n_out) return fail(GVX_ERR_INVALID, "null output argument");
  *xyz_out = nullptr;
  *n_out = 0;
  Cloud pts;
  int rc = load_cloud(path, use_model_path, shift_to_zero, offset_xyz, scaling, &pts);
  if (rc) return rc;
  if (!pts.empty()) {
    float* buf = (float*)malloc(pts.size() * sizeof(P3));
    if (!buf) return fail(GVX_ERR_NOMEM, "malloc(%zu) failed", pts.size() * sizeof(P3));
    memcpy(buf, pts.data(), pts.size() * sizeof(P3));
    *xyz_out = buf;
  }
  *n_out = pts.size();
  return GVX_OK;
}

void gvx_pointcloud_file_free(float* xyz) { free(xyz); }

int gvx_insert_pointcloud_file(gvx_handle_t map, const char* path, int use_model_path, int meaning, int shift_to_zero,
                               const float* offset_xyz, float scaling, int* any_outside) {
  Cloud pts;
  int rc = load_cloud(path, use_model_path, shift_to_zero, offset_xyz, scaling, &pts);
  if (rc) return rc;
  int outside = 0;  // the cloud lives on this stack frame: the insert must have consumed it before returning
  rc = gvx_insert_points(map, pts.empty() ? nullptr : &pts[0].x, pts.size(), GVX_MEM_HOST, meaning, &outside);
  if (any_outside) *any_outside = outside;
  return rc;
}

int gvx_list_insert_pointcloud_file(gvx_handle_t list, const char* path, int use_model_path, int meaning,
                                    int shift_to_zero, const float* offset_xyz, float scaling) {
  Cloud pts;
  int rc = load_cloud(path, use_model_path, shift_to_zero, offset_xyz, scaling, &pts);
  if (rc) return rc;
  return gvx_list_insert_points(list, pts.empty() ? nullptr : &pts[0].x, pts.size(), GVX_MEM_HOST, meaning);
}

}  // extern "C"
