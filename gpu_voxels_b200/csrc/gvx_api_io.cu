// gvx_api_io.cu -- the reference's on-disk formats for dense maps and voxel lists (host code; include/gvx.h "files").
//
//   map   TemplateVoxelMap::writeToDisk / readFromDisk, GV/voxelmap/TemplateVoxelMap.hpp:666-799 (bin_mode):
//         int32 MapType | float voxel_side_length | uint32 dim.x, dim.y, dim.z | raw voxels (z*dx*dy + y*dx + x)
//   list  TemplateVoxelList::writeToDisk / readFromDisk, GV/voxellist/TemplateVoxelList.hpp:444-526:
//         int32 MapType | uint32 ref_dim.x, .y, .z | float side | uint32 num_voxels | ids[n] uint32 |
//         coords[n] 3 x uint32 | voxels[n] 32 B
// Files written by the reference load here and vice versa (MapType values are the reference's enum order,
// GV/helpers/common_defines.h:61-75).
#include <cstdio>
#include <memory>
#include <new>
#include <vector>
#include "gvx_context.h"

using namespace gvx;

namespace {
constexpr int32_t kFileBitVoxelList = 1;  // MT_BITVECTOR_VOXELLIST

struct File {
  FILE* f;
  explicit File(const char* path, const char* mode) : f(path ? fopen(path, mode) : nullptr) {}
  ~File() { if (f) fclose(f); }
  bool put(const void* p, size_t n) { return n == 0 || fwrite(p, 1, n, f) == n; }
  bool get(void* p, size_t n) { return n == 0 || fread(p, 1, n, f) == n; }
  // bytes between the read position and the end of the file
  uint64_t remaining() {
    const long here = ftell(f);
    if (here < 0 || fseek(f, 0, SEEK_END) != 0) return 0;
    const long end = ftell(f);
    fseek(f, here, SEEK_SET);
    return end > here ? (uint64_t)(end - here) : 0;
  }
};

struct MapInfo { int type; uint32_t dx, dy, dz; float side; uint64_t bytes; bool sharded; };
int map_info(gvx_handle_t h, MapInfo* o) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  *o = MapInfo{m->type, m->dx, m->dy, m->dz, m->side, (uint64_t)m->bytes(), m->sharded()};
  return GVX_OK;
}
struct ListInfo { uint32_t dx, dy, dz; float side; uint64_t size; };
int list_info(gvx_handle_t h, ListInfo* o) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  *o = ListInfo{l->dx, l->dy, l->dz, l->side, l->size};
  return GVX_OK;
}
int fail_locked(int code, const char* what, const char* path) {
  std::lock_guard<ApiMutex> lk(g.mu);
  return fail(code, "%s %s", what, path ? path : "(null)");
}
}  // namespace

extern "C" {

int gvx_map_write_file(gvx_handle_t h, const char* path) {
  MapInfo mi;
  int rc = map_info(h, &mi);
  if (rc) return rc;
  if (mi.sharded) return fail_locked(GVX_ERR_UNSUPPORTED, "a z-slab map holds only part of the grid; not writing", path);
  File out(path, "wb");
  if (!out.f) return fail_locked(GVX_ERR_IO, "Write to file failed:", path);
  std::unique_ptr<char[]> buf(new (std::nothrow) char[mi.bytes]);
  if (!buf) return fail_locked(GVX_ERR_NOMEM, "no host memory for the voxels of", path);
  if ((rc = gvx_map_download(h, buf.get()))) return rc;
  const int32_t type = mi.type;
  const uint32_t dims[3] = {mi.dx, mi.dy, mi.dz};
  if (!(out.put(&type, 4) && out.put(&mi.side, 4) && out.put(dims, 12) && out.put(buf.get(), mi.bytes)))
    return fail_locked(GVX_ERR_IO, "short write to", path);
  return GVX_OK;
}

int gvx_map_read_file(gvx_handle_t h, const char* path) {
  MapInfo mi;
  int rc = map_info(h, &mi);
  if (rc) return rc;
  if (mi.sharded) return fail_locked(GVX_ERR_UNSUPPORTED, "a z-slab map holds only part of the grid; not reading", path);
  File in(path, "rb");
  if (!in.f) return fail_locked(GVX_ERR_IO, "Error in reading file", path);
  int32_t type; float side; uint32_t dims[3];
  if (!(in.get(&type, 4) && in.get(&side, 4) && in.get(dims, 12))) return fail_locked(GVX_ERR_IO, "truncated header in", path);
  if (type != mi.type) return fail_locked(GVX_ERR_INVALID, "Voxelmap type does not match!", path);
  // a differing voxel side length is only a warning in the reference ("Continuing though...", :760-763)
  if (dims[0] != mi.dx || dims[1] != mi.dy || dims[2] != mi.dz)
    return fail_locked(GVX_ERR_DIM_MISMATCH, "Read from file failed: map dimension does not match current object:", path);
  if (in.remaining() < mi.bytes) return fail_locked(GVX_ERR_IO, "truncated voxel data in", path);
  std::unique_ptr<char[]> buf(new (std::nothrow) char[mi.bytes]);
  if (!buf) return fail_locked(GVX_ERR_NOMEM, "no host memory for the voxels of", path);
  if (!in.get(buf.get(), mi.bytes)) return fail_locked(GVX_ERR_IO, "truncated voxel data in", path);
  return gvx_map_upload(h, buf.get());
}

int gvx_list_write_file(gvx_handle_t h, const char* path) {
  ListInfo li;
  int rc = list_info(h, &li);
  if (rc) return rc;
  File out(path, "wb");
  if (!out.f) return fail_locked(GVX_ERR_IO, "Write to file failed:", path);
  const size_t n = (size_t)li.size;
  std::vector<uint32_t> ids, coords;
  std::vector<uint8_t> vox;
  try { ids.resize(n); coords.resize(3 * n); vox.resize(32 * n); }
  catch (const std::bad_alloc&) { return fail_locked(GVX_ERR_NOMEM, "no host memory for the entries of", path); }
  if (n && (rc = gvx_list_download(h, ids.data(), coords.data(), vox.data()))) return rc;
  const int32_t type = kFileBitVoxelList;
  const uint32_t dims[3] = {li.dx, li.dy, li.dz};
  const uint32_t num = (uint32_t)n;
  if (!(out.put(&type, 4) && out.put(dims, 12) && out.put(&li.side, 4) && out.put(&num, 4) && out.put(ids.data(), 4 * n) &&
        out.put(coords.data(), 12 * n) && out.put(vox.data(), 32 * n)))
    return fail_locked(GVX_ERR_IO, "short write to", path);
  return GVX_OK;
}

int gvx_list_read_file(gvx_handle_t h, const char* path) {
  ListInfo li;
  int rc = list_info(h, &li);
  if (rc) return rc;
  File in(path, "rb");
  if (!in.f) return fail_locked(GVX_ERR_IO, "Read from file failed:", path);
  int32_t type; uint32_t dims[3]; float side; uint32_t num;
  if (!in.get(&type, 4)) return fail_locked(GVX_ERR_IO, "truncated header in", path);
  if (type != kFileBitVoxelList) return fail_locked(GVX_ERR_INVALID, "Read from file failed: The map type does not match current object:", path);
  if (!in.get(dims, 12)) return fail_locked(GVX_ERR_IO, "truncated header in", path);
  if (dims[0] != li.dx || dims[1] != li.dy || dims[2] != li.dz)
    return fail_locked(GVX_ERR_DIM_MISMATCH, "Read from file failed: reference map dimension does not match current object:", path);
  if (!in.get(&side, 4)) return fail_locked(GVX_ERR_IO, "truncated header in", path);
  if (side != li.side) return fail_locked(GVX_ERR_INVALID, "Read from file failed: Voxel side length does not match current object:", path);
  if (!in.get(&num, 4)) return fail_locked(GVX_ERR_IO, "truncated header in", path);
  const size_t n = num;
  // the entry count of the header is checked against the bytes that are there before anything is allocated for it
  if (in.remaining() < (uint64_t)n * 48) return fail_locked(GVX_ERR_IO, "truncated list data in", path);
  try {
    std::vector<uint32_t> ids(n), coords(3 * n);
    std::vector<uint8_t> vox(32 * n);
    if (!(in.get(ids.data(), 4 * n) && in.get(coords.data(), 12 * n) && in.get(vox.data(), 32 * n)))
      return fail_locked(GVX_ERR_IO, "truncated list data in", path);
    return gvx_list_upload(h, ids.data(), coords.data(), vox.data(), n);
  } catch (const std::bad_alloc&) {  // no exception crosses the C ABI
    return fail_locked(GVX_ERR_NOMEM, "no host memory for the entries of", path);
  }
}

}  // extern "C"
