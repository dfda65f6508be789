// gvx_api_list.cu -- C ABI of the BitVectorVoxelList path (include/gvx.h, "voxel lists" section).
#include "gvx_context.h"

using namespace gvx;

namespace {

void list_free(VoxList& l) {
  for (int k = 0; k < 2; ++k) {
    if (l.ids[k]) cudaFree(l.ids[k]);
    if (l.coords[k]) cudaFree(l.coords[k]);
    if (l.voxels[k]) cudaFree(l.voxels[k]);
    l.ids[k] = l.coords[k] = nullptr; l.voxels[k] = nullptr; l.cap[k] = 0;
  }
  if (l.keys_in) cudaFree(l.keys_in);
  if (l.keys_out) cudaFree(l.keys_out);
  if (l.vals_in) cudaFree(l.vals_in);
  if (l.vals_out) cudaFree(l.vals_out);
  if (l.new_meaning) cudaFree(l.new_meaning);
  if (l.tmp) cudaFree(l.tmp);
  l.keys_in = l.keys_out = l.vals_in = l.vals_out = nullptr; l.new_meaning = nullptr; l.tmp = nullptr;
  l.tmp_bytes = 0; l.scratch_cap = l.meaning_cap = 0; l.size = 0;
}

// storage buffer `k` with room for `n` entries (contents are NOT preserved: only ever called on the non-current buffer
// or on an empty list)
int list_reserve(VoxList& l, int k, uint64_t n) {
  if (n <= l.cap[k]) return GVX_OK;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (l.ids[k]) cudaFree(l.ids[k]);
  if (l.coords[k]) cudaFree(l.coords[k]);
  if (l.voxels[k]) cudaFree(l.voxels[k]);
  l.ids[k] = l.coords[k] = nullptr; l.voxels[k] = nullptr; l.cap[k] = 0;
  const uint64_t cap = n + n / 4 + 1024;
  GVX_CUDA(cudaMalloc(&l.ids[k], cap * 4));
  GVX_CUDA(cudaMalloc(&l.coords[k], cap * 12));
  GVX_CUDA(cudaMalloc(&l.voxels[k], cap * kBitVoxelBytes));
  l.cap[k] = cap;
  return GVX_OK;
}

int list_reserve_scratch(VoxList& l, uint64_t total, uint64_t n_new) {
  if (total > l.scratch_cap) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    if (l.keys_in) cudaFree(l.keys_in);
    if (l.keys_out) cudaFree(l.keys_out);
    if (l.vals_in) cudaFree(l.vals_in);
    if (l.vals_out) cudaFree(l.vals_out);
    if (l.tmp) cudaFree(l.tmp);
    l.keys_in = l.keys_out = l.vals_in = l.vals_out = nullptr; l.tmp = nullptr; l.scratch_cap = 0;
    const uint64_t cap = total + total / 4 + 1024;
    GVX_CUDA(cudaMalloc(&l.keys_in, cap * 4));
    GVX_CUDA(cudaMalloc(&l.keys_out, cap * 4));
    GVX_CUDA(cudaMalloc(&l.vals_in, cap * 4));
    GVX_CUDA(cudaMalloc(&l.vals_out, cap * 4));
    l.tmp_bytes = list_sort_temp_bytes(cap) + 256;
    GVX_CUDA(cudaMalloc(&l.tmp, l.tmp_bytes));
    l.scratch_cap = cap;
  }
  if (n_new > l.meaning_cap) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    if (l.new_meaning) cudaFree(l.new_meaning);
    l.new_meaning = nullptr; l.meaning_cap = 0;
    const uint64_t cap = n_new + n_new / 4 + 1024;
    GVX_CUDA(cudaMalloc(&l.new_meaning, cap));
    l.meaning_cap = cap;
  }
  return GVX_OK;
}

int list_insert(VoxList& l, const float* xyz_d, uint64_t n, const MetaDesc& meta, int single_meaning) {
  const uint64_t total = l.size + n;
  if (total >= 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "voxel list would exceed 2^32-1 entries");
  const int nxt = l.cur ^ 1;
  int rc = list_reserve(l, nxt, total);
  if (rc) return rc;
  rc = list_reserve_scratch(l, total, n);
  if (rc) return rc;
  ListBuild b{};
  b.old_ids = l.ids[l.cur]; b.old_coords = l.coords[l.cur]; b.old_voxels = l.voxels[l.cur]; b.old_size = l.size;
  b.xyz_d = xyz_d; b.n_new = n;
  b.dx = l.dx; b.dy = l.dy; b.side = l.side;
  b.meta = meta; b.single_meaning = single_meaning;
  b.keys_in = l.keys_in; b.keys_out = l.keys_out; b.vals_in = l.vals_in; b.vals_out = l.vals_out;
  b.new_meaning = l.new_meaning; b.tmp = l.tmp; b.tmp_bytes = l.tmp_bytes;
  b.out_ids = l.ids[nxt]; b.out_coords = l.coords[nxt]; b.out_voxels = l.voxels[nxt];
  GVX_CUDA(launch_list_build(lctx(), b));
  // the list size is a host value in the reference too (make_unique resizes the vectors)
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  l.size = g.result_h->count;
  l.cur = nxt;
  ++l.version;
  return GVX_OK;
}

// TemplateVoxelList<BitVoxel>::merge (GV/voxellist/TemplateVoxelList.hpp:548-600): append the other list's entries
// (ids + getVoxelIndexSigned(offset), coordinates + offset, voxels or the single new meaning), then make_unique
int list_merge(VoxList& l, const VoxList& o, const int32_t off3[3], int new_meaning) {
  const uint64_t n = o.size, total = l.size + n;
  if (total >= 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "voxel list would exceed 2^32-1 entries");
  if (n == 0) return GVX_OK;
  const int nxt = l.cur ^ 1;
  int rc = list_reserve(l, nxt, total);
  if (rc) return rc;
  rc = list_reserve_scratch(l, total, 0);
  if (rc) return rc;
  ListBuild b{};
  b.old_ids = l.ids[l.cur]; b.old_coords = l.coords[l.cur]; b.old_voxels = l.voxels[l.cur]; b.old_size = l.size;
  b.n_new = n;
  b.add_ids = o.ids[o.cur]; b.add_coords = o.coords[o.cur]; b.add_voxels = o.voxels[o.cur];
  // getVoxelIndexSigned (GV/voxelmap/kernels/VoxelMapOperations.h:61-74) in ptrdiff_t, truncated to the uint32 id
  const int64_t lin = (int64_t)off3[2] * l.dx * l.dy + (int64_t)off3[1] * l.dx + (int64_t)off3[0];
  b.add_id_off = (uint32_t)lin;
  for (int k = 0; k < 3; ++k) b.add_coord_off[k] = off3[k];
  b.add_meaning = new_meaning;
  b.dx = l.dx; b.dy = l.dy; b.side = l.side;
  b.single_meaning = 0;
  b.keys_in = l.keys_in; b.keys_out = l.keys_out; b.vals_in = l.vals_in; b.vals_out = l.vals_out;
  b.new_meaning = l.new_meaning; b.tmp = l.tmp; b.tmp_bytes = l.tmp_bytes;
  b.out_ids = l.ids[nxt]; b.out_coords = l.coords[nxt]; b.out_voxels = l.voxels[nxt];
  GVX_CUDA(launch_list_build(lctx(), b));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  l.size = g.result_h->count;
  l.cur = nxt;
  ++l.version;
  return GVX_OK;
}

struct ListCollideArgs {
  VoxList* l; Map* m; int64_t off;
};
int list_collide_prepare(gvx_handle_t hl, gvx_handle_t hm, const int32_t off3[3], ListCollideArgs* a) {
  a->l = find_list(hl);
  a->m = find_map(hm);
  if (!a->l || !a->m) return fail(GVX_ERR_INVALID, "unknown list or map handle");
  // TemplateVoxelList::collisionCheckWithCollider: differing reference dimensions -> SSIZE_MAX
  // (GV/voxellist/TemplateVoxelList.hpp:737-742)
  if (a->l->dx != a->m->dx || a->l->dy != a->m->dy || a->l->dz != a->m->dz)
    return fail(GVX_ERR_DIM_MISMATCH, "list and map have different reference dimensions");
  a->off = off3 ? (int64_t)off3[2] * a->m->dx * a->m->dy + (int64_t)off3[1] * a->m->dx + (int64_t)off3[0] : 0;
  return GVX_OK;
}

// Entries [i0, i1) of the sorted id array whose map voxel id + off lies inside the map's z-slab ("list ids
// are sorted, so shard boundaries are two binary searches"): a rank then reads size / world ids instead of all of them.
// Whole maps: the full list (the kernel's range test handles offsets). One tiny kernel + one sync per (list contents,
// slab, offset), then cached.
int list_slab_range(VoxList& l, const Map& m, int64_t off, uint64_t* i0, uint64_t* i1) {
  *i0 = 0; *i1 = l.size;
  if (!m.sharded() || l.size == 0) return GVX_OK;
  const int64_t first = (int64_t)m.z0 * m.dx * m.dy, end = first + (int64_t)m.n_voxels();
  // id + off in [first, end)  <=>  id in [first - off, end - off), clamped to the uint32 id range
  const int64_t lo_s = first - off, hi_s = end - off;
  const uint64_t lo = lo_s < 0 ? 0 : (uint64_t)lo_s, hi = hi_s < 0 ? 0 : (uint64_t)hi_s;
  if (!l.ids_exported && l.range.version == l.version && l.range.lo == lo && l.range.hi == hi) {
    *i0 = l.range.i0; *i1 = l.range.i1;
    return GVX_OK;
  }
  GVX_CUDA(launch_list_bounds(lctx(), l.ids[l.cur], l.size, lo, hi));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  *i0 = g.result_h->count;
  *i1 = g.result_h->per_meaning[0];
  if (*i1 < *i0) *i1 = *i0;
  l.range.version = l.version; l.range.lo = lo; l.range.hi = hi; l.range.i0 = *i0; l.range.i1 = *i1;
  return GVX_OK;
}

// the one place that launches list x map: slab range of the list, summary of the map, launch
int list_launch(const LaunchCtx& c, int mode, const ListCollideArgs& a, float thr, bool mark, const uint32_t* mask8,
                unsigned long long* count_d) {
  uint64_t i0, i1;
  int rc = list_slab_range(*a.l, *a.m, a.off, &i0, &i1);
  if (rc) return rc;
  GVX_CUDA(launch_collide_list_map(c, mode, a.l->ids[a.l->cur] + i0, a.l->voxels[a.l->cur] + i0 * kBitVoxelBytes, i1 - i0,
                                   a.m->data, a.m->type, a.m->n_total(), (uint64_t)a.m->z0 * a.m->dx * a.m->dy,
                                   a.m->n_voxels(), float_to_probability(thr), thr, a.off, mark, mask8, count_d,
                                   summary_for_collide(a.m)));
  return GVX_OK;
}

int list_collide(int mode, gvx_handle_t hl, gvx_handle_t hm, float thr, const int32_t off3[3], int mark,
                 const uint8_t* mask_in32, uint8_t* mask_out32, uint64_t* counts256, uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  ListCollideArgs a;
  int rc = list_collide_prepare(hl, hm, off3, &a);
  if (rc) return rc;
  if (mode == 2 && !mask_in32) return fail(GVX_ERR_INVALID, "null types_to_check mask");
  uint32_t mk[8] = {0};
  if (mask_in32) memcpy(mk, mask_in32, 32);
  if (a.l->size == 0) {  // the reference launches nothing useful on an empty list; result is 0
    if (count) *count = 0;
    if (mode == 1 && mask_out32) memset(mask_out32, 0, 32);
    if (mode == 3 && counts256) memset(counts256, 0, 256 * sizeof(uint64_t));
    return GVX_OK;
  }
  if ((rc = list_launch(lctx(), mode, a, thr, mark != 0, mk, nullptr))) return rc;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (shard_active() && a.m->sharded()) {  // GVX_SHARD=z: the list is replicated, the map is this rank's slab -> global result
    uint64_t red[1 + 256 + 4];
    red[0] = g.result_h->count;
    if (mode == 3) memcpy(red + 1, g.result_h->per_meaning, 256 * sizeof(uint64_t));
    else memset(red + 1, 0, 256 * sizeof(uint64_t));
    memcpy(red + 257, g.result_h->mask, 32);
    if ((rc = shard_allreduce(red, 257, 4))) return rc;
    if (count) *count = red[0];
    if (mode == 1 && mask_out32) memcpy(mask_out32, red + 257, 32);
    if (mode == 3 && counts256) memcpy(counts256, red + 1, 256 * sizeof(uint64_t));
    return GVX_OK;
  }
  if (count) *count = g.result_h->count;
  if (mode == 1 && mask_out32) memcpy(mask_out32, g.result_h->mask, 32);  // cleared first, BitVoxelList.hpp:163
  if (mode == 3 && counts256) memcpy(counts256, g.result_h->per_meaning, 256 * sizeof(uint64_t));
  return GVX_OK;
}

// lane variant of the counting list x map collide (gvx_api_lanes.cu)
int lane_list_collide(gvx_handle_t lane, gvx_handle_t hl, gvx_handle_t hm, float thr, const int32_t off3[3],
                      uint64_t* count) {
  std::shared_ptr<Lane> L;
  int rc = lane_acquire(lane, &L);
  if (rc) return rc;
  std::lock_guard<std::mutex> ll(L->mu);
  if (L->dead) return fail(GVX_ERR_INVALID, "lane %llu was destroyed", (unsigned long long)lane);
  {
    std::lock_guard<std::mutex> lk(g.mu.m);
    GVX_NEED_INIT();
    if ((rc = lane_thread_device())) return rc;
    ListCollideArgs a;
    if ((rc = list_collide_prepare(hl, hm, off3, &a))) return rc;
    if (shard_active() && a.m->sharded())
      return fail(GVX_ERR_UNSUPPORTED, "lanes are per process: under GVX_SHARD use the serial entry points");
    if (a.l->size == 0) { if (count) *count = 0; return GVX_OK; }
    LaunchCtx c;
    if ((rc = lane_join_locked(*L, &c))) return rc;
    uint32_t mk[8] = {0};
    // (explicit slab maps: list_launch finds the list's range inside the slab on the library's stream, cached after)
    if ((rc = list_launch(c, 0, a, thr, false, mk, nullptr))) return rc;
    if ((rc = lane_launched_locked(*L))) return rc;
  }
  GVX_CUDA(cudaStreamSynchronize(L->stream));
  if (count) *count = L->result_h->count;
  return GVX_OK;
}

// list x list: both handles, same reference dimensions, no offset (the reference refuses offsets between lists:
// "Offset for VoxelList operation not supported", BitVoxelList.hpp:413-417)
int list_list_prepare(gvx_handle_t ha, gvx_handle_t hb, VoxList** a, VoxList** b) {
  *a = find_list(ha);
  *b = find_list(hb);
  if (!*a || !*b) return fail(GVX_ERR_INVALID, "unknown list handle");
  if ((*a)->dx != (*b)->dx || (*a)->dy != (*b)->dy || (*a)->dz != (*b)->dz)
    return fail(GVX_ERR_DIM_MISMATCH, "lists have different reference dimensions");
  return GVX_OK;
}

int list_list(int mode, gvx_handle_t ha, gvx_handle_t hb, uint32_t margin, uint8_t* types_out32, uint64_t* counts256,
              uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList *a, *b;
  int rc = list_list_prepare(ha, hb, &a, &b);
  if (rc) return rc;
  if (margin > 31) return fail(GVX_ERR_INVALID, "window size %u: the reference's 64-bit window supports at most 31", margin);
  if (count) *count = 0;
  if (mode == 1 && types_out32) memset(types_out32, 0, 32);  // types_in_collision is assigned, not OR-ed (:122)
  if (mode == 3 && counts256) memset(counts256, 0, 256 * sizeof(uint64_t));
  if (a->size == 0 || b->size == 0) return GVX_OK;
  GVX_CUDA(launch_list_list(lctx(), mode, a->ids[a->cur], a->voxels[a->cur], a->size, b->ids[b->cur], b->voxels[b->cur],
                            b->size, margin, nullptr));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (count) *count = g.result_h->count;
  if (mode == 1 && types_out32) memcpy(types_out32, g.result_h->mask, 32);
  if (mode == 3 && counts256) memcpy(counts256, g.result_h->per_meaning, 256 * sizeof(uint64_t));
  return GVX_OK;
}

}  // namespace

namespace gvx {
void lists_free_all() {
  for (auto& kv : g.lists) list_free(kv.second);
  g.lists.clear();
}
}  // namespace gvx

extern "C" {

int gvx_list_create(gvx_handle_t* h, uint32_t dx, uint32_t dy, uint32_t dz, float side) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!h) return fail(GVX_ERR_INVALID, "null handle pointer");
  if (!dx || !dy || !dz || !(side > 0.0f)) return fail(GVX_ERR_INVALID, "bad list geometry %ux%ux%u side %g", dx, dy, dz, side);
  VoxList l;
  l.dx = dx; l.dy = dy; l.dz = dz; l.side = side;
  const uint64_t id = g.next_handle++;
  g.lists.emplace(id, l);
  *h = id;
  return GVX_OK;
}

int gvx_list_destroy(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  list_free(*l);
  g.lists.erase(h);
  return GVX_OK;
}

int gvx_list_clear(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  l->size = 0;
  ++l->version;
  return GVX_OK;
}

int gvx_list_size(gvx_handle_t h, uint64_t* size) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l || !size) return fail(GVX_ERR_INVALID, "unknown list handle or null pointer");
  *size = l->size;
  return GVX_OK;
}

int gvx_list_merge(gvx_handle_t h, gvx_handle_t other, const int32_t voxel_offset[3], int new_meaning) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  VoxList* o = find_list(other);
  if (!l || !o) return fail(GVX_ERR_INVALID, "unknown list handle");
  if (new_meaning < -1 || new_meaning > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", new_meaning);
  static const int32_t zero[3] = {0, 0, 0};
  if (l == o) {  // merging a list into itself: the reference appends a copy of its own entries; with no offset and
                 // no new meaning that changes nothing, otherwise the source would be overwritten while it is read
    if ((!voxel_offset || (!voxel_offset[0] && !voxel_offset[1] && !voxel_offset[2])) && new_meaning < 0) return GVX_OK;
    return fail(GVX_ERR_UNSUPPORTED, "merging a list into itself with an offset or a new meaning");
  }
  return list_merge(*l, *o, voxel_offset ? voxel_offset : zero, new_meaning);
}

int gvx_list_insert_points(gvx_handle_t h, const float* xyz, uint64_t n, int mem_kind, int meaning) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (meaning < 0 || meaning > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meaning);
  if (n == 0) return GVX_OK;
  if (!xyz) return fail(GVX_ERR_INVALID, "null point pointer");
  const float* xyz_d = nullptr;
  int rc = points_to_device(xyz, n, mem_kind, &xyz_d);
  if (rc) return rc;
  MetaDesc d{};
  return list_insert(*l, xyz_d, n, d, meaning);
}

int gvx_list_insert_meta(gvx_handle_t h, const float* xyz, int mem_kind, const uint32_t* cloud_sizes, uint32_t n_clouds,
                         const int32_t* meanings, uint32_t n_meanings) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (!cloud_sizes || !meanings || n_clouds == 0) return fail(GVX_ERR_INVALID, "null/empty cloud description");
  if (n_meanings != 1 && n_meanings != n_clouds)
    return fail(GVX_ERR_INVALID, "number of meanings (%u) differs from number of sub-clouds (%u)", n_meanings, n_clouds);
  for (uint32_t k = 0; k < n_meanings; ++k)
    if (meanings[k] < 0 || meanings[k] > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meanings[k]);
  const size_t off_mg = ((size_t)n_clouds + 1) * 4, total = off_mg + (size_t)n_meanings * 4;
  int slot; void *dh, *dd;
  int rc = desc_acquire(total, &slot, &dh, &dd);
  if (rc) return rc;
  uint32_t* prefix = (uint32_t*)dh;
  uint64_t acc = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) { prefix[c] = (uint32_t)acc; acc += cloud_sizes[c]; }
  if (acc > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "MetaPointCloud larger than 2^32-1 points");
  prefix[n_clouds] = (uint32_t)acc;
  memcpy((char*)dh + off_mg, meanings, (size_t)n_meanings * 4);
  if (acc == 0) return GVX_OK;
  if (!xyz) return fail(GVX_ERR_INVALID, "null point pointer");
  rc = desc_commit(slot, total);
  if (rc) return rc;
  const float* xyz_d = nullptr;
  rc = points_to_device(xyz, acc, mem_kind, &xyz_d);
  if (rc) return rc;
  MetaDesc d{};
  d.prefix_d = (const uint32_t*)dd;
  d.meanings_d = (const int32_t*)((char*)dd + off_mg);
  d.n_clouds = n_clouds;
  d.n_meanings = n_meanings;
  return list_insert(*l, xyz_d, acc, d, 0);
}

int gvx_list_download(gvx_handle_t h, uint32_t* ids, uint32_t* coords, uint8_t* voxels) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (l->size) {
    if (ids) GVX_CUDA(cudaMemcpyAsync(ids, l->ids[l->cur], l->size * 4, cudaMemcpyDeviceToHost, g.stream));
    if (coords) GVX_CUDA(cudaMemcpyAsync(coords, l->coords[l->cur], l->size * 12, cudaMemcpyDeviceToHost, g.stream));
    if (voxels) GVX_CUDA(cudaMemcpyAsync(voxels, l->voxels[l->cur], l->size * kBitVoxelBytes, cudaMemcpyDeviceToHost, g.stream));
  }
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_list_upload(gvx_handle_t h, const uint32_t* ids, const uint32_t* coords, const uint8_t* voxels, uint64_t n) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (n && (!ids || !voxels)) return fail(GVX_ERR_INVALID, "null ids/voxels");
  if (n >= 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "voxel list would exceed 2^32-1 entries");
  for (uint64_t i = 1; i < n; ++i)
    if (ids[i] <= ids[i - 1]) return fail(GVX_ERR_INVALID, "ids must be strictly increasing (entry %llu)", (unsigned long long)i);
  l->size = 0;
  ++l->version;
  int rc = list_reserve(*l, l->cur, n);
  if (rc) return rc;
  if (n) {
    GVX_CUDA(cudaMemcpyAsync(l->ids[l->cur], ids, n * 4, cudaMemcpyHostToDevice, g.stream));
    if (coords) GVX_CUDA(cudaMemcpyAsync(l->coords[l->cur], coords, n * 12, cudaMemcpyHostToDevice, g.stream));
    else GVX_CUDA(cudaMemsetAsync(l->coords[l->cur], 0, n * 12, g.stream));
    GVX_CUDA(cudaMemcpyAsync(l->voxels[l->cur], voxels, n * kBitVoxelBytes, cudaMemcpyHostToDevice, g.stream));
    GVX_CUDA(cudaStreamSynchronize(g.stream));
  }
  l->size = n;
  return GVX_OK;
}

int gvx_list_device_ptrs(gvx_handle_t h, void** ids_d, void** coords_d, void** voxels_d, uint64_t* size) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (ids_d) { *ids_d = l->ids[l->cur]; l->ids_exported = true; }  // the caller may rewrite ids: no range caching
  if (coords_d) *coords_d = l->coords[l->cur];
  if (voxels_d) *voxels_d = l->voxels[l->cur];
  if (size) *size = l->size;
  return GVX_OK;
}

int gvx_collide_list_map(gvx_handle_t l, gvx_handle_t m, float thr, const int32_t off3[3], int mark, uint64_t* count) {
  return list_collide(0, l, m, thr, off3, mark, nullptr, nullptr, nullptr, count);
}
int gvx_collide_list_map_types(gvx_handle_t l, gvx_handle_t m, float thr, const int32_t off3[3], int mark,
                               uint8_t mask_out32[GVX_BITVOXEL_BYTES], uint64_t* count) {
  return list_collide(1, l, m, thr, off3, mark, nullptr, mask_out32, nullptr, count);
}
int gvx_collide_list_map_mask(gvx_handle_t l, gvx_handle_t m, float thr, const int32_t off3[3], int mark,
                              const uint8_t mask_in32[GVX_BITVOXEL_BYTES], uint64_t* count) {
  return list_collide(2, l, m, thr, off3, mark, mask_in32, nullptr, nullptr, count);
}
int gvx_collide_list_map_per_meaning(gvx_handle_t l, gvx_handle_t m, float thr, const int32_t off3[3], int mark,
                                     uint64_t counts256[256], uint64_t* total) {
  return list_collide(3, l, m, thr, off3, mark, nullptr, nullptr, counts256, total);
}
int gvx_collide_list_map_async(gvx_handle_t hl, gvx_handle_t hm, float thr, const int32_t off3[3], int mark,
                               uint64_t* count_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  ListCollideArgs a;
  int rc = list_collide_prepare(hl, hm, off3, &a);
  if (rc) return rc;
  if (a.l->size == 0) {
    if (count_d) GVX_CUDA(cudaMemsetAsync(count_d, 0, 8, g.stream));
    return GVX_OK;
  }
  if ((rc = list_launch(lctx(), 0, a, thr, mark != 0, nullptr, (unsigned long long*)count_d))) return rc;
  return GVX_OK;
}

int gvx_collide_list_map_per_meaning_async(gvx_handle_t hl, gvx_handle_t hm, float thr, const int32_t off3[3], int mark,
                                           uint64_t* counts257_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!counts257_d) return fail(GVX_ERR_INVALID, "null result pointer");
  ListCollideArgs a;
  int rc = list_collide_prepare(hl, hm, off3, &a);
  if (rc) return rc;
  if (a.l->size == 0) {
    GVX_CUDA(cudaMemsetAsync(counts257_d, 0, 257 * sizeof(uint64_t), g.stream));
    return GVX_OK;
  }
  LaunchCtx c = lctx();
  c.result_hmap_d = nullptr;  // the result stays on the device
  return list_launch(c, 3, a, thr, mark != 0, nullptr, (unsigned long long*)counts257_d);  // the kernel writes all 257 words
}

int gvx_collide_list_map_per_meaning_allreduce(gvx_handle_t hl, gvx_handle_t hm, float thr, int mark, uint64_t* counts257_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!counts257_d) return fail(GVX_ERR_INVALID, "null result pointer");
  ListCollideArgs a;
  int rc = list_collide_prepare(hl, hm, nullptr, &a);
  if (rc) return rc;
  // a collective: every rank launches, also with an empty list or an empty slab range (size 0)
  LaunchCtx c = lctx_allreduce();
  if (c.peer.world > 1) c.peer.vec_epoch = ++g.peer_vec_epoch;  // the vector area keeps a sequence of its own
  c.result_hmap_d = nullptr;  // the result stays on the device
  return list_launch(c, 3, a, thr, mark != 0, nullptr, (unsigned long long*)counts257_d);  // the kernel writes all 257 words
}

int gvx_collide_list_map_allreduce(gvx_handle_t hl, gvx_handle_t hm, float thr, int mark, uint64_t* count,
                                   uint64_t* count_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  ListCollideArgs a;
  int rc = list_collide_prepare(hl, hm, nullptr, &a);
  if (rc) return rc;
  // an empty list still has to take part in the collective: launch with size 0
  if ((rc = list_launch(lctx_allreduce(), 0, a, thr, mark != 0, nullptr, (unsigned long long*)count_d))) return rc;
  if (count) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    *count = g.result_h->count;
    if (*count == ~0ull) return fail(GVX_ERR_CUDA, "peer all-reduce timed out: a rank did not arrive");
  }
  return GVX_OK;
}

int gvx_collide_list_list(gvx_handle_t a, gvx_handle_t b, uint64_t* count) {
  return list_list(0, a, b, 0, nullptr, nullptr, count);
}
int gvx_collide_list_list_types(gvx_handle_t a, gvx_handle_t b, uint8_t types_out32[GVX_BITVOXEL_BYTES], uint64_t* count) {
  return list_list(1, a, b, 0, types_out32, nullptr, count);
}
int gvx_collide_list_list_bitcheck(gvx_handle_t a, gvx_handle_t b, uint32_t margin, uint64_t* count) {
  return list_list(2, a, b, margin, nullptr, nullptr, count);
}
int gvx_collide_list_list_per_meaning(gvx_handle_t a, gvx_handle_t b, uint64_t counts256[256], uint64_t* total) {
  return list_list(3, a, b, 0, nullptr, counts256, total);
}

int gvx_list_subtract(gvx_handle_t ha, gvx_handle_t hb) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList *a, *b;
  int rc = list_list_prepare(ha, hb, &a, &b);
  if (rc) return rc;
  if (a->size == 0 || b->size == 0) return GVX_OK;
  if (a == b) { a->size = 0; ++a->version; return GVX_OK; }
  const int nxt = a->cur ^ 1;
  if ((rc = list_reserve(*a, nxt, a->size))) return rc;
  if ((rc = list_reserve_scratch(*a, a->size, 0))) return rc;
  GVX_CUDA(launch_list_list(lctx(), 4, a->ids[a->cur], a->voxels[a->cur], a->size, b->ids[b->cur], b->voxels[b->cur],
                            b->size, 0, a->keys_in));
  GVX_CUDA(launch_exclusive_sum(lctx(), a->keys_in, a->keys_out, a->size, a->tmp, a->tmp_bytes));
  GVX_CUDA(launch_list_compact(lctx(), a->ids[a->cur], a->coords[a->cur], a->voxels[a->cur], a->size, a->keys_in,
                               a->keys_out, a->ids[nxt], a->coords[nxt], a->voxels[nxt]));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  a->size -= g.result_h->count;  // the stencil kernel counted the overlapping entries
  a->cur = nxt;
  ++a->version;
  return GVX_OK;
}

int gvx_list_equals(gvx_handle_t ha, gvx_handle_t hb, int* equal) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* a = find_list(ha);
  VoxList* b = find_list(hb);
  if (!a || !b || !equal) return fail(GVX_ERR_INVALID, "unknown list handle or null result");
  *equal = 0;
  if (a->dx != b->dx || a->dy != b->dy || a->dz != b->dz || a->side != b->side || a->size != b->size) return GVX_OK;
  if (a->size == 0 || a == b) { *equal = 1; return GVX_OK; }
  GVX_CUDA(launch_list_diff(lctx(), a->ids[a->cur], a->coords[a->cur], a->voxels[a->cur], b->ids[b->cur],
                            b->coords[b->cur], b->voxels[b->cur], a->size));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  *equal = g.result_h->count == 0;
  return GVX_OK;
}

int gvx_list_shift_sv_ids(gvx_handle_t h, uint32_t shift) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  VoxList* l = find_list(h);
  if (!l) return fail(GVX_ERR_INVALID, "unknown list handle %llu", (unsigned long long)h);
  if (shift > 56) return fail(GVX_ERR_INVALID, "Maximum shift size is 56! Higher shift number requested. Not performing shift operation.");
  if (l->size == 0) return GVX_OK;
  GVX_CUDA(launch_shift_sv_ids(lctx(), l->voxels[l->cur], l->size, shift, nullptr));
  return GVX_OK;
}

int gvx_lane_collide_list_map(gvx_handle_t lane, gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3],
                              uint64_t* count) {
  return lane_list_collide(lane, list, map, thr, off3, count);
}

}  // extern "C"
