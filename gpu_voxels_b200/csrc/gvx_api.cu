// gvx_api.cu -- the C ABI (include/gvx.h): context, handle table, host glue. No torch types.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/gvx.h"
#include "gvx_internal.h"

#include "gvx_context.h"

namespace gvx {
Context g;
thread_local std::string t_err;
}  // namespace gvx

using namespace gvx;

namespace gvx {
int mailbox_create_locked(void* ipc_handle64) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle is exchanged as 64 bytes");
  if (!ipc_handle64) return fail(GVX_ERR_INVALID, "null handle buffer");
  if (!g.peer_mailbox) {
    GVX_CUDA(cudaMalloc((void**)&g.peer_mailbox, kPeerMailboxBytes));
    GVX_CUDA(cudaMemset(g.peer_mailbox, 0, kPeerMailboxBytes));
    GVX_CUDA(cudaDeviceSynchronize());
  }
  cudaIpcMemHandle_t h;
  GVX_CUDA(cudaIpcGetMemHandle(&h, g.peer_mailbox));
  memcpy(ipc_handle64, &h, 64);
  return GVX_OK;
}

int mailbox_disconnect_locked();
int mailbox_connect_locked(int rank, int world, const void* all_handles) {
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return fail(GVX_ERR_INVALID, "bad rank %d / world %d", rank, world);
  if (!g.peer_mailbox) return fail(GVX_ERR_INVALID, "gvx_peer_mailbox_create has not been called");
  if (world > 1 && !all_handles) return fail(GVX_ERR_INVALID, "null handles");
  for (int r = 0; r < world; ++r) {
    if (r == rank) { g.peer_ptr[r] = g.peer_mailbox; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)all_handles + 64 * r, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      g.peer_world = r; g.peer_rank = rank;  // unmap what has been mapped so far
      mailbox_disconnect_locked();
      return fail(GVX_ERR_CUDA, "cudaIpcOpenMemHandle for rank %d: %s", r, cudaGetErrorString(e));
    }
    g.peer_ptr[r] = (PeerSlot*)p;
  }
  g.peer_rank = rank;
  g.peer_world = world;
  // the epoch is NOT reset: the mailbox (allocated once per process) keeps the 16-bit tags of the previous session and
  // a count restarted at 0 would match them after a disconnect + connect. Ranks of one session issue the same
  // collectives, so they all hold the same epoch and simply continue the sequence.
  return GVX_OK;
}

int mailbox_disconnect_locked() {
  cudaStreamSynchronize(g.stream);
  for (int r = 0; r < g.peer_world; ++r)
    if (r != g.peer_rank && g.peer_ptr[r]) cudaIpcCloseMemHandle(g.peer_ptr[r]);
  for (int r = 0; r < kMaxPeers; ++r) g.peer_ptr[r] = nullptr;
  g.peer_world = 1; g.peer_rank = 0;
  return GVX_OK;
}

}  // namespace gvx

extern "C" {

static int shutdown_locked(std::vector<std::shared_ptr<Lane>>* lanes);

int gvx_init(int device) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (g.ready) return GVX_OK;
  int n_dev = 0;
  cudaError_t e = cudaGetDeviceCount(&n_dev);
  if (e != cudaSuccess || n_dev == 0)
    return fail(GVX_ERR_CUDA, "no CUDA device available (%s)", e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
  if (device < 0) {  // GVX_SHARD=z: the rank's device, GVX_DEVICES[local rank] or the local rank itself
    const char* mode = getenv("GVX_SHARD");
    if (mode && strcmp(mode, "z") == 0) {
      const char* lr = getenv("GVX_LOCAL_RANK");
      if (!lr) lr = getenv("LOCAL_RANK");
      if (!lr) lr = getenv("OMPI_COMM_WORLD_LOCAL_RANK");
      if (!lr) lr = getenv("GVX_RANK");
      if (!lr) lr = getenv("RANK");
      int local = lr ? atoi(lr) : 0;
      std::vector<int> devs;
      if (const char* dl = getenv("GVX_DEVICES"))
        for (const char* p = dl; *p;) { devs.push_back(atoi(p)); while (*p && *p != ',') ++p; if (*p == ',') ++p; }
      device = devs.empty() ? local % n_dev : devs[(size_t)local % devs.size()];
    }
  }
  if (device >= 0) GVX_CUDA(cudaSetDevice(device));
  GVX_CUDA(cudaGetDevice(&g.device));
  GVX_CUDA(cudaDeviceGetAttribute(&g.sm_count, cudaDevAttrMultiProcessorCount, g.device));
  GVX_CUDA(cudaStreamCreateWithFlags(&g.own_stream, cudaStreamNonBlocking));
  g.stream = g.own_stream;
  GVX_CUDA(cudaMalloc(&g.scratch_d, sizeof(ReduceScratch)));
  GVX_CUDA(cudaMemset(g.scratch_d, 0, sizeof(ReduceScratch)));
  GVX_CUDA(cudaMalloc(&g.result_d, sizeof(ResultBlock)));
  GVX_CUDA(cudaMemset(g.result_d, 0, sizeof(ResultBlock)));
  GVX_CUDA(cudaHostAlloc(&g.result_h, sizeof(ResultBlock), cudaHostAllocMapped));
  memset(g.result_h, 0, sizeof(ResultBlock));
  GVX_CUDA(cudaHostGetDevicePointer((void**)&g.result_hmap_d, g.result_h, 0));
  GVX_CUDA(cudaDeviceSynchronize());
  g.ready = true;
  // GVX_SHARD=z: join the job (rendezvous, peer mailboxes); from then on gvx_map_create makes this rank's z-slab and
  // collision results are global (gvx_api_shard.cu)
  const int rc = shard_init_from_env();
  if (rc != GVX_OK) g.ready = false;
  return rc;
}

int gvx_shutdown(void) {
  std::vector<std::shared_ptr<Lane>> lanes;
  const int rc = shutdown_locked(&lanes);
  lanes_release(lanes);  // outside g.mu: lock order lane.mu -> g.mu
  return rc;
}

static int shutdown_locked(std::vector<std::shared_ptr<Lane>>* lanes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (!g.ready) return GVX_OK;
  cudaDeviceSynchronize();
  lanes_detach_all_locked(lanes);
  {  // peers' slabs mapped for offset collides: every rank unmaps, then a barrier, then the owners free
    bool any_open = false;
    for (auto& kv : g.maps)
      if (kv.second.peers_open) { any_open |= kv.second.peer_world == 0; shard_map_close_peers(&kv.second); }
    if (any_open) shard_barrier();
  }
  if (g.halo_counts_d) { cudaFree(g.halo_counts_d); g.halo_counts_d = nullptr; }
  shard_detach();
  for (auto& kv : g.maps) {
    if (kv.second.owned && kv.second.data) cudaFree(kv.second.data);
    if (kv.second.occ) cudaFree(kv.second.occ);
    if (kv.second.coarse) cudaFree(kv.second.coarse);
  }
  g.maps.clear();
  lists_free_all();
  dispatch_free_all();
  for (auto& kv : g.buffers) if (kv.second.p) cudaFree(kv.second.p);
  g.buffers.clear();
  for (void** p : {&g.bat_bits, &g.bat_seen, &g.bat_tr, &g.bat_counts}) { if (*p) cudaFree(*p); *p = nullptr; }
  for (void** p : {&g.bat_tr_h, &g.bat_counts_h, &g.bat_fk_h}) { if (*p) cudaFreeHost(*p); *p = nullptr; }
  if (g.bat_fk) { cudaFree(g.bat_fk); g.bat_fk = nullptr; }
  g.bat_bits_cap = g.bat_seen_cap = g.bat_tr_cap = g.bat_tr_h_cap = g.bat_counts_cap = g.bat_counts_h_cap = 0;
  g.bat_fk_cap = g.bat_fk_h_cap = 0;
  if (g.stage_d) cudaFree(g.stage_d);
  if (g.stage_ev) { cudaEventDestroy(g.stage_ev); g.stage_ev = nullptr; }
  if (g.copy_stream) {
    cudaStreamSynchronize(g.copy_stream);
    for (int k = 0; k < 2; ++k) {
      if (g.pipe_d[k]) cudaFree(g.pipe_d[k]);
      cudaEventDestroy(g.pipe_ready[k]); cudaEventDestroy(g.pipe_consumed[k]);
      g.pipe_d[k] = nullptr; g.pipe_cap[k] = 0; g.pipe_ready[k] = g.pipe_consumed[k] = nullptr;
    }
    cudaStreamDestroy(g.copy_stream);
    g.copy_stream = nullptr; g.pipe_next = 0; g.pipe_pending = -1;
  }
  for (int k = 0; k < Context::kDescSlots; ++k) {
    if (g.desc_d[k]) cudaFree(g.desc_d[k]);
    if (g.desc_h[k]) cudaFreeHost(g.desc_h[k]);
    if (g.desc_ev[k]) cudaEventDestroy(g.desc_ev[k]);
    g.desc_d[k] = g.desc_h[k] = nullptr; g.desc_cap[k] = 0; g.desc_ev[k] = nullptr;
  }
  for (int r = 0; r < g.peer_world; ++r)
    if (r != g.peer_rank && g.peer_ptr[r]) cudaIpcCloseMemHandle(g.peer_ptr[r]);
  for (int r = 0; r < kMaxPeers; ++r) g.peer_ptr[r] = nullptr;
  g.peer_world = 1; g.peer_rank = 0;
  if (g.peer_mailbox) { cudaFree(g.peer_mailbox); g.peer_mailbox = nullptr; }
  cudaFree(g.scratch_d); cudaFree(g.result_d); cudaFreeHost(g.result_h);
  cudaStreamDestroy(g.own_stream);
  g.stage_d = nullptr; g.stage_cap = 0;
  g.ready = false;
  return GVX_OK;
}

const char* gvx_last_error(void) { return t_err.c_str(); }

int gvx_set_stream(void* cuda_stream) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  // NULL selects the library's own non-blocking stream; pass cudaStreamLegacy / cudaStreamPerThread (or any created
  // stream) explicitly to share a stream with the host program.
  g.stream = cuda_stream ? (cudaStream_t)cuda_stream : g.own_stream;
  return GVX_OK;
}

int gvx_wait_event(void* cuda_event) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!cuda_event) return fail(GVX_ERR_INVALID, "null event");
  GVX_CUDA(cudaStreamWaitEvent(g.stream, (cudaEvent_t)cuda_event, 0));
  return GVX_OK;
}

int gvx_sync(void) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_device_info(int* sm_count, uint64_t* free_bytes, uint64_t* total_bytes, char* name64) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  size_t f = 0, t = 0;
  GVX_CUDA(cudaMemGetInfo(&f, &t));
  if (sm_count) *sm_count = g.sm_count;
  if (free_bytes) *free_bytes = f;
  if (total_bytes) *total_bytes = t;
  if (name64) {
    cudaDeviceProp p;
    GVX_CUDA(cudaGetDeviceProperties(&p, g.device));
    strncpy(name64, p.name, 63);
    name64[63] = 0;
  }
  return GVX_OK;
}

uint64_t gvx_launch_count(void) { return g.launches; }
uint64_t gvx_summary_fallback_count(void) { return g.summary_fallbacks; }

int gvx_set_l2_fetch_granularity(int bytes, int* previous) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  size_t cur = 0;
  GVX_CUDA(cudaDeviceGetLimit(&cur, cudaLimitMaxL2FetchGranularity));
  if (previous) *previous = (int)cur;
  if (bytes == 0) return GVX_OK;
  if (bytes != 32 && bytes != 64 && bytes != 128) return fail(GVX_ERR_INVALID, "granularity must be 32, 64 or 128");
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  GVX_CUDA(cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)bytes));
  return GVX_OK;
}

// ---- device buffers (PointCloud / MetaPointCloud storage of the C++ facade) ---------------------------------
int gvx_buffer_create(gvx_handle_t* h, uint64_t bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!h) return fail(GVX_ERR_INVALID, "null handle pointer");
  Context::DevBuffer b{nullptr, bytes};
  if (bytes) {
    cudaError_t e = cudaMalloc(&b.p, bytes);
    if (e != cudaSuccess) { cudaGetLastError(); return fail(GVX_ERR_NOMEM, "cudaMalloc(%llu) failed: %s", (unsigned long long)bytes, cudaGetErrorString(e)); }
  }
  const uint64_t id = g.next_handle++;
  g.buffers.emplace(id, b);
  *h = id;
  return GVX_OK;
}
int gvx_buffer_destroy(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  auto it = g.buffers.find(h);
  if (it == g.buffers.end()) return fail(GVX_ERR_INVALID, "unknown buffer handle %llu", (unsigned long long)h);
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (it->second.p) GVX_CUDA(cudaFree(it->second.p));
  g.buffers.erase(it);
  return GVX_OK;
}
static int buffer_range(gvx_handle_t h, uint64_t offset, uint64_t bytes, Context::DevBuffer** out) {
  auto it = g.buffers.find(h);
  if (it == g.buffers.end()) return fail(GVX_ERR_INVALID, "unknown buffer handle %llu", (unsigned long long)h);
  if (offset > it->second.bytes || bytes > it->second.bytes - offset)
    return fail(GVX_ERR_INVALID, "range [%llu,+%llu) outside the buffer (%llu bytes)", (unsigned long long)offset,
                (unsigned long long)bytes, (unsigned long long)it->second.bytes);
  *out = &it->second;
  return GVX_OK;
}
int gvx_buffer_upload(gvx_handle_t h, uint64_t offset, const void* host, uint64_t bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Context::DevBuffer* b;
  int rc = buffer_range(h, offset, bytes, &b);
  if (rc) return rc;
  if (!bytes) return GVX_OK;
  if (!host) return fail(GVX_ERR_INVALID, "null host pointer");
  GVX_CUDA(cudaMemcpyAsync((char*)b->p + offset, host, bytes, cudaMemcpyHostToDevice, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));  // pageable host memory may be reused by the caller right away
  return GVX_OK;
}
int gvx_buffer_download(gvx_handle_t h, uint64_t offset, void* host, uint64_t bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Context::DevBuffer* b;
  int rc = buffer_range(h, offset, bytes, &b);
  if (rc) return rc;
  if (!bytes) return GVX_OK;
  if (!host) return fail(GVX_ERR_INVALID, "null host pointer");
  GVX_CUDA(cudaMemcpyAsync(host, (const char*)b->p + offset, bytes, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}
int gvx_buffer_device_ptr(gvx_handle_t h, void** data_d, uint64_t* bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  auto it = g.buffers.find(h);
  if (it == g.buffers.end()) return fail(GVX_ERR_INVALID, "unknown buffer handle %llu", (unsigned long long)h);
  if (data_d) *data_d = it->second.p;
  if (bytes) *bytes = it->second.bytes;
  return GVX_OK;
}

// ---- dense maps ---------------------------------------------------------------------------------------
static int map_make(gvx_handle_t* h, int map_type, void* wrap, uint32_t dx, uint32_t dy, uint32_t dz, float side,
                    uint32_t z0, uint32_t z1) {
  GVX_NEED_INIT();
  if (!h) return fail(GVX_ERR_INVALID, "null handle pointer");
  if (map_type != GVX_MT_BITVECTOR_VOXELMAP && map_type != GVX_MT_PROBAB_VOXELMAP)
    return fail(GVX_ERR_UNSUPPORTED, "map type %d is not a dense voxel map of this path", map_type);
  if (!dx || !dy || !dz || !(side > 0.0f) || z0 >= z1 || z1 > dz)
    return fail(GVX_ERR_INVALID, "bad map geometry %ux%ux%u side %g slab [%u,%u)", dx, dy, dz, side, z0, z1);
  Map m{map_type, dx, dy, dz, z0, z1, side, wrap, wrap == nullptr};
  if (wrap) {
    if ((uintptr_t)wrap & 31) return fail(GVX_ERR_INVALID, "wrapped map pointer must be 32-byte aligned");
  } else {
    cudaError_t e = cudaMalloc(&m.data, m.bytes());
    if (e != cudaSuccess) { cudaGetLastError(); return fail(GVX_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", m.bytes(), cudaGetErrorString(e)); }
    GVX_CUDA(cudaMemsetAsync(m.data, map_type == GVX_MT_PROBAB_VOXELMAP ? 0x80 : 0x00, m.bytes(), g.stream));
    {  // summary, 1 bit per voxel: "occupied" for BitVectorVoxels, "known" (!= UNKNOWN) for ProbabilisticVoxels
      e = cudaMalloc((void**)&m.occ, m.occ_bytes());
      if (e != cudaSuccess) { cudaGetLastError(); cudaFree(m.data); return fail(GVX_ERR_NOMEM, "cudaMalloc(summary) failed: %s", cudaGetErrorString(e)); }
      GVX_CUDA(cudaMemsetAsync(m.occ, 0, m.occ_bytes(), g.stream));
      m.occ_valid = true;
      // coarse mask (1 bit per 4096 voxels), allocated here so that no collide call ever allocates; built lazily
      if (cudaMalloc((void**)&m.coarse, m.coarse_bytes()) != cudaSuccess) { cudaGetLastError(); m.coarse = nullptr; }
    }
  }
  const uint64_t id = g.next_handle++;
  g.maps.emplace(id, m);
  *h = id;
  return GVX_OK;
}

int gvx_map_create(gvx_handle_t* h, int map_type, uint32_t dx, uint32_t dy, uint32_t dz, float side) {
  std::lock_guard<ApiMutex> lk(g.mu);
  uint32_t z0 = 0, z1 = dz;
  const bool slab = shard_active() && dz >= (uint32_t)g.shard_world;
  if (slab) shard_slab(dz, &z0, &z1);  // GVX_SHARD=z: this rank's slab
  const int rc = map_make(h, map_type, nullptr, dx, dy, dz, side, z0, z1);
  if (rc == GVX_OK && slab) find_map(*h)->shard_made = true;
  return rc;
}
int gvx_map_create_slab(gvx_handle_t* h, int map_type, uint32_t dx, uint32_t dy, uint32_t dz, float side, uint32_t z0,
                        uint32_t z1) {
  std::lock_guard<ApiMutex> lk(g.mu);
  return map_make(h, map_type, nullptr, dx, dy, dz, side, z0, z1);
}
int gvx_map_wrap(gvx_handle_t* h, int map_type, void* data_d, uint32_t dx, uint32_t dy, uint32_t dz, float side,
                 uint32_t z0, uint32_t z1) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (!data_d) return fail(GVX_ERR_INVALID, "null device pointer");
  return map_make(h, map_type, data_d, dx, dy, dz, side, z0, z1);
}
int gvx_map_destroy(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (m->peers_open) {
    const bool spmd = m->peer_world == 0;
    shard_map_close_peers(m);
    // GVX_SHARD=z: collective like the mapping was -- nobody frees a slab a peer still has mapped. Explicitly connected
    // maps (gvx_map_connect_peers): the host program disconnects on every rank and synchronises them before destroying
    if (spmd) { int rc = shard_barrier(); if (rc) return rc; }
  }
  if (m->owned) GVX_CUDA(cudaFree(m->data));
  if (m->occ) GVX_CUDA(cudaFree(m->occ));
  if (m->coarse) GVX_CUDA(cudaFree(m->coarse));
  g.maps.erase(h);
  return GVX_OK;
}
static int map_clear_locked(Map* m) {
  GVX_CUDA(cudaMemsetAsync(m->data, m->type == kMapProb ? 0x80 : 0x00, m->bytes(), g.stream));
  if (m->occ) {  // an empty map has an empty summary: clearing re-validates it (unless the raw pointer is out)
    GVX_CUDA(cudaMemsetAsync(m->occ, 0, m->occ_bytes(), g.stream));
    m->revalidate();
  }
  return GVX_OK;
}
int gvx_map_clear(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  return map_clear_locked(m);
}
int gvx_map_clear_bits(gvx_handle_t h, const uint8_t mask32[GVX_BITVOXEL_BYTES]) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m || !mask32) return fail(GVX_ERR_INVALID, "unknown map handle or null mask");
  if (m->type != kMapBit) return fail(GVX_ERR_UNSUPPORTED, "clearBits needs a BitVectorVoxelMap");
  uint32_t w[8];
  memcpy(w, mask32, 32);
  GVX_CUDA(launch_clear_bits(lctx(), m->data, m->n_voxels(), w, m->occ));
  m->revalidate();  // the kernel re-derives every summary word
  return GVX_OK;
}
int gvx_map_clear_meaning(gvx_handle_t h, int meaning) {
  {
    std::lock_guard<ApiMutex> lk(g.mu);
    GVX_NEED_INIT();
    Map* m = find_map(h);
    if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
    if (meaning < 0 || meaning > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meaning);
    if (m->type == kMapProb) {
      if (meaning != GVX_BVM_OCCUPIED) return fail(GVX_ERR_UNSUPPORTED, "ProbVoxelMap only supports eBVM_OCCUPIED");
      return map_clear_locked(m);
    }
  }
  uint8_t mask[32] = {0};
  mask[meaning >> 3] = (uint8_t)(1u << (meaning & 7));
  return gvx_map_clear_bits(h, mask);
}
int gvx_map_device_ptr(gvx_handle_t h, void** data_d, uint64_t* first_voxel, uint64_t* n_voxels) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (data_d) {
    *data_d = m->data;
    // the caller may write voxels behind the library's back, now or later: the summary stays untrusted -- also across
    // clearMap / clearBits / upload -- until gvx_map_refresh_summary
    m->occ_valid = false;
    m->raw_exported = true;
    m->touch();
  }
  if (first_voxel) *first_voxel = (uint64_t)m->z0 * m->dx * m->dy;
  if (n_voxels) *n_voxels = m->n_voxels();
  return GVX_OK;
}
int gvx_map_device_ptr_const(gvx_handle_t h, const void** data_d, uint64_t* first_voxel, uint64_t* n_voxels) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (data_d) *data_d = m->data;  // read-only view: the summary stays valid
  if (first_voxel) *first_voxel = (uint64_t)m->z0 * m->dx * m->dy;
  if (n_voxels) *n_voxels = m->n_voxels();
  return GVX_OK;
}
int gvx_map_info(gvx_handle_t h, int* map_type, uint32_t dims3[3], float* side, uint32_t* z0, uint32_t* z1,
                 uint64_t* bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (map_type) *map_type = m->type;
  if (dims3) { dims3[0] = m->dx; dims3[1] = m->dy; dims3[2] = m->dz; }
  if (side) *side = m->side;
  if (z0) *z0 = m->z0;
  if (z1) *z1 = m->z1;
  if (bytes) *bytes = m->bytes();
  return GVX_OK;
}
int gvx_map_download(gvx_handle_t h, void* host_buf) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m || !host_buf) return fail(GVX_ERR_INVALID, "unknown map handle or null buffer");
  GVX_CUDA(cudaMemcpyAsync(host_buf, m->data, m->bytes(), cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}
int gvx_map_upload(gvx_handle_t h, const void* host_buf) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m || !host_buf) return fail(GVX_ERR_INVALID, "unknown map handle or null buffer");
  GVX_CUDA(cudaMemcpyAsync(m->data, host_buf, m->bytes(), cudaMemcpyHostToDevice, g.stream));
  if (m->occ) {
    // Bit: bit = !noneButEmpty; Prob: bit = occupancy >= -127, i.e. anything but UNKNOWN (-128)
    GVX_CUDA(launch_env_bitmap(lctx(), m->data, m->type, m->n_voxels(), -127, m->occ));
    m->revalidate();
  }
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_map_refresh_summary(gvx_handle_t h) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (!m->occ) return GVX_OK;  // wrapped memory: no summary
  GVX_CUDA(launch_env_bitmap(lctx(), m->data, m->type, m->n_voxels(), -127, m->occ));
  m->raw_exported = false;  // the caller declares its direct writes finished
  m->touch();
  m->occ_valid = true;
  m->fallback_logged = false;
  return GVX_OK;
}

int gvx_map_summary_download(gvx_handle_t h, uint32_t* words, uint64_t n_words, int* valid) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (valid) *valid = m->occ_valid ? 1 : 0;
  if (!m->occ || !words) return GVX_OK;
  const uint64_t have = (m->n_voxels() + 31) / 32;
  if (n_words > have) n_words = have;
  GVX_CUDA(cudaMemcpyAsync(words, m->occ, n_words * 4, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_map_coarse_download(gvx_handle_t h, uint32_t* words, uint64_t n_words, int* valid) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  const bool ok = m->coarse && m->occ_valid && m->coarse_version == m->version;
  if (valid) *valid = ok ? 1 : 0;
  if (!ok || !words) return GVX_OK;
  const uint64_t have = m->coarse_bytes() / 4;
  if (n_words > have) n_words = have;
  GVX_CUDA(cudaMemcpyAsync(words, m->coarse, n_words * 4, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_set_option(int option, int value) {
  std::lock_guard<ApiMutex> lk(g.mu);
  switch (option) {
    case GVX_OPT_USE_SUMMARY: g.use_summary = value != 0; return GVX_OK;
    case GVX_OPT_PDL: g.pdl = value != 0; return GVX_OK;
    case GVX_OPT_USE_COARSE: g.use_coarse = value != 0; return GVX_OK;
    case GVX_OPT_COLLIDE_TMA:
      if (value < -1 || value > 5) return fail(GVX_ERR_INVALID, "GVX_OPT_COLLIDE_TMA must be in [-1, 5]");
      g.collide_tma = value; return GVX_OK;
    case GVX_OPT_DH_CHECK_BITS:
      if (value < 8 || value > 52) return fail(GVX_ERR_INVALID, "GVX_OPT_DH_CHECK_BITS must be in [8, 52]");
      g.dh_check_bits = value; return GVX_OK;
    default: return fail(GVX_ERR_INVALID, "unknown option %d", option);
  }
}

// ---- voxelisation -------------------------------------------------------------------------------------
static int read_flag(int* any_outside) {
  if (!any_outside) return GVX_OK;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  *any_outside = g.result_h->flag ? 1 : 0;
  return GVX_OK;
}

int gvx_insert_points(gvx_handle_t h, const float* xyz, uint64_t n, int mem_kind, int meaning, int* any_outside) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (meaning < 0 || meaning > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meaning);
  if (m->type == kMapProb && meaning != GVX_BVM_OCCUPIED)
    return fail(GVX_ERR_UNSUPPORTED, "ProbVoxelMap only supports eBVM_OCCUPIED");
  if (any_outside) *any_outside = 0;
  if (n == 0) return GVX_OK;
  if (!xyz) return fail(GVX_ERR_INVALID, "null point pointer");
  const float* xyz_d = nullptr;
  int rc = points_to_device(xyz, n, mem_kind, &xyz_d);
  if (rc) return rc;
  GVX_CUDA(launch_insert_points(lctx(), m->view(), xyz_d, n, meaning, any_outside != nullptr));
  return read_flag(any_outside);
}

int gvx_insert_meta(gvx_handle_t h, const float* xyz, int mem_kind, const uint32_t* cloud_sizes, uint32_t n_clouds,
                    const int32_t* meanings, uint32_t n_meanings, const float* transforms, int* any_outside) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (!cloud_sizes || !meanings || n_clouds == 0) return fail(GVX_ERR_INVALID, "null/empty cloud description");
  if (n_meanings != 1 && n_meanings != n_clouds)
    return fail(GVX_ERR_INVALID, "number of meanings (%u) differs from number of sub-clouds (%u)", n_meanings, n_clouds);
  for (uint32_t k = 0; k < n_meanings; ++k) {
    if (meanings[k] < 0 || meanings[k] > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meanings[k]);
    if (m->type == kMapProb && meanings[k] != GVX_BVM_OCCUPIED)
      return fail(GVX_ERR_UNSUPPORTED, "ProbVoxelMap only supports eBVM_OCCUPIED");
  }
  if (any_outside) *any_outside = 0;
  // descriptor = prefix[n_clouds+1] | meanings[n_meanings] | transforms[n_clouds*16], one async H2D
  const size_t off_mg = ((size_t)n_clouds + 1) * 4, off_tr = off_mg + (size_t)n_meanings * 4;
  const size_t total = off_tr + (transforms ? (size_t)n_clouds * 64 : 0);
  int slot; void *dh, *dd;
  int rc = desc_acquire(total, &slot, &dh, &dd);
  if (rc) return rc;
  uint32_t* prefix = (uint32_t*)dh;
  uint64_t acc = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) { prefix[c] = (uint32_t)acc; acc += cloud_sizes[c]; }
  if (acc > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "MetaPointCloud larger than 2^32-1 points");
  prefix[n_clouds] = (uint32_t)acc;
  memcpy((char*)dh + off_mg, meanings, (size_t)n_meanings * 4);
  if (transforms) memcpy((char*)dh + off_tr, transforms, (size_t)n_clouds * 64);
  if (acc == 0) return GVX_OK;
  if (!xyz) return fail(GVX_ERR_INVALID, "null point pointer");
  rc = desc_commit(slot, total);
  if (rc) return rc;
  const float* xyz_d = nullptr;
  rc = points_to_device(xyz, acc, mem_kind, &xyz_d);
  if (rc) return rc;
  MetaDesc d;
  d.prefix_d = (const uint32_t*)dd;
  d.meanings_d = (const int32_t*)((char*)dd + off_mg);
  d.transforms_d = transforms ? (const float*)((char*)dd + off_tr) : nullptr;
  d.n_clouds = n_clouds;
  d.n_meanings = n_meanings;
  GVX_CUDA(launch_insert_meta(lctx(), m->view(), xyz_d, acc, d, any_outside != nullptr));
  return read_flag(any_outside);
}

int gvx_transform_points(const float* in_xyz, float* out_xyz, uint64_t n, int mem_kind, const float M16[16]) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (n == 0) return GVX_OK;
  if (!in_xyz || !out_xyz || !M16) return fail(GVX_ERR_INVALID, "null pointer");
  Affine12 M;
  memcpy(M.a, M16, 48);
  if (mem_kind == GVX_MEM_DEVICE) {
    GVX_CUDA(launch_transform(lctx(), in_xyz, out_xyz, n, M));
    return GVX_OK;
  }
  const float* in_d = nullptr;
  // host in, host out, synchronous: the staging buffer doubles as the output, so no pipelined upload here
  int rc = points_to_device(in_xyz, n, mem_kind == GVX_MEM_HOST_PINNED ? GVX_MEM_HOST : mem_kind, &in_d);
  if (rc) return rc;
  GVX_CUDA(launch_transform(lctx(), in_d, (float*)g.stage_d, n, M));
  GVX_CUDA(cudaMemcpyAsync(out_xyz, g.stage_d, (size_t)n * 12, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

int gvx_update_logodds(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind, const float pose12[12],
                       int delta_occ) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (m->type != kMapProb) return fail(GVX_ERR_UNSUPPORTED, "insertSensorData needs a ProbVoxelMap");
  if (delta_occ < -127 || delta_occ > 127) return fail(GVX_ERR_INVALID, "delta %d outside int8 range", delta_occ);
  if (n == 0) return GVX_OK;
  if (!xyz_sensor || !pose12) return fail(GVX_ERR_INVALID, "null pointer");
  Affine12 P;
  memcpy(P.a, pose12, 48);
  const float* xyz_d = nullptr;
  int rc = points_to_device(xyz_sensor, n, mem_kind, &xyz_d);
  if (rc) return rc;
  GVX_CUDA(launch_update_logodds(lctx(), m->view(), xyz_d, n, P, delta_occ));
  return GVX_OK;
}

int gvx_insert_sensor_data(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind, const float pose12[12],
                           int enable_raycasting, gvx_handle_t robot_map, int robot_bit, int delta_free,
                           int delta_occ) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (m->type != kMapProb) return fail(GVX_ERR_UNSUPPORTED, "insertSensorData needs a ProbVoxelMap");
  if (delta_occ < 0 || delta_occ > 127) return fail(GVX_ERR_INVALID, "delta_occ %d outside [0, 127]", delta_occ);
  if (delta_free > 0 || delta_free < -127) return fail(GVX_ERR_INVALID, "delta_free %d outside [-127, 0]", delta_free);
  const void* robot_d = nullptr;
  if (robot_map) {
    Map* r = find_map(robot_map);
    if (!r) return fail(GVX_ERR_INVALID, "unknown robot map handle %llu", (unsigned long long)robot_map);
    if (r->type != kMapBit) return fail(GVX_ERR_UNSUPPORTED, "cut_real_robot needs a BitVectorVoxelMap of the robot");
    if (r->dx != m->dx || r->dy != m->dy || r->dz != m->dz) return fail(GVX_ERR_DIM_MISMATCH, "robot map dimensions differ");
    if (r->sharded()) return fail(GVX_ERR_UNSUPPORTED, "the robot map of cut_real_robot must be whole on every rank (a ray's hit voxel may lie in another slab)");
    if (robot_bit < 0 || robot_bit > 255) return fail(GVX_ERR_INVALID, "robot bit %d out of range", robot_bit);
    robot_d = r->data;
  }
  if (n == 0) return GVX_OK;
  if (!xyz_sensor || !pose12) return fail(GVX_ERR_INVALID, "null pointer");
  Affine12 P;
  memcpy(P.a, pose12, 48);
  const float* xyz_d = nullptr;
  int rc = points_to_device(xyz_sensor, n, mem_kind, &xyz_d);
  if (rc) return rc;
  // defined order: all free-space decrements of the frame, then all hits (two launches = the barrier between them)
  if (enable_raycasting && delta_free != 0)
    GVX_CUDA(launch_raycast_free(lctx(), m->view(), xyz_d, n, P, delta_free, robot_d, (uint32_t)robot_bit));
  GVX_CUDA(launch_update_logodds(lctx(), m->view(), xyz_d, n, P, delta_occ, robot_d, (uint32_t)robot_bit));
  return GVX_OK;
}

int gvx_update_logodds_multi(gvx_handle_t h, const float* xyz_sensor, int mem_kind, const uint32_t* frame_sizes,
                             uint32_t n_frames, const float* poses12, int delta_occ) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (m->type != kMapProb) return fail(GVX_ERR_UNSUPPORTED, "insertSensorData needs a ProbVoxelMap");
  if (delta_occ < -127 || delta_occ > 127) return fail(GVX_ERR_INVALID, "delta %d outside int8 range", delta_occ);
  if (!n_frames) return GVX_OK;
  if (!frame_sizes || !poses12) return fail(GVX_ERR_INVALID, "null pointer");
  if (n_frames <= (uint32_t)kMaxFramesByValue) {  // the usual case: everything travels in the kernel parameters
    FramesByValue fv;
    uint64_t acc = 0;
    for (uint32_t f = 0; f < n_frames; ++f) { fv.prefix[f] = (uint32_t)acc; acc += frame_sizes[f]; }
    if (acc > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "more than 2^32-1 sensor points in one call");
    fv.prefix[n_frames] = (uint32_t)acc;
    memcpy(fv.poses, poses12, (size_t)n_frames * 48);
    if (acc == 0) return GVX_OK;
    // scheduling hint only: a frame whose sensor stands inside this slab will do most of the atomics; its CTAs go first
    uint32_t k = 0;
    const float zlo = (float)m->z0 * m->side, zhi = (float)m->z1 * m->side;
    for (int pass = 0; pass < 2; ++pass)
      for (uint32_t f = 0; f < n_frames; ++f) {
        const float sz = poses12[12 * f + 11];
        if (((sz >= zlo && sz < zhi) ? 0 : 1) == pass) fv.order[k++] = (uint8_t)f;
      }
    if (!xyz_sensor) return fail(GVX_ERR_INVALID, "null point pointer");
    const float* xyz_d = nullptr;
    int rc = points_to_device(xyz_sensor, acc, mem_kind, &xyz_d);
    if (rc) return rc;
    GVX_CUDA(launch_update_logodds_multi(lctx(), m->view(), xyz_d, acc, nullptr, nullptr, n_frames, delta_occ, &fv));
    return GVX_OK;
  }
  const size_t off_pose = ((size_t)n_frames + 1) * 4, total = off_pose + (size_t)n_frames * 48;
  int slot; void *dh, *dd;
  int rc = desc_acquire(total, &slot, &dh, &dd);
  if (rc) return rc;
  uint32_t* prefix = (uint32_t*)dh;
  uint64_t acc = 0;
  for (uint32_t f = 0; f < n_frames; ++f) { prefix[f] = (uint32_t)acc; acc += frame_sizes[f]; }
  if (acc > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "more than 2^32-1 sensor points in one call");
  prefix[n_frames] = (uint32_t)acc;
  memcpy((char*)dh + off_pose, poses12, (size_t)n_frames * 48);
  if (acc == 0) return GVX_OK;
  if (!xyz_sensor) return fail(GVX_ERR_INVALID, "null point pointer");
  if ((rc = desc_commit(slot, total))) return rc;
  const float* xyz_d = nullptr;
  if ((rc = points_to_device(xyz_sensor, acc, mem_kind, &xyz_d))) return rc;
  GVX_CUDA(launch_update_logodds_multi(lctx(), m->view(), xyz_d, acc, (const uint32_t*)dd,
                                       (const float*)((const char*)dd + off_pose), n_frames, delta_occ));
  return GVX_OK;
}

// ---- collision counting, map x map --------------------------------------------------------------------
static int collide_prepare(gvx_handle_t ha, gvx_handle_t hb, const int32_t off3[3], Map** a, Map** b, int64_t* off,
                           int64_t* i0, int64_t* i1, bool allow_halo = false) {
  *a = find_map(ha);
  *b = find_map(hb);
  if (!*a || !*b) return fail(GVX_ERR_INVALID, "unknown map handle");
  Map *A = *a, *B = *b;
  if (A->dx != B->dx || A->dy != B->dy || A->dz != B->dz || A->z0 != B->z0 || A->z1 != B->z1)
    return fail(GVX_ERR_DIM_MISMATCH, "maps differ in dimensions or slab");
  int64_t o = 0;
  if (off3) o = (int64_t)off3[2] * A->dx * A->dy + (int64_t)off3[1] * A->dx + (int64_t)off3[0];  // getVoxelIndexSigned
  // an offset shifts the LINEAR index: on z-slab maps the partner voxel of a slab's first / last layers (any offset) or
  // of every voxel (z offsets) lives in another rank's slab. Under GVX_SHARD=z the synchronous collides read it from the
  // peer over CUDA IPC (collide_halo below); every other combination is refused
  const bool spmd_halo = shard_active() && A->shard_made && B->shard_made;
  const bool explicit_halo = B->peers_open && B->peer_world > 0;  // gvx_map_connect_peers
  if (o != 0 && A->sharded() && !(allow_halo && (spmd_halo || explicit_halo)))
    return fail(GVX_ERR_UNSUPPORTED, "offset collide on a z-slab map needs the peers' slabs: supported by the synchronous "
                                     "collideWith under GVX_SHARD=z, or after gvx_map_connect_peers on the other map "
                                     "(replicas only otherwise)");
  const int64_t n = (int64_t)A->n_voxels();
  *off = o;
  *i0 = o < 0 ? -o : 0;
  *i1 = o > 0 ? n - o : n;
  if (*i1 < *i0) *i1 = *i0;
  return GVX_OK;
}

// A marking collide at the int8 threshold -128 (a nonsense threshold that converts to it, see launch_collide_map_map)
// lets UNKNOWN voxels of a ProbVoxelMap collide and writes 127 into them: they become known behind the summary's back.
// Re-derive the "known" summary from the voxels on the stream (one pass; the threshold is an error case, not a path).
static int fixup_summary_after_mark(Map* a, int8_t thr_i, bool mark) {
  if (mark && thr_i == -128 && a->type == kMapProb && a->occ && a->occ_valid) {
    GVX_CUDA(launch_env_bitmap(lctx(), a->data, a->type, a->n_voxels(), -127, a->occ));
    a->touch();
  }
  return GVX_OK;
}

// Offset collide on z-slab maps under GVX_SHARD=z (collective). The reference shifts THIS map's pointer
// (GV/voxelmap/TemplateVoxelMap.hpp:486-519: getVoxelPtrSignedOffset on m_dev_data): voxel i + off of `a` meets voxel i of
// `b`, for every grid index i with both in range. This rank owns the a-voxels of its slab (it counts and marks them);
// their partners i = j - off lie in some ranks' slabs of `b`. For every rank s whose b-slab holds partners: one ordinary
// offset-collide launch over that sub-range, with b = rank s's slab mapped through CUDA IPC (loads over NVLink; own
// slab: local) and the offset rewritten slab-to-slab. Counts land in one device word per peer slab, are summed on the
// host and all-reduced.
// The launches of an offset collide for the rank that owns layers [z_bounds[rank], z_bounds[rank + 1]) of `a`: for every
// rank s whose b-slab holds partners, segment s = {i_begin, i_end, off_local}: the kernel runs over b's LOCAL indices
// [i_begin, i_end) of rank s's slab and meets b_s[i'] with a_local[i' + off_local]. i_begin == i_end: nothing to do with
// rank s. Pure arithmetic (no CUDA): exported as gvx_halo_plan and pinned by brute force on the CPU
// (tests/test_shard_host.py).
static void halo_plan(int64_t layer, uint32_t dz, const uint32_t* z_bounds, int world, int rank, int64_t off,
                      int64_t* seg3 /* world x {i_begin, i_end, off_local} */) {
  const int64_t n_tot = layer * dz;
  const int64_t A0 = (int64_t)z_bounds[rank] * layer, A1 = (int64_t)z_bounds[rank + 1] * layer;
  // grid indices i of b with a partner i + off inside this rank's slab of a (and inside the grid)
  const int64_t lo_valid = std::max<int64_t>(A0 - off, off < 0 ? -off : 0), hi_valid = std::min<int64_t>(A1 - off, off > 0 ? n_tot - off : n_tot);
  for (int s = 0; s < world; ++s) {
    const int64_t B0 = (int64_t)z_bounds[s] * layer, B1 = (int64_t)z_bounds[s + 1] * layer;
    const int64_t lo = std::max(lo_valid, B0), hi = std::min(hi_valid, B1);
    seg3[3 * s] = lo < hi ? lo - B0 : 0;
    seg3[3 * s + 1] = lo < hi ? hi - B0 : 0;
    seg3[3 * s + 2] = B0 + off - A0;  // a's local index = i + off - A0 = i' + B0 + off - A0
  }
}

int gvx_halo_plan(uint32_t dx, uint32_t dy, uint32_t dz, const uint32_t* z_bounds, int world, int rank,
                  const int32_t off3[3], int64_t* seg3) {
  if (!z_bounds || !off3 || !seg3 || world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
    return fail(GVX_ERR_INVALID, "gvx_halo_plan: bad arguments");
  if (z_bounds[0] != 0 || z_bounds[world] != dz) return fail(GVX_ERR_INVALID, "z_bounds must run from 0 to dz");
  for (int r = 0; r < world; ++r)
    if (z_bounds[r] > z_bounds[r + 1]) return fail(GVX_ERR_INVALID, "z_bounds must not decrease");
  const int64_t layer = (int64_t)dx * dy;
  halo_plan(layer, dz, z_bounds, world, rank, (int64_t)off3[2] * layer + (int64_t)off3[1] * dx + off3[0], seg3);
  return GVX_OK;
}

static int collide_halo(Map* a, Map* b, int64_t off, int8_t thr, int thr_b_or_same, bool mark, uint64_t* count) {
  if (a == b && mark)
    return fail(GVX_ERR_UNSUPPORTED, "marking a z-slab map while colliding it with a shifted copy of itself: the peers read "
                                     "what this rank marks");
  // explicit slab maps (gvx_map_connect_peers): the host program owns the communicator -- it orders the ranks around
  // the call and all-reduces the LOCAL count this returns. GVX_SHARD=z: both happen here.
  const bool spmd = !(b->peers_open && b->peer_world > 0);
  int rc = GVX_OK;
  if (spmd) {
    if ((rc = shard_map_open_peers(b))) return rc;  // first time: collective CUDA IPC exchange
    // what the peers are about to read must be complete on every rank, and nobody may be ahead by a whole call
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    if ((rc = shard_barrier())) return rc;
  }
  const int world = spmd ? g.shard_world : b->peer_world;
  if (!g.halo_counts_d) GVX_CUDA(cudaMalloc((void**)&g.halo_counts_d, kMaxPeers * sizeof(unsigned long long)));
  GVX_CUDA(cudaMemsetAsync(g.halo_counts_d, 0, kMaxPeers * sizeof(unsigned long long), g.stream));
  uint32_t zb[kMaxPeers + 1];
  for (int s = 0; s < world; ++s) {
    uint32_t z1;
    if (spmd) shard_slab_of(s, b->dz, &zb[s], &z1);
    else zb[s] = b->peer_z[s];
  }
  zb[world] = b->dz;
  const int rank = spmd ? g.shard_rank : b->peer_rank;
  if (zb[rank] != a->z0 || zb[rank + 1] != a->z1) return fail(GVX_ERR_INVALID, "slab [%u,%u) is not rank %d's share of the partition", a->z0, a->z1, rank);
  int64_t seg[3 * kMaxPeers];
  halo_plan((int64_t)a->dx * a->dy, a->dz, zb, world, rank, off, seg);
  for (int s = 0; s < world; ++s) {
    if (seg[3 * s] >= seg[3 * s + 1]) continue;
    LaunchCtx c = lctx();
    c.result_hmap_d = nullptr;  // device-only result, one word per peer slab
    GVX_CUDA(launch_collide_map_map(c, a->data, a->type, b->peer_data[s], b->type, seg[3 * s], seg[3 * s + 1], seg[3 * s + 2],
                                    thr, mark, false, g.halo_counts_d + s, nullptr, nullptr, nullptr, nullptr, thr_b_or_same));
  }
  if ((rc = fixup_summary_after_mark(a, thr, mark))) return rc;
  unsigned long long parts[kMaxPeers];
  GVX_CUDA(cudaMemcpyAsync(parts, g.halo_counts_d, sizeof parts, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  uint64_t c = 0;
  for (int s = 0; s < world; ++s) c += parts[s];
  if (spmd && (rc = shard_allreduce(&c, 1, 0))) return rc;  // also: every rank's reads are done before anyone mutates `b` again
  if (count) *count = c;
  return GVX_OK;
}

int gvx_collide_map_map(gvx_handle_t ha, gvx_handle_t hb, float thr, const int32_t off3[3], int mark, uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, off3, &a, &b, &off, &i0, &i1, true);
  if (rc) return rc;
  if (off != 0 && a->sharded()) return collide_halo(a, b, off, float_to_probability(thr), kSameThreshold, mark != 0, count);
  // GVX_SHARD=z: the maps are this rank's slabs and the caller expects the count over the whole grid -- all-reduced
  // inside the collide kernel when the peers' mailboxes are mapped, else through the host segment
  const bool global = shard_active() && a->sharded();
  const bool fused = global && g.shard_fused && g.peer_world > 1;
  const uint32_t *ca = nullptr, *cb = nullptr;
  if (off == 0 && ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb)))) return rc;
  GVX_CUDA(launch_collide_map_map(fused ? lctx_allreduce() : lctx(), a->data, a->type, b->data, b->type, i0, i1, off,
                                  float_to_probability(thr), mark != 0, false, nullptr,
                                  summary_for_collide(a), summary_for_collide(b), ca, cb));
  if ((rc = fixup_summary_after_mark(a, float_to_probability(thr), mark != 0))) return rc;
  if (count || global) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    uint64_t c = g.result_h->count;
    if (fused && c == ~0ull) return fail(GVX_ERR_CUDA, "peer all-reduce timed out: a rank did not arrive");
    if (global && !fused && (rc = shard_allreduce(&c, 1, 0))) return rc;
    if (count) *count = c;
  }
  return GVX_OK;
}

int gvx_collide_map_map_thresholds(gvx_handle_t ha, gvx_handle_t hb, int threshold1, int threshold2,
                                   const int32_t off3[3], int mark, uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (threshold1 < -128 || threshold1 > 127 || threshold2 < -128 || threshold2 > 127)
    return fail(GVX_ERR_INVALID, "thresholds (%d, %d) outside the int8 range of Probability", threshold1, threshold2);
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, off3, &a, &b, &off, &i0, &i1, true);
  if (rc) return rc;
  if (off != 0 && a->sharded()) return collide_halo(a, b, off, (int8_t)threshold1, threshold2, mark != 0, count);
  const bool global = shard_active() && a->sharded();
  const bool fused = global && g.shard_fused && g.peer_world > 1;
  const uint32_t *ca = nullptr, *cb = nullptr;
  if (off == 0 && ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb)))) return rc;
  GVX_CUDA(launch_collide_map_map(fused ? lctx_allreduce() : lctx(), a->data, a->type, b->data, b->type, i0, i1, off,
                                  (int8_t)threshold1, mark != 0, false, nullptr, summary_for_collide(a),
                                  summary_for_collide(b), ca, cb, threshold2));
  if ((rc = fixup_summary_after_mark(a, (int8_t)threshold1, mark != 0))) return rc;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  uint64_t c = g.result_h->count;
  if (fused && c == ~0ull) return fail(GVX_ERR_CUDA, "peer all-reduce timed out: a rank did not arrive");
  if (global && !fused && (rc = shard_allreduce(&c, 1, 0))) return rc;
  if (count) *count = c;
  return GVX_OK;
}

int gvx_collide_map_map_async(gvx_handle_t ha, gvx_handle_t hb, float thr, const int32_t off3[3], int mark,
                              uint64_t* count_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, off3, &a, &b, &off, &i0, &i1);
  if (rc) return rc;
  LaunchCtx c = lctx();
  if (count_d) c.result_hmap_d = nullptr;  // the result stays on the device: nothing is written to mapped host memory
  const uint32_t *ca = nullptr, *cb = nullptr;
  if (off == 0 && ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb)))) return rc;
  GVX_CUDA(launch_collide_map_map(c, a->data, a->type, b->data, b->type, i0, i1, off, float_to_probability(thr),
                                  mark != 0, false, (unsigned long long*)count_d,
                                  summary_for_collide(a), summary_for_collide(b), ca, cb));
  return fixup_summary_after_mark(a, float_to_probability(thr), mark != 0);
}

// ---- fused all-reduce over peer memory ----------------------------------------------------------------
int gvx_peer_mailbox_create(void* ipc_handle64) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  return mailbox_create_locked(ipc_handle64);
}
int gvx_peer_mailbox_connect(int rank, int world, const void* all_handles) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  return mailbox_connect_locked(rank, world, all_handles);
}
int gvx_peer_mailbox_disconnect(void) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (!g.ready) return GVX_OK;
  return mailbox_disconnect_locked();
}

int gvx_collide_map_map_allreduce(gvx_handle_t ha, gvx_handle_t hb, float thr, int mark, uint64_t* count,
                                  uint64_t* count_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, nullptr, &a, &b, &off, &i0, &i1);
  if (rc) return rc;
  LaunchCtx c = lctx_allreduce();
  if (!count && count_d) c.result_hmap_d = nullptr;  // device-only result
  const uint32_t *ca = nullptr, *cb = nullptr;
  if ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb))) return rc;
  GVX_CUDA(launch_collide_map_map(c, a->data, a->type, b->data, b->type, i0, i1, 0,
                                  float_to_probability(thr), mark != 0, false, (unsigned long long*)count_d,
                                  summary_for_collide(a), summary_for_collide(b), ca, cb));
  if ((rc = fixup_summary_after_mark(a, float_to_probability(thr), mark != 0))) return rc;
  if (count) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));
    *count = g.result_h->count;
    if (*count == ~0ull) return fail(GVX_ERR_CUDA, "peer all-reduce timed out: a rank did not arrive");
  }
  return GVX_OK;
}

int gvx_collide_map_map_any(gvx_handle_t ha, gvx_handle_t hb, float thr, int* flag) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, nullptr, &a, &b, &off, &i0, &i1);
  if (rc) return rc;
  LaunchCtx c = lctx();
  c.any = true;  // the kernels leave as soon as one CTA has found a collision; the count is then a lower bound
  const uint32_t *ca = nullptr, *cb = nullptr;
  if ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb))) return rc;
  GVX_CUDA(launch_collide_map_map(c, a->data, a->type, b->data, b->type, i0, i1, 0, float_to_probability(thr),
                                  false, false, nullptr,
                                  summary_for_collide(a), summary_for_collide(b), ca, cb));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  uint64_t any = g.result_h->count != 0;
  if (shard_active() && a->sharded() && (rc = shard_allreduce(&any, 0, 1))) return rc;  // OR over the ranks' slabs
  if (flag) *flag = any != 0;
  return GVX_OK;
}

// Lane variants (gvx_api_lanes.cu): the same kernels on the lane's stream and result blocks. g.mu covers validation and
// the launch only; the wait for the result runs outside it, so queries of different lanes overlap on the GPU.
static int lane_collide_map_map(gvx_handle_t lane, gvx_handle_t ha, gvx_handle_t hb, float thr, const int32_t off3[3],
                                bool any, uint64_t* count) {
  std::shared_ptr<Lane> L;
  int rc = lane_acquire(lane, &L);
  if (rc) return rc;
  std::lock_guard<std::mutex> ll(L->mu);
  if (L->dead) return fail(GVX_ERR_INVALID, "lane %llu was destroyed", (unsigned long long)lane);
  {
    std::lock_guard<std::mutex> lk(g.mu.m);
    GVX_NEED_INIT();
    if ((rc = lane_thread_device())) return rc;
    Map *a, *b;
    int64_t off, i0, i1;
    if ((rc = collide_prepare(ha, hb, off3, &a, &b, &off, &i0, &i1))) return rc;
    if (shard_active() && a->sharded())
      return fail(GVX_ERR_UNSUPPORTED, "lanes are per process: under GVX_SHARD the global result needs a collective, which "
                                       "every rank must issue in the same order -- use the serial entry points");
    const uint32_t *ca = nullptr, *cb = nullptr;
    if (off == 0 && ((rc = coarse_for_collide(a, &ca)) || (rc = coarse_for_collide(b, &cb)))) return rc;
    const uint32_t *sa = summary_for_collide(a), *sb = summary_for_collide(b);
    LaunchCtx c;
    if ((rc = lane_join_locked(*L, &c))) return rc;  // after the coarse masks: they are built on the library's stream
    c.any = any;
    GVX_CUDA(launch_collide_map_map(c, a->data, a->type, b->data, b->type, i0, i1, off, float_to_probability(thr), false,
                                    false, nullptr, sa, sb, ca, cb));
    if ((rc = lane_launched_locked(*L))) return rc;
  }
  GVX_CUDA(cudaStreamSynchronize(L->stream));
  if (count) *count = L->result_h->count;
  return GVX_OK;
}

int gvx_lane_collide_map_map(gvx_handle_t lane, gvx_handle_t ha, gvx_handle_t hb, float thr, const int32_t off3[3],
                             uint64_t* count) {
  return lane_collide_map_map(lane, ha, hb, thr, off3, false, count);
}

int gvx_lane_collide_map_map_any(gvx_handle_t lane, gvx_handle_t ha, gvx_handle_t hb, float thr, int* flag) {
  uint64_t c = 0;
  const int rc = lane_collide_map_map(lane, ha, hb, thr, nullptr, true, &c);
  if (rc == GVX_OK && flag) *flag = c != 0;
  return rc;
}

int gvx_collide_map_map_types(gvx_handle_t ha, gvx_handle_t hb, float thr, int mark,
                              uint8_t mask_inout32[GVX_BITVOXEL_BYTES], uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, nullptr, &a, &b, &off, &i0, &i1);
  if (rc) return rc;
  if (a->type != kMapBit || b->type != kMapProb)
    return fail(GVX_ERR_UNSUPPORTED, "collideWithTypes is implemented for BitVectorVoxelMap x ProbVoxelMap");
  GVX_CUDA(launch_collide_map_map(lctx(), a->data, a->type, b->data, b->type, i0, i1, 0, float_to_probability(thr),
                                  mark != 0, true, nullptr,
                                  summary_for_collide(a), summary_for_collide(b)));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  uint64_t red[5] = {g.result_h->count, 0, 0, 0, 0};
  memcpy(red + 1, g.result_h->mask, 32);
  if (shard_active() && a->sharded() && (rc = shard_allreduce(red, 1, 4))) return rc;  // count summed, mask OR-ed
  if (count) *count = red[0];
  if (mask_inout32) {
    const uint8_t* m = (const uint8_t*)(red + 1);
    for (int k = 0; k < 32; ++k) mask_inout32[k] |= m[k];
  }
  return GVX_OK;
}

int gvx_collide_map_map_bitcheck(gvx_handle_t ha, gvx_handle_t hb, uint32_t margin, uint32_t sv_offset, int mark,
                                 uint8_t types_inout32[GVX_BITVOXEL_BYTES], uint64_t* count) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map *a, *b;
  int64_t off, i0, i1;
  int rc = collide_prepare(ha, hb, nullptr, &a, &b, &off, &i0, &i1);
  if (rc) return rc;
  if (a->type != kMapBit || b->type != kMapBit)
    return fail(GVX_ERR_UNSUPPORTED, "the bit-window check needs two BitVectorVoxelMaps");
  // the reference prints an error for a window above 32 and then shifts by a negative amount; 32 itself shifts a
  // 64-bit word by 64. Both are undefined there and refused here. The same for an offset that shifts by >= 64.
  if (margin > 31) return fail(GVX_ERR_INVALID, "window size %u: the reference's 64-bit window supports at most 31", margin);
  if (sv_offset > 0xFFFFu) return fail(GVX_ERR_INVALID, "sv_offset %u exceeds the kernel's uint16", sv_offset);
  const uint32_t high = (sv_offset >> 3) & 0xFFu;
  if (high >= 32 + margin) return fail(GVX_ERR_INVALID, "sv_offset %u shifts the 64-bit window by 64 or more", sv_offset);
  // candidate filters: `other` contributes only swept-volume bits, a subset of its occupancy summary; `this` also
  // contributes bits 0..3, and eBVM_FREE (not in the summary) reaches a swept-volume bit when margin + high >= 4
  const uint32_t* occ_b = summary_for_collide(b);
  const uint32_t* occ_a = (g.use_summary && occ_b && margin + high < 4) ? a->occ_if_valid() : nullptr;
  if (mark) a->touch();  // a marked voxel that held only eBVM_FREE becomes occupied: the kernel updates a's summary
  GVX_CUDA(launch_collide_bitcheck(lctx(), a->data, b->data, i1 - i0, margin, sv_offset, mark != 0, occ_a, occ_b,
                                   a->occ_valid ? a->occ : nullptr, nullptr));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  uint64_t red[5] = {g.result_h->count, 0, 0, 0, 0};
  memcpy(red + 1, g.result_h->mask, 32);
  if (shard_active() && a->sharded() && (rc = shard_allreduce(red, 1, 4))) return rc;  // count summed, mask OR-ed
  if (count) *count = red[0];
  if (types_inout32) {
    const uint8_t* m = (const uint8_t*)(red + 1);
    for (int k = 0; k < 32; ++k) types_inout32[k] |= m[k];
  }
  return GVX_OK;
}

int gvx_map_shift_sv_ids(gvx_handle_t h, uint32_t shift) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (m->type != kMapBit) return fail(GVX_ERR_UNSUPPORTED, "shiftLeftSweptVolumeIDs needs a BitVectorVoxelMap");
  if (shift > 56) return fail(GVX_ERR_INVALID, "Maximum shift size is 56! Higher shift number requested. Not performing shift operation.");
  GVX_CUDA(launch_shift_sv_ids(lctx(), m->data, m->n_voxels(), shift, m->occ));
  m->revalidate();  // the kernel re-derives every summary word
  return GVX_OK;
}

}  // extern "C"
