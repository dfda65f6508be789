// gvx_context.h -- process-wide context, handle tables and error plumbing shared by the gvx_api*.cu files.
#pragma once
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../include/gvx.h"
#include "gvx_internal.h"

namespace gvx {

struct Map {
  int type;
  uint32_t dx, dy, dz, z0, z1;
  float side;
  void* data;
  bool owned;
  // Summary: 1 bit per voxel, bit i&31 of word i>>5, kept in sync by every mutating entry point (inserts set bits,
  // clears reset or recompute). BitVectorVoxelMap: bit = !noneButEmpty(voxel) (any meaning except eBVM_FREE), exact
  // occupancy, 1/256 of the map. ProbVoxelMap: bit = "known" (occupancy != UNKNOWN_PROBABILITY), a superset of
  // "occupancy >= threshold" for every threshold, 1/8 of the map. Collide kernels read summaries first and touch voxels
  // only where they must. nullptr for wrapped (caller-owned) memory; occ_valid drops when the raw device pointer is
  // handed out.
  uint32_t* occ = nullptr;
  bool occ_valid = false;
  // GVX_SHARD=z: made collectively by gvx_map_create (every rank holds a slab of the same grid, created in the same
  // order); peer_data[r] = rank r's slab mapped through CUDA IPC once an offset collide needed it (gvx_api_shard.cu)
  bool shard_made = false;
  bool peers_open = false;
  void* peer_data[kMaxPeers] = {};
  // ... or connected by the host program for an explicit slab map (gvx_map_connect_peers): world and slab boundaries as
  // given there (peer_world == 0: the GVX_SHARD partition applies)
  int peer_world = 0, peer_rank = 0;
  uint32_t peer_z[kMaxPeers + 1] = {};
  // the writable raw device pointer has been handed out (gvx_map_device_ptr): the caller may write voxels at any time,
  // so NO library call re-validates the summary until gvx_map_refresh_summary says the writes are over
  bool raw_exported = false;
  void revalidate() { touch(); if (occ && !raw_exported) occ_valid = true; }
  bool fallback_logged = false;
  // Coarse mask: 1 bit per kCoarseVoxels (4096) voxels = per 32 consecutive 128-bit words of the summary (what one warp
  // of k_collide_summary2 loads per step); bit clear = that whole chunk of the summary is zero. Built lazily -- on the
  // second collide of a map that has not been modified in between (a robot volume, a static environment) -- and valid
  // only while `version` stands still: every entry point that can change the summary bumps it (touch()).
  uint32_t* coarse = nullptr;
  uint64_t version = 1, coarse_version = 0;
  uint32_t quiet_collides = 0;
  void touch() { ++version; quiet_collides = 0; }
  size_t coarse_bytes() const { return (((occ_bytes() / 16 + 31) / 32 + 31) / 32) * 4 + 16; }
  size_t occ_bytes() const { return ((n_voxels() + 31) / 32 + 3) / 4 * 16; }  // padded to 16 B
  uint32_t* occ_if_valid() const { return occ_valid ? occ : nullptr; }
  uint64_t n_voxels() const { return (uint64_t)dx * dy * (z1 - z0); }
  uint64_t n_total() const { return (uint64_t)dx * dy * dz; }
  bool sharded() const { return z0 != 0 || z1 != dz; }
  size_t voxel_bytes() const { return type == kMapProb ? 1 : kBitVoxelBytes; }
  size_t bytes() const { return n_voxels() * voxel_bytes(); }
  // the view handed to MUTATING kernels: taking it counts as a modification
  MapView view() { touch(); return MapView{data, type, dx, dy, dz, z0, z1, side, occ_valid ? occ : nullptr}; }
};

// BitVectorVoxelList storage: SoA id / coord / voxel like the reference (GV/voxellist/TemplateVoxelList.h:250-252),
// double-buffered so a build reads the old list and writes the merged one; plus the sort scratch.
struct VoxList {
  uint32_t dx = 0, dy = 0, dz = 0;
  float side = 0;
  uint64_t size = 0;
  int cur = 0;
  uint32_t* ids[2] = {nullptr, nullptr};
  uint32_t* coords[2] = {nullptr, nullptr};
  uint8_t* voxels[2] = {nullptr, nullptr};
  uint64_t cap[2] = {0, 0};
  uint32_t *keys_in = nullptr, *keys_out = nullptr, *vals_in = nullptr, *vals_out = nullptr;
  uint8_t* new_meaning = nullptr;
  void* tmp = nullptr;
  size_t tmp_bytes = 0;
  uint64_t scratch_cap = 0, meaning_cap = 0;
  // z-slab maps: the entries whose map voxel lies in a slab are ONE contiguous range of the sorted ids, found with two
  // binary searches on the device and remembered per (list contents, slab, offset). `version` moves with every change
  // of the entries; a list whose id array has been handed out (gvx_list_device_ptrs) is never cached.
  uint64_t version = 1;
  bool ids_exported = false;
  struct { uint64_t version = 0, lo = 0, hi = 0, i0 = 0, i1 = 0; } range;
  uint64_t n_total() const { return (uint64_t)dx * dy * dz; }
};

// One collision lane: its own stream, accumulators and result blocks, so that a query running on it shares nothing
// mutable with the library's stream or another lane. `mu` serialises the callers of ONE lane (held across the wait for
// the result); lock order lane.mu -> g.mu.
struct Lane {
  std::mutex mu;
  bool dead = false;  // destroyed while a caller still held the pointer
  cudaStream_t stream = nullptr;
  ReduceScratch* scratch_d = nullptr;
  ResultBlock* result_d = nullptr;
  ResultBlock* result_h = nullptr;       // pinned, mapped
  ResultBlock* result_hmap_d = nullptr;  // device alias of result_h
  cudaEvent_t join_ev = nullptr;  // recorded on the library's stream at every lane launch: the lane reads what was enqueued before it
  cudaEvent_t done_ev = nullptr;  // recorded on the lane's stream by the next serial call: what the library's stream waits for
  bool pending = false;           // launched since the library's stream last waited for this lane (under g.mu)
  uint64_t joined_serial_calls = ~0ull;  // g.serial_calls when the lane's stream last waited for the library's stream
};

// The library mutex. Every serial entry point takes it through std::lock_guard<ApiMutex>: lock() also orders the
// library's stream behind the lane queries launched since the last serial call, so a mutation enqueued by the caller can
// never overtake a lane kernel that is still reading the map. Lane entry points lock `m` directly (no fence).
struct ApiMutex {
  std::mutex m;
  void lock();  // gvx_api_lanes.cu
  void unlock() { m.unlock(); }
};

struct Context {
  bool ready = false;
  int device = 0;
  int sm_count = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t stream = nullptr;
  ReduceScratch* scratch_d = nullptr;
  ResultBlock* result_d = nullptr;
  ResultBlock* result_h = nullptr;       // pinned, mapped
  ResultBlock* result_hmap_d = nullptr;  // device alias of result_h
  // staging: device buffer for host point clouds, pinned + device buffers for small descriptors
  void* stage_d = nullptr; size_t stage_cap = 0;
  cudaEvent_t stage_ev = nullptr;  // "the H2D copy out of a page-locked GVX_MEM_HOST buffer is done"
  // GVX_MEM_HOST_PINNED: two device slots filled on a copy stream; ready = "slot holds the cloud" (copy stream),
  // consumed = "everything that read the slot has been enqueued before this mark" (main stream)
  cudaStream_t copy_stream = nullptr;
  void* pipe_d[2] = {}; size_t pipe_cap[2] = {};
  cudaEvent_t pipe_ready[2] = {}; cudaEvent_t pipe_consumed[2] = {};
  int pipe_next = 0, pipe_pending = -1;
  // ring of small descriptor slots (pinned host + device + "H2D done" event): no stream sync per call
  static constexpr int kDescSlots = 8;
  void* desc_h[kDescSlots] = {}; void* desc_d[kDescSlots] = {}; size_t desc_cap[kDescSlots] = {};
  cudaEvent_t desc_ev[kDescSlots] = {}; int desc_next = 0;
  // batch-collide scratch (gvx_api_batch.cu): env occupancy bitmap, per-configuration "seen" bitmaps, transforms,
  // counts (device + pinned host)
  void* bat_bits = nullptr; size_t bat_bits_cap = 0;
  void* bat_seen = nullptr; size_t bat_seen_cap = 0;
  void* bat_tr = nullptr; size_t bat_tr_cap = 0;
  void* bat_tr_h = nullptr; size_t bat_tr_h_cap = 0;
  void* bat_counts = nullptr; size_t bat_counts_cap = 0;
  void* bat_counts_h = nullptr; size_t bat_counts_h_cap = 0;
  void* bat_fk = nullptr; size_t bat_fk_cap = 0;      // DH links | joint values | ambiguous list (device)
  void* bat_fk_h = nullptr; size_t bat_fk_h_cap = 0;  // the same, pinned
  int dh_check_bits = 50;  // GVX_OPT_DH_CHECK_BITS: bracket of the double sin/cos before rounding to float
  size_t bat_seen_budget = (size_t)4 << 30;  // bytes of "seen" scratch per wave of configurations
  // fused all-reduce over NVLink peer memory (gvx_peer_*): this rank's mailbox, the peers' mailboxes mapped through
  // CUDA IPC, and the collective sequence number
  PeerSlot* peer_mailbox = nullptr;
  PeerSlot* peer_ptr[kMaxPeers] = {};
  int peer_rank = 0, peer_world = 1;
  unsigned long long peer_epoch = 0;
  unsigned long long peer_vec_epoch = 0;  // collectives that used the mailbox's vector area (PeerReduce::vec_epoch)
  // sensor-frame dispatch over NVLink peer memory (gvx_peer_dispatch_*, gvx_dispatch.cu): this rank's head + inboxes,
  // the peers' allocations mapped through CUDA IPC, the collective sequence number, a mapped host word for time-outs
  unsigned char* disp_mem = nullptr;
  unsigned char* disp_ptr[kMaxPeers] = {};
  uint32_t disp_capacity = 0;
  int disp_world_alloc = 0, disp_rank = 0, disp_world = 1;
  bool disp_connected = false;
  unsigned long long disp_epoch = 0;
  uint32_t* disp_err_h = nullptr; uint32_t* disp_err_hmap_d = nullptr;
  // z-slab sharding behind the API (gvx_api_shard.cu): one process per GPU, GVX_SHARD=z in the environment
  int shard_rank = 0, shard_world = 1;
  bool shard_fused = false;  // the peers' mailboxes are mapped: map x map counts are all-reduced inside the collide kernel
  bool pdl = true;       // GVX_OPT_PDL
  int collide_tma = -1;  // GVX_OPT_COLLIDE_TMA: -1 auto (TMA where it measured faster), 0 never, 1..5 always (shape k)
  bool use_summary = true;  // GVX_OPT_USE_SUMMARY: collide kernels may read occupancy summaries instead of voxels
  bool use_coarse = true;   // GVX_OPT_USE_COARSE: ... and skip 4096-voxel chunks a map's coarse mask says are empty
  unsigned long long* halo_counts_d = nullptr;  // one count per peer slab of an offset collide on z-slab maps (gvx_api.cu collide_halo)
  uint64_t launches = 0;
  uint64_t summary_fallbacks = 0;  // collides that streamed voxels because a library-owned map's summary was invalid
  uint64_t next_handle = 1;
  std::unordered_map<uint64_t, Map> maps;
  std::unordered_map<uint64_t, VoxList> lists;
  struct DevBuffer { void* p; uint64_t bytes; };
  std::unordered_map<uint64_t, DevBuffer> buffers;  // gvx_buffer_*: device-resident point clouds of the C++ facade
  // collision lanes (gvx_api_lanes.cu): read-only collision queries of several host threads that overlap on the GPU
  std::unordered_map<uint64_t, std::shared_ptr<Lane>> lanes;
  bool lanes_pending = false;  // a lane launch since the library's stream last waited for the lanes
  uint64_t serial_calls = 0;   // serial entry points so far (ApiMutex::lock)
  ApiMutex mu;
};

extern Context g;
// gvx_api_lanes.cu
int lane_thread_device();
int lane_acquire(gvx_handle_t h, std::shared_ptr<Lane>* out);
int lane_join_locked(Lane& L, LaunchCtx* c);
int lane_launched_locked(Lane& L);
void lanes_detach_all_locked(std::vector<std::shared_ptr<Lane>>* out);
void lanes_release(std::vector<std::shared_ptr<Lane>>& v);
// gvx_api_shard.cu
int shard_init_from_env();
void shard_detach();
int shard_allreduce(uint64_t* v, int n_sum, int n_or);  // host-side, through the job's shared-memory segment
void shard_slab(uint32_t dz, uint32_t* z0, uint32_t* z1);
void shard_slab_of(int rank, uint32_t dz, uint32_t* z0, uint32_t* z1);
int shard_allgather(const void* mine, size_t bytes, void* all);
int shard_barrier();
int shard_map_open_peers(Map* m);     // collective
void shard_map_close_peers(Map* m);   // local
inline bool shard_active() { return g.shard_world > 1; }
// gvx_api.cu: the peer mailbox entry points without taking g.mu
int mailbox_create_locked(void* ipc_handle64);
int mailbox_connect_locked(int rank, int world, const void* all_handles);
int mailbox_disconnect_locked();
void lists_free_all();  // gvx_api_list.cu
void dispatch_free_all();  // gvx_api_dispatch.cu
extern thread_local std::string t_err;

inline int fail(int code, const char* fmt, ...) {
  char buf[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buf, sizeof buf, fmt, ap);
  va_end(ap);
  t_err = buf;
  return code;
}
#define GVX_CUDA(x)                                                                                   \
  do {                                                                                                \
    cudaError_t e_ = (x);                                                                             \
    if (e_ != cudaSuccess) return fail(GVX_ERR_CUDA, "%s: %s (%s:%d)", #x, cudaGetErrorString(e_), __FILE__, __LINE__); \
  } while (0)
#define GVX_NEED_INIT() \
  do { if (!g.ready) return fail(GVX_ERR_NOT_INITIALIZED, "gvx_init has not been called"); } while (0)

inline LaunchCtx lctx() {
  LaunchCtx c{g.stream, g.sm_count, g.scratch_d, g.result_d, g.result_hmap_d, &g.launches, PeerReduce{}};
  c.peer.world = 1;
  c.tma = g.collide_tma;
  c.pdl = g.pdl;
  return c;
}
// launch context of a counting kernel that all-reduces its count over the connected peers (next epoch)
inline LaunchCtx lctx_allreduce() {
  LaunchCtx c = lctx();
  if (g.peer_world > 1) {
    for (int r = 0; r < g.peer_world; ++r) c.peer.peers[r] = g.peer_ptr[r];
    c.peer.rank = g.peer_rank;
    c.peer.world = g.peer_world;
    c.peer.epoch = ++g.peer_epoch;
  }
  return c;
}
// the summary a collide may use for map m; counts (and logs once per map) the silent cliff where a library-owned map is
// streamed only because its raw pointer was handed out and gvx_map_refresh_summary has not been called since
inline const uint32_t* summary_for_collide(Map* m) {
  if (!g.use_summary) return nullptr;
  if (m->occ && !m->occ_valid) {
    ++g.summary_fallbacks;
    if (!m->fallback_logged) {
      m->fallback_logged = true;
      fprintf(stderr, "[gvx] collide streams the voxels of a %ux%ux%u map: its summary is invalid since the raw device "
                      "pointer was handed out (gvx_map_device_ptr); call gvx_map_refresh_summary after writing\n",
              m->dx, m->dy, m->dz);
    }
  }
  return m->occ_if_valid();
}
// the coarse mask a summary collide may use for map m (nullptr: none). Builds it on the stream when the map has been
// collided twice without a modification in between; never allocates (the mask is allocated with the map).
inline int coarse_for_collide(Map* m, const uint32_t** out) {
  *out = nullptr;
  if (!g.use_summary || !g.use_coarse || !m->coarse || !m->occ || !m->occ_valid) return GVX_OK;
  if (m->coarse_version == m->version) { *out = m->coarse; return GVX_OK; }
  if (++m->quiet_collides < 2) return GVX_OK;
  GVX_CUDA(cudaMemsetAsync(m->coarse, 0, m->coarse_bytes(), g.stream));
  GVX_CUDA(launch_coarse_build(lctx(), m->occ, (long long)(m->occ_bytes() / 16), m->coarse));
  m->coarse_version = m->version;
  ++g.serial_calls;  // work on the library's stream that did not come through a serial entry point: lanes must join it
  *out = m->coarse;
  return GVX_OK;
}
inline Map* find_map(gvx_handle_t h) {
  auto it = g.maps.find(h);
  return it == g.maps.end() ? nullptr : &it->second;
}
inline VoxList* find_list(gvx_handle_t h) {
  auto it = g.lists.find(h);
  return it == g.lists.end() ? nullptr : &it->second;
}
// grow-only device / pinned-host scratch buffers
inline int ensure_dev(void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return GVX_OK;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (*p) { cudaFree(*p); *p = nullptr; *cap = 0; }
  cudaError_t e = cudaMalloc(p, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(GVX_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", bytes, cudaGetErrorString(e)); }
  *cap = bytes;
  return GVX_OK;
}
inline int ensure_pinned(void** p, size_t* cap, size_t bytes) {
  if (bytes <= *cap) return GVX_OK;
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (*p) { cudaFreeHost(*p); *p = nullptr; *cap = 0; }
  GVX_CUDA(cudaHostAlloc(p, bytes, cudaHostAllocDefault));
  *cap = bytes;
  return GVX_OK;
}
inline int ensure_stage(size_t bytes) {
  if (bytes <= g.stage_cap) return GVX_OK;
  if (g.stage_d) { GVX_CUDA(cudaStreamSynchronize(g.stream)); GVX_CUDA(cudaFree(g.stage_d)); g.stage_d = nullptr; g.stage_cap = 0; }
  size_t cap = bytes + bytes / 4 + (1u << 20);
  GVX_CUDA(cudaMalloc(&g.stage_d, cap));
  g.stage_cap = cap;
  return GVX_OK;
}
// Acquire the next descriptor slot with room for `bytes`; waits only if that slot's previous H2D copy (8 calls
// ago) has not finished. Fill *h, then call desc_commit to copy it to *d on the stream.
inline int desc_acquire(size_t bytes, int* slot, void** h, void** d) {
  const int k = g.desc_next;
  g.desc_next = (g.desc_next + 1) % Context::kDescSlots;
  if (!g.desc_ev[k]) GVX_CUDA(cudaEventCreateWithFlags(&g.desc_ev[k], cudaEventDisableTiming));
  GVX_CUDA(cudaEventSynchronize(g.desc_ev[k]));
  if (bytes > g.desc_cap[k]) {
    GVX_CUDA(cudaStreamSynchronize(g.stream));  // a kernel may still read the old device slot
    if (g.desc_h[k]) cudaFreeHost(g.desc_h[k]);
    if (g.desc_d[k]) cudaFree(g.desc_d[k]);
    g.desc_h[k] = g.desc_d[k] = nullptr; g.desc_cap[k] = 0;
    const size_t cap = bytes * 2 + 4096;
    GVX_CUDA(cudaHostAlloc(&g.desc_h[k], cap, cudaHostAllocDefault));
    GVX_CUDA(cudaMalloc(&g.desc_d[k], cap));
    g.desc_cap[k] = cap;
  }
  *slot = k; *h = g.desc_h[k]; *d = g.desc_d[k];
  return GVX_OK;
}
inline int desc_commit(int slot, size_t bytes) {
  GVX_CUDA(cudaMemcpyAsync(g.desc_d[slot], g.desc_h[slot], bytes, cudaMemcpyHostToDevice, g.stream));
  GVX_CUDA(cudaEventRecord(g.desc_ev[slot], g.stream));
  return GVX_OK;
}
// Points to device: host clouds go through the staging buffer (one async H2D, no per-call cudaMalloc/cudaFree
// as in TemplateVoxelMap::insertPointCloud(std::vector), GV/voxelmap/TemplateVoxelMap.hpp:563-576).
inline int points_to_device(const float* xyz, uint64_t n, int mem_kind, const float** out_d) {
  if (mem_kind == GVX_MEM_DEVICE) { *out_d = xyz; return GVX_OK; }
  const size_t bytes = (size_t)n * 12;
  if (mem_kind == GVX_MEM_HOST_PINNED) {
    if (!g.copy_stream) {
      GVX_CUDA(cudaStreamCreateWithFlags(&g.copy_stream, cudaStreamNonBlocking));
      for (int k = 0; k < 2; ++k) {
        GVX_CUDA(cudaEventCreateWithFlags(&g.pipe_ready[k], cudaEventDisableTiming));
        GVX_CUDA(cudaEventCreateWithFlags(&g.pipe_consumed[k], cudaEventDisableTiming));
      }
    }
    // the kernels that read the slot filled by the previous call are on g.stream by now: mark the slot reusable
    if (g.pipe_pending >= 0) { GVX_CUDA(cudaEventRecord(g.pipe_consumed[g.pipe_pending], g.stream)); g.pipe_pending = -1; }
    const int k = g.pipe_next;
    g.pipe_next ^= 1;
    if (bytes > g.pipe_cap[k]) {
      GVX_CUDA(cudaStreamSynchronize(g.copy_stream));
      GVX_CUDA(cudaStreamSynchronize(g.stream));
      if (g.pipe_d[k]) { cudaFree(g.pipe_d[k]); g.pipe_d[k] = nullptr; g.pipe_cap[k] = 0; }
      const size_t cap = bytes + bytes / 4 + (1u << 20);
      cudaError_t e = cudaMalloc(&g.pipe_d[k], cap);
      if (e != cudaSuccess) { cudaGetLastError(); return fail(GVX_ERR_NOMEM, "cudaMalloc(%zu) failed: %s", cap, cudaGetErrorString(e)); }
      g.pipe_cap[k] = cap;
    }
    GVX_CUDA(cudaStreamWaitEvent(g.copy_stream, g.pipe_consumed[k], 0));  // never recorded yet = nothing to wait for
    if (bytes) GVX_CUDA(cudaMemcpyAsync(g.pipe_d[k], xyz, bytes, cudaMemcpyHostToDevice, g.copy_stream));
    GVX_CUDA(cudaEventRecord(g.pipe_ready[k], g.copy_stream));
    GVX_CUDA(cudaStreamWaitEvent(g.stream, g.pipe_ready[k], 0));
    g.pipe_pending = k;
    *out_d = (const float*)g.pipe_d[k];
    return GVX_OK;
  }
  if (mem_kind != GVX_MEM_HOST) return fail(GVX_ERR_INVALID, "bad mem_kind %d", mem_kind);
  int rc = ensure_stage(bytes);
  if (rc) return rc;
  if (bytes) {
    GVX_CUDA(cudaMemcpyAsync(g.stage_d, xyz, bytes, cudaMemcpyHostToDevice, g.stream));
    // GVX_MEM_HOST promises "consumed or staged when the call returns". From pageable memory the runtime has staged the
    // bytes when cudaMemcpyAsync returns; from page-locked memory (cudaHostAlloc / cudaHostRegister / torch pinned) the
    // DMA reads the caller's buffer later, so wait for it here. Callers that want the overlap say GVX_MEM_HOST_PINNED.
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, xyz) != cudaSuccess) { cudaGetLastError(); at.type = cudaMemoryTypeUnregistered; }
    if (at.type == cudaMemoryTypeHost) {
      if (!g.stage_ev) GVX_CUDA(cudaEventCreateWithFlags(&g.stage_ev, cudaEventDisableTiming));
      GVX_CUDA(cudaEventRecord(g.stage_ev, g.stream));
      GVX_CUDA(cudaEventSynchronize(g.stage_ev));
    }
  }
  *out_d = (const float*)g.stage_d;
  return GVX_OK;
}
// DefaultCollider::floatToProbability (GV/voxel/DefaultCollider.hpp:94-98), evaluated on the host like the
// reference does: thr*254 - 127 in fp32, C truncation to int32, low 8 bits.
inline int8_t float_to_probability(float thr) {
  const float tmp = (thr * (float(127) - float(-127))) + float(-127);
  return (int8_t)(int32_t)tmp;
}

}  // namespace gvx
