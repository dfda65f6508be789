// gvx_api_dispatch.cu -- C ABI glue of the sensor-frame dispatch over NVLink peer memory (kernel: gvx_dispatch.cu).
#include "gvx_context.h"

using namespace gvx;

namespace gvx {

static size_t dispatch_bytes(int world, uint64_t capacity) {
  return kDispatchHeadBytes + (size_t)2 * world * capacity * sizeof(uint32_t);
}

static void dispatch_unmap() {
  for (int r = 0; r < g.disp_world; ++r)
    if (r != g.disp_rank && g.disp_ptr[r]) cudaIpcCloseMemHandle(g.disp_ptr[r]);
  for (int r = 0; r < kMaxPeers; ++r) g.disp_ptr[r] = nullptr;
  g.disp_world = 1; g.disp_rank = 0; g.disp_connected = false;
}

void dispatch_free_all() {
  dispatch_unmap();
  if (g.disp_mem) { cudaFree(g.disp_mem); g.disp_mem = nullptr; }
  if (g.disp_err_h) { cudaFreeHost(g.disp_err_h); g.disp_err_h = nullptr; g.disp_err_hmap_d = nullptr; }
  g.disp_capacity = 0; g.disp_world_alloc = 0; g.disp_epoch = 0;
}

static int dispatch_alloc(int world, uint64_t capacity) {
  if (g.disp_mem && g.disp_world_alloc == world && g.disp_capacity == capacity) return GVX_OK;
  if (g.disp_connected) return fail(GVX_ERR_INVALID, "dispatch buffers are mapped by peers: disconnect before resizing");
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  if (g.disp_mem) { cudaFree(g.disp_mem); g.disp_mem = nullptr; }
  const size_t bytes = dispatch_bytes(world, capacity);
  cudaError_t e = cudaMalloc((void**)&g.disp_mem, bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return fail(GVX_ERR_NOMEM, "cudaMalloc(%zu) for the dispatch inboxes failed: %s", bytes, cudaGetErrorString(e)); }
  GVX_CUDA(cudaMemset(g.disp_mem, 0, kDispatchHeadBytes));
  GVX_CUDA(cudaDeviceSynchronize());
  if (!g.disp_err_h) {
    GVX_CUDA(cudaHostAlloc((void**)&g.disp_err_h, 64, cudaHostAllocMapped));
    GVX_CUDA(cudaHostGetDevicePointer((void**)&g.disp_err_hmap_d, g.disp_err_h, 0));
  }
  *g.disp_err_h = 0;
  g.disp_capacity = (uint32_t)capacity;
  g.disp_world_alloc = world;
  g.disp_epoch = 0;
  return GVX_OK;
}

}  // namespace gvx

extern "C" {

int gvx_peer_dispatch_create(int world, uint64_t capacity_points, void* ipc_handle64) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (world < 1 || world > kMaxPeers) return fail(GVX_ERR_INVALID, "world %d outside [1, %d]", world, kMaxPeers);
  if (capacity_points == 0 || capacity_points > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "capacity must be in [1, 2^32)");
  if (!ipc_handle64) return fail(GVX_ERR_INVALID, "null handle buffer");
  int rc = dispatch_alloc(world, capacity_points);
  if (rc) return rc;
  cudaIpcMemHandle_t h;
  GVX_CUDA(cudaIpcGetMemHandle(&h, g.disp_mem));
  memcpy(ipc_handle64, &h, 64);
  return GVX_OK;
}

int gvx_peer_dispatch_connect(int rank, int world, const void* all_handles) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (!g.disp_mem) return fail(GVX_ERR_INVALID, "gvx_peer_dispatch_create has not been called");
  if (world != g.disp_world_alloc || rank < 0 || rank >= world) return fail(GVX_ERR_INVALID, "bad rank %d / world %d (created for %d)", rank, world, g.disp_world_alloc);
  if (world > 1 && !all_handles) return fail(GVX_ERR_INVALID, "null handles");
  if (g.disp_connected) return fail(GVX_ERR_INVALID, "already connected");
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  for (int r = 0; r < world; ++r) {
    if (r == rank) { g.disp_ptr[r] = g.disp_mem; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const char*)all_handles + 64 * r, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      g.disp_world = r; g.disp_rank = rank;  // unmap what has been mapped so far
      dispatch_unmap();
      return fail(GVX_ERR_CUDA, "cudaIpcOpenMemHandle for rank %d: %s", r, cudaGetErrorString(e));
    }
    g.disp_ptr[r] = (unsigned char*)p;
  }
  g.disp_rank = rank;
  g.disp_world = world;
  // disp_epoch continues (it restarts only with a fresh, zeroed allocation in dispatch_alloc): ready[] tags and ack[]
  // credits of the previous session are still in the heads, and a restarted count would match them
  g.disp_connected = true;
  return GVX_OK;
}

int gvx_peer_dispatch_disconnect(void) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (!g.ready) return GVX_OK;
  cudaStreamSynchronize(g.stream);
  dispatch_unmap();
  return GVX_OK;
}

int gvx_update_logodds_dispatch(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind,
                                const float pose12[12], int delta_occ, const uint32_t* z_bounds) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(h);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)h);
  if (m->type != kMapProb) return fail(GVX_ERR_UNSUPPORTED, "insertSensorData needs a ProbVoxelMap");
  if (delta_occ < -127 || delta_occ > 127) return fail(GVX_ERR_INVALID, "delta %d outside int8 range", delta_occ);
  if (n && (!xyz_sensor || !pose12)) return fail(GVX_ERR_INVALID, "null pointer");
  if (g.disp_err_h && *g.disp_err_h) {
    const unsigned e = *g.disp_err_h;
    *g.disp_err_h = 0;
    return fail(GVX_ERR_CUDA, "an earlier sensor-frame dispatch failed (%s)", (e & 1u) ? "a rank did not arrive in time" : "an inbox overflowed");
  }
  const int world = g.disp_connected ? g.disp_world : 1;
  if (!g.disp_connected) {  // nobody to exchange with: the plain update (a slab map keeps its own points)
    if (n == 0) return GVX_OK;
    Affine12 P;
    memcpy(P.a, pose12, 48);
    const float* xyz_d = nullptr;
    int rc = points_to_device(xyz_sensor, n, mem_kind, &xyz_d);
    if (rc) return rc;
    GVX_CUDA(launch_update_logodds(lctx(), m->view(), xyz_d, n, P, delta_occ));
    return GVX_OK;
  }
  // a collective: EVERY rank launches, also with n == 0, in the same order
  if (!z_bounds) return fail(GVX_ERR_INVALID, "z_bounds (world + 1 slab boundaries) is required when ranks are connected");
  if (n > g.disp_capacity) return fail(GVX_ERR_INVALID, "%llu points exceed the dispatch capacity %u", (unsigned long long)n, g.disp_capacity);
  PeerDispatch pd{};
  uint32_t widest = 0;
  for (int r = 0; r <= world; ++r) {
    pd.z_bounds[r] = z_bounds[r];
    if (r && z_bounds[r] < z_bounds[r - 1]) return fail(GVX_ERR_INVALID, "z_bounds must not decrease");
    if (r && z_bounds[r] - z_bounds[r - 1] > widest) widest = z_bounds[r] - z_bounds[r - 1];
  }
  if (z_bounds[0] != 0 || z_bounds[world] != m->dz || z_bounds[g.disp_rank] != m->z0 || z_bounds[g.disp_rank + 1] != m->z1)
    return fail(GVX_ERR_DIM_MISMATCH, "z_bounds do not describe this map: rank %d holds [%u, %u) of %u", g.disp_rank, m->z0, m->z1, m->dz);
  if ((uint64_t)m->dx * m->dy * widest > (1ull << 32))  // slab-local indices 0 .. 2^32 - 1
    return fail(GVX_ERR_UNSUPPORTED, "a slab of more than 2^32 voxels does not fit the 32-bit dispatch entries");
  for (int r = 0; r < world; ++r) pd.base[r] = g.disp_ptr[r];
  pd.capacity = g.disp_capacity;
  pd.rank = g.disp_rank;
  pd.world = world;
  pd.epoch = ++g.disp_epoch;
  Affine12 P{};
  if (pose12) memcpy(P.a, pose12, 48);
  const float* xyz_d = nullptr;
  if (n) {
    int rc = points_to_device(xyz_sensor, n, mem_kind, &xyz_d);
    if (rc) return rc;
  }
  GVX_CUDA(launch_update_logodds_dispatch(lctx(), m->view(), xyz_d, n, P, delta_occ, pd, g.disp_err_hmap_d));
  return GVX_OK;
}

}  // extern "C"
