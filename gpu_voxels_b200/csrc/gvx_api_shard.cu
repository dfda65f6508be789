// gvx_api_shard.cu -- z-slab sharding BEHIND the drop-in API, one process per GPU, no torch / MPI / NCCL needed.
//
// The reference is single-GPU: GpuVoxels::addMap allocates the whole grid on one device (GV/GpuVoxels.cpp:164-290) and
// refuses maps above 4 GiB (GV/voxelmap/TemplateVoxelMap.hpp:65-75). Here the SAME caller code runs as N processes
// (SPMD: every process makes the same calls with the same clouds), started with
//     GVX_SHARD=z GVX_WORLD=N GVX_RANK=r [GVX_LOCAL_RANK=l] [GVX_DEVICES=0,1,..] [GVX_RENDEZVOUS=name]
// (RANK / WORLD_SIZE / LOCAL_RANK of torchrun and OMPI_COMM_WORLD_* of mpirun are understood as well). Then
//   * gvx_init picks the rank's device and joins the job's POSIX shared-memory segment (the rendezvous);
//   * gvx_map_create makes the rank's z-slab of every map (slab [z0,z1) = voxels z0*dx*dy .. z1*dx*dy-1 of
//     the reference layout), inserts keep the points of the slab, lists stay replicated;
//   * every call that returns a collision result returns the GLOBAL result on every rank: the count of
//     map x map collideWith is all-reduced inside the collide kernel over NVLink peer memory (CUDA IPC mailboxes, handles
//     exchanged through the segment); masks, per-meaning counts and the other counts go through a host-side
//     all-reduce in the segment (they are host values of synchronous calls anyway).
// Host code only; the kernels are the ones every other entry point uses.
#include <cerrno>
#include <chrono>
#include <cstdlib>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <thread>
#include <unistd.h>

#include "gvx_context.h"

using namespace gvx;

namespace gvx {

constexpr uint64_t kShardMagic = 0x6776785f73686172ull;  // "gvx_shar"
constexpr int kShardWords = 272;                          // 1 + 256 + 8 result words and some room

struct ShardSeg {
  volatile uint64_t magic;  // kShardMagic once rank 0 has zeroed and initialised the segment
  volatile uint64_t world;
  volatile uint64_t arrived[kMaxPeers];  // pid of rank r once it has attached
  volatile uint64_t acked[kMaxPeers];    // rank 0 echoes arrived[r]: "you are in the segment of THIS job, not a stale one"
  volatile uint64_t seq[kMaxPeers];      // sequence number of the last all-reduce rank r has published
  uint64_t data[2][kMaxPeers][kShardWords];
  unsigned char ipc[kMaxPeers][64];
};

static ShardSeg* s_seg = nullptr;
static std::string s_name;
static uint64_t s_seq = 0;

static const char* env_first(std::initializer_list<const char*> names) {
  for (const char* n : names) {
    const char* v = getenv(n);
    if (v && *v) return v;
  }
  return nullptr;
}

static bool wait_until(const std::function<bool()>& cond, double seconds) {
  const auto t0 = std::chrono::steady_clock::now();
  for (unsigned spin = 0;; ++spin) {
    if (cond()) return true;
    if (spin > 2000) std::this_thread::sleep_for(std::chrono::microseconds(50));
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > seconds) return false;
  }
}

static int shard_attach(int rank, int world, const std::string& name) {
  const std::string path = "/" + name;
  const size_t bytes = sizeof(ShardSeg);
  const uint64_t me = (uint64_t)getpid();
  if (rank == 0) {
    shm_unlink(path.c_str());  // a segment left behind by a crashed job of the same name
    const int fd = shm_open(path.c_str(), O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) return fail(GVX_ERR_IO, "shm_open(%s) failed: %s", path.c_str(), strerror(errno));
    if (ftruncate(fd, (off_t)bytes) != 0) { close(fd); return fail(GVX_ERR_IO, "ftruncate(%s) failed: %s", path.c_str(), strerror(errno)); }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return fail(GVX_ERR_IO, "mmap(%s) failed: %s", path.c_str(), strerror(errno));
    s_seg = (ShardSeg*)p;
    memset((void*)s_seg, 0, bytes);
    s_seg->world = (uint64_t)world;
    s_seg->arrived[0] = s_seg->acked[0] = me;
    __sync_synchronize();
    s_seg->magic = kShardMagic;
    // acknowledge every rank that shows up
    for (int r = 1; r < world; ++r) {
      if (!wait_until([&] { return s_seg->arrived[r] != 0; }, 120.0))
        return fail(GVX_ERR_IO, "rank %d of %d never joined the rendezvous %s", r, world, path.c_str());
      s_seg->acked[r] = s_seg->arrived[r];
    }
    __sync_synchronize();
    return GVX_OK;
  }
  const auto t0 = std::chrono::steady_clock::now();
  for (;;) {
    if (std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count() > 120.0)
      return fail(GVX_ERR_IO, "rank %d could not join the rendezvous %s (is rank 0 running?)", rank, path.c_str());
    const int fd = shm_open(path.c_str(), O_RDWR, 0600);
    if (fd < 0) { std::this_thread::sleep_for(std::chrono::milliseconds(2)); continue; }
    struct stat st;
    if (fstat(fd, &st) != 0 || (size_t)st.st_size < bytes) { close(fd); std::this_thread::sleep_for(std::chrono::milliseconds(2)); continue; }
    void* p = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    close(fd);
    if (p == MAP_FAILED) return fail(GVX_ERR_IO, "mmap(%s) failed: %s", path.c_str(), strerror(errno));
    ShardSeg* seg = (ShardSeg*)p;
    // a stale segment of an earlier job has no rank 0 that answers: leave it and look again
    bool ok = wait_until([&] { return seg->magic == kShardMagic; }, 1.0) && seg->world == (uint64_t)world;
    if (ok) {
      seg->arrived[rank] = me;
      __sync_synchronize();
      ok = wait_until([&] { return seg->acked[rank] == me; }, 3.0);
    }
    if (ok) { s_seg = seg; return GVX_OK; }
    munmap(p, bytes);
    std::this_thread::sleep_for(std::chrono::milliseconds(20));
  }
}

// sum of the first n_sum words, OR of the next n_or words, over all ranks; in place; every rank gets the result
int shard_allreduce(uint64_t* v, int n_sum, int n_or) {
  if (g.shard_world <= 1) return GVX_OK;
  if (!s_seg) return fail(GVX_ERR_NOT_INITIALIZED, "sharding is not connected");
  const int n = n_sum + n_or;
  if (n > kShardWords) return fail(GVX_ERR_INVALID, "host all-reduce of %d words (max %d)", n, kShardWords);
  const uint64_t e = ++s_seq;
  const int par = (int)(e & 1);
  // double-buffered by parity: a rank starts epoch e + 1 only after every rank has PUBLISHED e, and publishes e + 2 (same
  // buffer as e) only after every rank has published e + 1, i.e. has finished reading e
  for (int k = 0; k < n; ++k) s_seg->data[par][g.shard_rank][k] = v[k];
  __sync_synchronize();
  s_seg->seq[g.shard_rank] = e;
  for (int r = 0; r < g.shard_world; ++r)
    if (!wait_until([&] { return s_seg->seq[r] >= e; }, 120.0))
      return fail(GVX_ERR_IO, "host all-reduce: rank %d did not arrive (every rank must make the same calls)", r);
  __sync_synchronize();
  for (int k = 0; k < n; ++k) {
    uint64_t acc = 0;
    for (int r = 0; r < g.shard_world; ++r) {
      const uint64_t x = s_seg->data[par][r][k];
      acc = k < n_sum ? acc + x : (acc | x);
    }
    v[k] = acc;
  }
  return GVX_OK;
}

void shard_slab(uint32_t dz, uint32_t* z0, uint32_t* z1) {
  const uint32_t world = (uint32_t)g.shard_world, rank = (uint32_t)g.shard_rank;
  const uint32_t base = dz / world, rem = dz % world;
  *z0 = rank * base + (rank < rem ? rank : rem);
  *z1 = *z0 + base + (rank < rem ? 1u : 0u);
}

void shard_slab_of(int rank_, uint32_t dz, uint32_t* z0, uint32_t* z1) {
  const uint32_t world = (uint32_t)g.shard_world, rank = (uint32_t)rank_;
  const uint32_t base = dz / world, rem = dz % world;
  *z0 = rank * base + (rank < rem ? rank : rem);
  *z1 = *z0 + base + (rank < rem ? 1u : 0u);
}

// every rank's `bytes` bytes, in rank order, on every rank (the all-reduce's buffers and sequence numbers)
int shard_allgather(const void* mine, size_t bytes, void* all) {
  if (g.shard_world <= 1) { memcpy(all, mine, bytes); return GVX_OK; }
  if (!s_seg) return fail(GVX_ERR_NOT_INITIALIZED, "sharding is not connected");
  const size_t words = (bytes + 7) / 8;
  if (words > (size_t)kShardWords) return fail(GVX_ERR_INVALID, "host all-gather of %zu bytes (max %d)", bytes, kShardWords * 8);
  const uint64_t e = ++s_seq;
  const int par = (int)(e & 1);
  uint64_t buf[kShardWords] = {};
  memcpy(buf, mine, bytes);
  for (size_t k = 0; k < words; ++k) s_seg->data[par][g.shard_rank][k] = buf[k];
  __sync_synchronize();
  s_seg->seq[g.shard_rank] = e;
  for (int r = 0; r < g.shard_world; ++r)
    if (!wait_until([&] { return s_seg->seq[r] >= e; }, 120.0))
      return fail(GVX_ERR_IO, "host all-gather: rank %d did not arrive (every rank must make the same calls)", r);
  __sync_synchronize();
  for (int r = 0; r < g.shard_world; ++r) {
    for (size_t k = 0; k < words; ++k) buf[k] = s_seg->data[par][r][k];
    memcpy((unsigned char*)all + (size_t)r * bytes, buf, bytes);
  }
  return GVX_OK;
}

int shard_barrier() {
  uint64_t z = 0;
  return shard_allreduce(&z, 1, 0);
}

// Collective (every rank, same map, same order): map every peer's slab of map m into this process through CUDA IPC, so
// that a collide with an offset can read the voxels that live in a neighbour's slab (the halo) straight over NVLink.
// Opened on the first such collide of the map and kept until the map is destroyed.
int shard_map_open_peers(Map* m) {
  if (m->peers_open) return GVX_OK;
  if (!shard_active() || !m->shard_made || !m->owned)
    return fail(GVX_ERR_UNSUPPORTED, "offset collide on a z-slab map needs the peers' slabs: only maps made by "
                                     "gvx_map_create under GVX_SHARD=z can be mapped (replicas only otherwise)");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
  unsigned char mine[64] = {}, all[kMaxPeers][64];
  uint64_t bad = 0;
  cudaIpcMemHandle_t hd;
  if (cudaIpcGetMemHandle(&hd, m->data) != cudaSuccess) { cudaGetLastError(); bad = 1; }
  else memcpy(mine, &hd, 64);
  int rc = shard_allgather(mine, 64, all);
  if (rc) return rc;
  if ((rc = shard_allreduce(&bad, 1, 0))) return rc;
  if (!bad) {
    for (int r = 0; r < g.shard_world && !bad; ++r) {
      if (r == g.shard_rank) { m->peer_data[r] = m->data; continue; }
      cudaIpcMemHandle_t hr;
      memcpy(&hr, all[r], 64);
      void* p = nullptr;
      if (cudaIpcOpenMemHandle(&p, hr, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { cudaGetLastError(); bad = 1; }
      else m->peer_data[r] = p;
    }
    if ((rc = shard_allreduce(&bad, 1, 0))) return rc;  // all ranks agree; also: nobody proceeds before all have mapped
  }
  if (bad) {
    shard_map_close_peers(m);
    return fail(GVX_ERR_UNSUPPORTED, "a rank could not map a peer's slab through CUDA IPC (offset collides on z-slab maps "
                                     "need peer access between the ranks' GPUs)");
  }
  m->peers_open = true;
  return GVX_OK;
}

// local: drop the mappings of the peers' slabs (the owner frees its slab only after a barrier, gvx_map_destroy)
void shard_map_close_peers(Map* m) {
  for (int r = 0; r < kMaxPeers; ++r) {
    if (m->peer_data[r] && m->peer_data[r] != m->data) cudaIpcCloseMemHandle(m->peer_data[r]);
    m->peer_data[r] = nullptr;
  }
  m->peers_open = false;
  m->peer_world = 0;
}

void shard_detach() {
  if (s_seg) {
    munmap((void*)s_seg, sizeof(ShardSeg));
    s_seg = nullptr;
    if (g.shard_rank == 0 && !s_name.empty()) shm_unlink(("/" + s_name).c_str());
  }
  g.shard_world = 1; g.shard_rank = 0; s_seq = 0;
}

// called by gvx_init (g.mu held, context ready): join the job when the environment asks for it
int shard_init_from_env() {
  int rank, world, rc;
  if (g.shard_world > 1) {  // joined explicitly (gvx_shard_connect_host) before gvx_init: only the mailboxes are missing
    rank = g.shard_rank; world = g.shard_world;
  } else {
    const char* mode = getenv("GVX_SHARD");
    if (!mode || strcmp(mode, "z") != 0) return GVX_OK;
    const char* w = env_first({"GVX_WORLD", "WORLD_SIZE", "OMPI_COMM_WORLD_SIZE"});
    const char* r = env_first({"GVX_RANK", "RANK", "OMPI_COMM_WORLD_RANK"});
    world = w ? atoi(w) : 1; rank = r ? atoi(r) : 0;
    if (world <= 1) return GVX_OK;
    if (world > kMaxPeers || rank < 0 || rank >= world) return fail(GVX_ERR_INVALID, "GVX_SHARD=z: bad rank %d / world %d", rank, world);
    const char* rv = getenv("GVX_RENDEZVOUS");
    const char* port = getenv("MASTER_PORT");
    s_name = rv && *rv ? rv : (std::string("gvx_") + (port && *port ? port : "default"));
    if ((rc = shard_attach(rank, world, s_name))) return rc;
    g.shard_rank = rank;
    g.shard_world = world;
  }
  // peer mailboxes for the in-kernel all-reduce of map x map counts: every rank publishes its CUDA IPC handle in the
  // segment; all ranks agree (host all-reduce) on whether every mapping succeeded, else counts go through the host too
  const char* fz = getenv("GVX_SHARD_FUSED");
  uint64_t ok = (fz && atoi(fz) == 0) ? 0 : 1;
  if (ok && mailbox_create_locked(s_seg->ipc[rank]) != GVX_OK) ok = 0;
  uint64_t bad = ok ? 0 : 1;
  if ((rc = shard_allreduce(&bad, 1, 0))) return rc;  // also the barrier behind which all handles are visible
  if (!bad) {
    unsigned char handles[kMaxPeers][64];
    memcpy(handles, s_seg->ipc, sizeof handles);
    bad = mailbox_connect_locked(rank, world, handles) == GVX_OK ? 0 : 1;
    if ((rc = shard_allreduce(&bad, 1, 0))) return rc;
    if (bad) mailbox_disconnect_locked();
  }
  g.shard_fused = !bad;
  return GVX_OK;
}

}  // namespace gvx

extern "C" {

int gvx_shard_info(int* rank, int* world, int* fused) {
  if (rank) *rank = g.shard_rank;
  if (world) *world = g.shard_world;
  if (fused) *fused = g.shard_fused ? 1 : 0;
  return GVX_OK;
}

int gvx_shard_allreduce_u64(uint64_t* values, int n_sum, int n_or) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (!values && n_sum + n_or > 0) return fail(GVX_ERR_INVALID, "null values");
  return shard_allreduce(values, n_sum, n_or);
}

int gvx_shard_allgather(const void* mine, uint32_t bytes, void* all) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if ((!mine || !all) && bytes > 0) return fail(GVX_ERR_INVALID, "null buffer");
  return shard_allgather(mine, bytes, all);
}

// ---- explicit slab maps: the host program exchanges the handles with its own communicator -------------------------------
int gvx_map_ipc_handle(gvx_handle_t hm, void* handle64_out) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(hm);
  if (!m || !handle64_out) return fail(GVX_ERR_INVALID, "unknown map handle or null output");
  if (!m->owned) return fail(GVX_ERR_UNSUPPORTED, "wrapped (caller-owned) memory: export it with cudaIpcGetMemHandle yourself");
  cudaIpcMemHandle_t hd;
  GVX_CUDA(cudaIpcGetMemHandle(&hd, m->data));
  memcpy(handle64_out, &hd, 64);
  return GVX_OK;
}

int gvx_map_connect_peers(gvx_handle_t hm, int rank, int world, const void* all_handles, const uint32_t* z_bounds) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(hm);
  if (!m || !all_handles || !z_bounds) return fail(GVX_ERR_INVALID, "unknown map handle or null argument");
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return fail(GVX_ERR_INVALID, "bad rank %d / world %d", rank, world);
  if (m->peers_open) return fail(GVX_ERR_INVALID, "the map's peers are connected already");
  if (z_bounds[0] != 0 || z_bounds[world] != m->dz || z_bounds[rank] != m->z0 || z_bounds[rank + 1] != m->z1)
    return fail(GVX_ERR_INVALID, "z_bounds do not describe this map: slab [%u,%u) of %u layers as rank %d of %d", m->z0, m->z1, m->dz, rank, world);
  for (int r = 0; r < world; ++r)
    if (z_bounds[r] > z_bounds[r + 1]) return fail(GVX_ERR_INVALID, "z_bounds must not decrease");
  for (int r = 0; r < world; ++r) {
    if (r == rank) { m->peer_data[r] = m->data; continue; }
    cudaIpcMemHandle_t hr;
    memcpy(&hr, (const unsigned char*)all_handles + 64 * (size_t)r, 64);
    void* p = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&p, hr, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      cudaGetLastError();
      shard_map_close_peers(m);
      return fail(GVX_ERR_UNSUPPORTED, "cudaIpcOpenMemHandle(rank %d's slab) failed: %s", r, cudaGetErrorString(e));
    }
    m->peer_data[r] = p;
  }
  m->peer_world = world;
  m->peer_rank = rank;
  for (int r = 0; r <= world; ++r) m->peer_z[r] = z_bounds[r];
  m->peers_open = true;
  return GVX_OK;
}

int gvx_map_disconnect_peers(gvx_handle_t hm) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(hm);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle");
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  shard_map_close_peers(m);
  return GVX_OK;
}

int gvx_shard_connect_host(int rank, int world, const char* rendezvous) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (g.shard_world > 1) return fail(GVX_ERR_INVALID, "already part of a sharded job");
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world) return fail(GVX_ERR_INVALID, "bad rank %d / world %d", rank, world);
  if (!rendezvous || !*rendezvous) return fail(GVX_ERR_INVALID, "empty rendezvous name");
  if (world == 1) return GVX_OK;
  s_name = rendezvous;
  int rc = shard_attach(rank, world, s_name);
  if (rc) return rc;
  g.shard_rank = rank;
  g.shard_world = world;
  g.shard_fused = false;
  return GVX_OK;
}

int gvx_shard_disconnect(void) {
  std::lock_guard<ApiMutex> lk(g.mu);
  shard_detach();
  return GVX_OK;
}

}  // extern "C"
