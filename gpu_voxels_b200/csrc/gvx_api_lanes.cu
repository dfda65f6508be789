// gvx_api_lanes.cu -- collision lanes: read-only collision queries from several host threads that overlap on the GPU.
//
// The reference serialises per MAP (every collide takes the two maps' m_mutex, GV/voxelmap/TemplateVoxelMap.hpp
// collideWith / collisionCheck: "lock_guard guard(this->m_mutex); lock_guard guard2(other->m_mutex)"), so two planner
// threads that check different robot maps against one environment run at the same time. The serial entry points of this
// library share ONE stream, ONE accumulator block and ONE result block under one mutex. A lane
// gives a thread its own stream + accumulators + result blocks:
//
//   * g.mu is held only while the query is validated and ENQUEUED (a few microseconds); the wait for the result happens
//     outside it, so N threads keep N kernels in flight.
//   * Ordering against the serial API is by events, not by locks. A lane launch that follows a serial call makes the
//     lane's stream wait for an event recorded on the library's stream (the lane sees every insert / clear enqueued
//     before the call; with no serial call in between there is nothing new to wait for), and the next serial entry
//     point -- any of them, ApiMutex::lock -- makes the library's stream wait for the lanes launched since (a mutation
//     can never overtake a lane kernel that is still reading the map).
//   * Lanes never write voxels (no `mark`), never take part in a collective (refused under GVX_SHARD=z), and allocate
//     nothing after gvx_lane_create.
#include "gvx_context.h"

namespace gvx {

// under g.mu.m: the library's stream waits for every lane launched since the last fence
static void lanes_fence_locked() {
  for (auto& kv : g.lanes) {
    Lane& L = *kv.second;
    if (!L.pending) continue;
    // the event is recorded here, not after every lane launch: lane launches happen under g.mu too, so "everything on
    // the lane's stream now" covers them all, and the lanes' hot path stays one launch long
    cudaError_t e = cudaEventRecord(L.done_ev, L.stream);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(g.stream, L.done_ev, 0);
    if (e != cudaSuccess) {  // lock() cannot fail: fall back to waiting on the host
      cudaGetLastError();
      cudaStreamSynchronize(L.stream);
    }
    L.pending = false;
  }
  g.lanes_pending = false;
}

void ApiMutex::lock() {
  m.lock();
  ++g.serial_calls;  // a serial entry point may enqueue work on the library's stream: lanes join it again (lane_join_locked)
  if (g.lanes_pending) lanes_fence_locked();
}

static void lane_release(Lane& L) {
  if (L.stream) cudaStreamSynchronize(L.stream);
  if (L.scratch_d) cudaFree(L.scratch_d);
  if (L.result_d) cudaFree(L.result_d);
  if (L.result_h) cudaFreeHost(L.result_h);
  if (L.join_ev) cudaEventDestroy(L.join_ev);
  if (L.done_ev) cudaEventDestroy(L.done_ev);
  if (L.stream) cudaStreamDestroy(L.stream);
  L.stream = nullptr; L.scratch_d = nullptr; L.result_d = L.result_h = L.result_hmap_d = nullptr;
  L.join_ev = L.done_ev = nullptr;
}

// the calling thread may never have touched CUDA: make the library's device current before any runtime call
int lane_thread_device() {
  int d = -1;
  GVX_CUDA(cudaGetDevice(&d));
  if (d != g.device) GVX_CUDA(cudaSetDevice(g.device));
  return GVX_OK;
}

int lane_acquire(gvx_handle_t h, std::shared_ptr<Lane>* out) {
  std::lock_guard<std::mutex> lk(g.mu.m);
  GVX_NEED_INIT();
  auto it = g.lanes.find(h);
  if (it == g.lanes.end()) return fail(GVX_ERR_INVALID, "unknown lane handle %llu", (unsigned long long)h);
  *out = it->second;
  return GVX_OK;
}

// under g.mu.m, before the lane's launch: the lane's stream waits for what the library's stream holds so far
int lane_join_locked(Lane& L, LaunchCtx* c) {
  // ... unless no serial entry point has run since this lane last joined: then the library's stream holds nothing new
  // (every serial call passes ApiMutex::lock, which counts it; gvx_set_stream is one of them)
  if (L.joined_serial_calls != g.serial_calls) {
    GVX_CUDA(cudaEventRecord(L.join_ev, g.stream));
    GVX_CUDA(cudaStreamWaitEvent(L.stream, L.join_ev, 0));
    L.joined_serial_calls = g.serial_calls;
  }
  *c = LaunchCtx{L.stream, g.sm_count, L.scratch_d, L.result_d, L.result_hmap_d, &g.launches, PeerReduce{}};
  c->peer.world = 1;
  c->tma = g.collide_tma;
  c->pdl = false;  // nothing precedes the kernel on the lane's stream but the event wait
  return GVX_OK;
}

// under g.mu.m, after the lane's launch
int lane_launched_locked(Lane& L) {
  L.pending = true;
  g.lanes_pending = true;
  return GVX_OK;
}

// gvx_shutdown, g.mu.m held and the device idle: hand the lanes out; the caller releases them after dropping g.mu
// (lock order lane.mu -> g.mu)
void lanes_detach_all_locked(std::vector<std::shared_ptr<Lane>>* out) {
  for (auto& kv : g.lanes) out->push_back(kv.second);
  g.lanes.clear();
  g.lanes_pending = false;
}
void lanes_release(std::vector<std::shared_ptr<Lane>>& v) {
  for (auto& p : v) {
    std::lock_guard<std::mutex> ll(p->mu);
    p->dead = true;
    lane_release(*p);
  }
  v.clear();
}

}  // namespace gvx

using namespace gvx;

extern "C" {

int gvx_lane_create(gvx_handle_t* lane) {
  if (!lane) return fail(GVX_ERR_INVALID, "null lane pointer");
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  int rc = lane_thread_device();
  if (rc) return rc;
  auto L = std::make_shared<Lane>();
  auto bail = [&](cudaError_t e, const char* what) {
    lane_release(*L);
    return fail(GVX_ERR_CUDA, "gvx_lane_create: %s: %s", what, cudaGetErrorString(e));
  };
  cudaError_t e;
  if ((e = cudaStreamCreateWithFlags(&L->stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
  if ((e = cudaMalloc(&L->scratch_d, sizeof(ReduceScratch))) != cudaSuccess) return bail(e, "accumulators");
  if ((e = cudaMemset(L->scratch_d, 0, sizeof(ReduceScratch))) != cudaSuccess) return bail(e, "accumulators");
  if ((e = cudaMalloc(&L->result_d, sizeof(ResultBlock))) != cudaSuccess) return bail(e, "result block");
  if ((e = cudaMemset(L->result_d, 0, sizeof(ResultBlock))) != cudaSuccess) return bail(e, "result block");
  if ((e = cudaHostAlloc(&L->result_h, sizeof(ResultBlock), cudaHostAllocMapped)) != cudaSuccess) return bail(e, "host result block");
  memset(L->result_h, 0, sizeof(ResultBlock));
  if ((e = cudaHostGetDevicePointer((void**)&L->result_hmap_d, L->result_h, 0)) != cudaSuccess) return bail(e, "host result block");
  if ((e = cudaEventCreateWithFlags(&L->join_ev, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
  if ((e = cudaEventCreateWithFlags(&L->done_ev, cudaEventDisableTiming)) != cudaSuccess) return bail(e, "event");
  if ((e = cudaDeviceSynchronize()) != cudaSuccess) return bail(e, "synchronize");
  const gvx_handle_t h = g.next_handle++;
  g.lanes.emplace(h, L);
  *lane = h;
  return GVX_OK;
}

int gvx_lane_destroy(gvx_handle_t lane) {
  std::shared_ptr<Lane> L;
  {
    std::lock_guard<std::mutex> lk(g.mu.m);
    GVX_NEED_INIT();
    auto it = g.lanes.find(lane);
    if (it == g.lanes.end()) return fail(GVX_ERR_INVALID, "unknown lane handle %llu", (unsigned long long)lane);
    L = it->second;
    if (L->pending) {  // the lane leaves the table: order the library's stream behind its last query now
      if (cudaEventRecord(L->done_ev, L->stream) != cudaSuccess || cudaStreamWaitEvent(g.stream, L->done_ev, 0) != cudaSuccess) {
        cudaGetLastError();
        cudaStreamSynchronize(L->stream);
      }
      L->pending = false;
    }
    g.lanes.erase(it);
  }
  std::lock_guard<std::mutex> ll(L->mu);  // a query still waiting on this lane finishes first
  int rc = lane_thread_device();
  if (rc) return rc;
  L->dead = true;
  lane_release(*L);
  return GVX_OK;
}

int gvx_lane_count(uint32_t* n) {
  std::lock_guard<std::mutex> lk(g.mu.m);
  GVX_NEED_INIT();
  if (n) *n = (uint32_t)g.lanes.size();
  return GVX_OK;
}

}  // extern "C"
