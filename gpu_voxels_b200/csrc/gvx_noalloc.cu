// gvx_noalloc.cu -- gvx_batch_robot_collide_noalloc: the planner batch (gvx_batch.cu, BASELINE.json configs[2]) for
// callers that cannot allocate -- code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// No counterpart in the reference.
//
// gvx_batch_robot_collide keeps grow-only cudaMalloc'd scratch in the context: the environment's 1-bit-per-voxel
// bitmap, one "seen" bitmap per configuration of a wave, the staged transforms and counts. Here ALL scratch is one
// file-scope __device__ array of kNoAllocScratchBytes (16 MiB, held by every GPU that loads the library until the
// process exits), the small descriptors (link prefix, eligibility) travel in the kernel parameters, and the caller
// passes the inputs and the output as device pointers. The call enqueues one memset, at most one bitmap kernel and one
// collide kernel on the library's stream and returns; it never calls cudaMalloc / cudaFree / cudaHostAlloc or any
// allocating library routine, and it does not synchronise.
//
// In HOST code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address comes from
// cudaGetSymbolAddress, once per device with that device current (scratch_ptr below), and is what the kernels get.
//
// Every call shares the array and finds the previous call's contents in it: the part that must start at zero (the
// "seen" bitmaps) is cleared on the stream first, the environment bitmap is rewritten in full. Two calls that overlap on
// the GPU (different streams via gvx_set_stream, or different host threads racing on one device) would race on it:
// calls must be ordered on one stream, which is what the library's single stream + mutex give by default.
#include "gvx_context.h"
#include "gvx_reduce.cuh"

using namespace gvx;

namespace gvx {

constexpr size_t kNoAllocScratchBytes = (size_t)16 << 20;
constexpr int kNoAllocMaxClouds = 16;
constexpr int kNoAllocMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

struct NoAllocDesc {  // by value in the kernel parameters: nothing to stage
  uint32_t prefix[kNoAllocMaxClouds + 1];  // exclusive prefix of the link cloud sizes
  uint8_t eligible[kNoAllocMaxClouds];     // 1: points of this link can collide (meaning vs. mask, see the entry point)
  uint32_t n_clouds;
  uint32_t dx, dy, dz;
  float side;
};

// device address of the scratch array on the CURRENT device, resolved once per device
static int scratch_ptr(unsigned char** out) {
  static unsigned char* ptr[kNoAllocMaxDevices] = {};
  int dev = -1;
  GVX_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) return fail(GVX_ERR_UNSUPPORTED, "device ordinal %d outside [0, %d)", dev, kNoAllocMaxDevices);
  if (!ptr[dev]) {
    void* p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess || !p) {
      cudaGetLastError();
      return fail(GVX_ERR_CUDA, "cudaGetSymbolAddress(g_noalloc_scratch) on device %d: %s", dev, cudaGetErrorString(e));
    }
    ptr[dev] = (unsigned char*)p;
  }
  *out = ptr[dev];
  return GVX_OK;
}

// k_batch_collide (gvx_batch.cu) with the descriptors in the parameters and the transforms read from the caller's
// device array: grid = (chunks, configurations); fused transform -> voxel index -> bit test in the environment bitmap ->
// atomicOr on the configuration's "seen" bitmap; the thread that flips a bit counts the voxel.
__global__ void __launch_bounds__(kThreads)
k_batch_collide_noalloc(const float* __restrict__ xyz, const float* __restrict__ transforms, const NoAllocDesc d,
                        const uint32_t* __restrict__ env_bits, uint32_t* __restrict__ seen,
                        unsigned long long seen_words, unsigned long long* __restrict__ counts) {
  __shared__ float sh_T[kNoAllocMaxClouds * 12];
  __shared__ uint32_t sh_cnt[kThreads / 32];
  const uint32_t cfg = blockIdx.y;
  const float* T = transforms + (size_t)cfg * d.n_clouds * 16;
  for (uint32_t k = threadIdx.x; k < d.n_clouds * 12; k += kThreads) sh_T[k] = T[(k / 12) * 16 + (k % 12)];
  __syncthreads();
  uint32_t* my_seen = seen + (size_t)cfg * seen_words;
  const uint32_t n = d.prefix[d.n_clouds];
  uint32_t cnt = 0;
  for (uint32_t p = blockIdx.x * kThreads + threadIdx.x; p < n; p += gridDim.x * kThreads) {
    const uint32_t c = find_cloud(d.prefix, d.n_clouds, p);
    if (!d.eligible[c]) continue;
    const float x = __ldg(xyz + 3 * (size_t)p), y = __ldg(xyz + 3 * (size_t)p + 1), z = __ldg(xyz + 3 * (size_t)p + 2);
    const float* M = sh_T + 12 * c;
    const float tx = affine_row(M, x, y, z), ty = affine_row(M + 4, x, y, z), tz = affine_row(M + 8, x, y, z);
    const uint32_t cx = coord_to_voxel(tx, d.side), cy = coord_to_voxel(ty, d.side), cz = coord_to_voxel(tz, d.side);
    if (cx >= d.dx || cy >= d.dy || cz >= d.dz) continue;
    const unsigned long long idx = ((unsigned long long)cz * d.dy + cy) * d.dx + cx;
    const uint32_t bit = 1u << (idx & 31);
    if (!(__ldg(env_bits + (idx >> 5)) & bit)) continue;
    const uint32_t old = atomicOr(my_seen + (idx >> 5), bit);
    cnt += !(old & bit);
  }
  const uint32_t wsum = __reduce_add_sync(0xffffffffu, cnt);
  if ((threadIdx.x & 31) == 0) sh_cnt[threadIdx.x >> 5] = wsum;
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned long long tot = 0;
#pragma unroll
    for (int w = 0; w < kThreads / 32; ++w) tot += sh_cnt[w];
    if (tot) atomicAdd(counts + cfg, tot);
  }
}

}  // namespace gvx

extern "C" {

uint64_t gvx_noalloc_scratch_bytes(void) { return kNoAllocScratchBytes; }

int gvx_batch_robot_collide_noalloc(gvx_handle_t env_map, const float* xyz_d, const uint32_t* cloud_sizes,
                                    uint32_t n_clouds, const int32_t* meanings, const float* transforms_d,
                                    uint32_t n_cfg, const uint8_t* mask32, float thr, uint64_t* counts_d) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(env_map);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)env_map);
  if (m->sharded()) return fail(GVX_ERR_UNSUPPORTED, "batch collide shards the BATCH across GPUs; the environment map must be whole");
  if (!cloud_sizes || !meanings || n_clouds == 0) return fail(GVX_ERR_INVALID, "null/empty cloud description");
  if (n_clouds > (uint32_t)kNoAllocMaxClouds)
    return fail(GVX_ERR_INVALID, "%u link clouds exceed the no-alloc variant's maximum of %d (use gvx_batch_robot_collide)", n_clouds, kNoAllocMaxClouds);
  if (n_cfg == 0) return GVX_OK;
  if (n_cfg > 65535u) return fail(GVX_ERR_INVALID, "%u configurations exceed the no-alloc variant's maximum of 65535 per call", n_cfg);
  if (!counts_d || !transforms_d) return fail(GVX_ERR_INVALID, "null transforms / counts pointer");
  NoAllocDesc d{};
  uint64_t acc = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) {
    if (meanings[c] < 0 || meanings[c] > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meanings[c]);
    d.prefix[c] = (uint32_t)acc;
    acc += cloud_sizes[c];
    // the rule of gvx_batch_robot_collide: without a mask (DefaultCollider) a robot voxel counts if any bit other than
    // eBVM_FREE is set (GV/helpers/BitVector.h:183-198); with a mask, if (mask & voxel) != 0 (VoxelListOperations.hpp:286)
    const int mg = meanings[c];
    const bool el = mask32 ? ((mask32[mg >> 3] >> (mg & 7)) & 1u) != 0 : (mg != GVX_BVM_FREE);
    d.eligible[c] = el ? 1 : 0;
  }
  if (acc > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "MetaPointCloud larger than 2^32-1 points");
  d.prefix[n_clouds] = (uint32_t)acc;
  d.n_clouds = n_clouds;
  d.dx = m->dx; d.dy = m->dy; d.dz = m->dz; d.side = m->side;
  if (acc && !xyz_d) return fail(GVX_ERR_INVALID, "null point pointer");
  // scratch layout: [environment bitmap | n_cfg "seen" bitmaps], every part a multiple of 256 bytes
  const uint64_t n_vox = m->n_voxels();
  const uint64_t words = ((n_vox + 31) / 32 + 63) / 64 * 64;
  const uint64_t need = (uint64_t)(1 + (uint64_t)n_cfg) * words * 4;
  if (need > kNoAllocScratchBytes)
    return fail(GVX_ERR_INVALID,
                "gvx_batch_robot_collide_noalloc: %u configurations of a %ux%ux%u grid need %llu bytes of scratch, the "
                "static array holds %llu (1 + n_cfg bitmaps of %llu bytes); split the batch or use gvx_batch_robot_collide",
                n_cfg, m->dx, m->dy, m->dz, (unsigned long long)need, (unsigned long long)kNoAllocScratchBytes,
                (unsigned long long)(words * 4));
  unsigned char* scratch = nullptr;
  int rc = scratch_ptr(&scratch);
  if (rc) return rc;
  uint32_t* env_bits = (uint32_t*)scratch;
  uint32_t* seen = env_bits + words;
  GVX_CUDA(cudaMemsetAsync(counts_d, 0, (size_t)n_cfg * sizeof(uint64_t), g.stream));
  if (acc == 0) return GVX_OK;
  // the last call's "seen" bits are still there: clear what this call uses
  GVX_CUDA(cudaMemsetAsync(seen, 0, (size_t)n_cfg * words * 4, g.stream));
  // the environment bitmap is rewritten in full (DefaultCollider's test of the environment voxel: BitVector ->
  // !noneButEmpty, Prob -> occupancy >= floatToProbability(thr)); a valid BitVector summary IS that bitmap
  const uint32_t* bits = (g.use_summary && m->type == kMapBit) ? m->occ_if_valid() : nullptr;
  if (!bits) {
    GVX_CUDA(launch_env_bitmap(lctx(), m->data, m->type, n_vox, float_to_probability(thr), env_bits));
    bits = env_bits;
  }
  const uint32_t n_points = (uint32_t)acc;
  uint32_t chunks = (n_points + kThreads * 8 - 1) / (kThreads * 8);
  const uint32_t want = (uint32_t)((g.sm_count * 32 + n_cfg - 1) / n_cfg);
  if (chunks > want) chunks = want;
  if (chunks < 1) chunks = 1;
  ++g.launches;
  k_batch_collide_noalloc<<<dim3(chunks, n_cfg), kThreads, 0, g.stream>>>(xyz_d, transforms_d, d, bits, seen, words,
                                                                          (unsigned long long*)counts_d);
  GVX_CUDA(cudaGetLastError());
  return GVX_OK;
}

}  // extern "C"
