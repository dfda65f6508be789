// gvx_api_batch.cu -- C ABI of the batched robot-configuration collision check (include/gvx.h, "batch" section).
#include "gvx_context.h"

using namespace gvx;

extern "C" {

int gvx_batch_set_scratch_budget(uint64_t bytes) {
  std::lock_guard<std::mutex> lk(g.mu);
  if (bytes < 4096) return fail(GVX_ERR_INVALID, "scratch budget below 4 KiB");
  g.bat_seen_budget = (size_t)bytes;
  return GVX_OK;
}

int gvx_batch_robot_collide(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes,
                            uint32_t n_clouds, const int32_t* meanings, const float* transforms, uint32_t n_cfg,
                            const uint8_t* mask32, float thr, uint64_t* counts) {
  std::lock_guard<std::mutex> lk(g.mu);
  GVX_NEED_INIT();
  Map* m = find_map(env);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)env);
  if (m->sharded()) return fail(GVX_ERR_UNSUPPORTED, "batch collide shards the BATCH across GPUs; the environment map must be whole");
  if (!cloud_sizes || !meanings || !n_clouds) return fail(GVX_ERR_INVALID, "null/empty cloud description");
  if (n_clouds > 256) return fail(GVX_ERR_UNSUPPORTED, "more than 256 sub-clouds per robot");
  if (n_cfg == 0) return GVX_OK;
  if (!transforms || !counts) return fail(GVX_ERR_INVALID, "null transforms or counts");
  uint64_t n_pts = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) n_pts += cloud_sizes[c];
  if (n_pts > 0xFFFFFFFFull) return fail(GVX_ERR_INVALID, "MetaPointCloud larger than 2^32-1 points");
  for (uint32_t c = 0; c < n_clouds; ++c)
    if (meanings[c] < 0 || meanings[c] > 255) return fail(GVX_ERR_INVALID, "meaning %d out of range", meanings[c]);
  if (n_pts == 0) { memset(counts, 0, (size_t)n_cfg * 8); return GVX_OK; }
  if (!base_xyz) return fail(GVX_ERR_INVALID, "null point pointer");

  const uint64_t n_vox = m->n_total();
  const uint64_t words = (n_vox + 31) / 32;
  int rc;
  if ((rc = ensure_dev(&g.bat_bits, &g.bat_bits_cap, words * 4))) return rc;
  // configurations per wave: as many "seen" bitmaps as fit the budget
  uint64_t wave = g.bat_seen_budget / (words * 4);
  if (wave < 1) wave = 1;
  if (wave > n_cfg) wave = n_cfg;
  if (wave > 65535) wave = 65535;
  if ((rc = ensure_dev(&g.bat_seen, &g.bat_seen_cap, wave * words * 4))) return rc;
  const size_t tr_bytes = (size_t)n_cfg * n_clouds * 64;
  const size_t hdr_bytes = ((size_t)n_clouds + 1) * 4 + ((n_clouds + 3) & ~3u);
  if ((rc = ensure_dev(&g.bat_tr, &g.bat_tr_cap, tr_bytes + hdr_bytes))) return rc;
  if ((rc = ensure_pinned(&g.bat_tr_h, &g.bat_tr_h_cap, tr_bytes + hdr_bytes))) return rc;
  if ((rc = ensure_dev(&g.bat_counts, &g.bat_counts_cap, (size_t)n_cfg * 8))) return rc;
  if ((rc = ensure_pinned(&g.bat_counts_h, &g.bat_counts_h_cap, (size_t)n_cfg * 8))) return rc;

  // descriptor: prefix | eligible | transforms, one pinned -> device copy (the previous call has been synchronised
  // before it returned, so the pinned buffer is free)
  uint32_t* prefix = (uint32_t*)g.bat_tr_h;
  uint8_t* elig = (uint8_t*)g.bat_tr_h + ((size_t)n_clouds + 1) * 4;
  uint32_t acc = 0;
  for (uint32_t c = 0; c < n_clouds; ++c) {
    prefix[c] = acc; acc += cloud_sizes[c];
    const int mg = meanings[c];
    // mask == NULL: DefaultCollider, the robot voxel counts if any bit other than eBVM_FREE is set
    // (GV/helpers/BitVector.h:183-198); mask != NULL: (mask & voxel) != 0 (VoxelListOperations.hpp:286)
    elig[c] = mask32 ? ((mask32[mg >> 3] >> (mg & 7)) & 1) : (mg != 0);
  }
  prefix[n_clouds] = acc;
  memcpy((char*)g.bat_tr_h + hdr_bytes, transforms, tr_bytes);
  GVX_CUDA(cudaMemcpyAsync(g.bat_tr, g.bat_tr_h, tr_bytes + hdr_bytes, cudaMemcpyHostToDevice, g.stream));
  const float* xyz_d = nullptr;
  if ((rc = points_to_device(base_xyz, n_pts, mem_kind, &xyz_d))) return rc;

  // environment occupancy, 1 bit per voxel: a BitVectorVoxelMap's summary is exactly that; otherwise build it
  const uint32_t* env_bits = (g.use_summary && m->type == kMapBit) ? m->occ_if_valid() : nullptr;
  if (!env_bits) {
    GVX_CUDA(launch_env_bitmap(lctx(), m->data, m->type, n_vox, float_to_probability(thr), (uint32_t*)g.bat_bits));
    env_bits = (const uint32_t*)g.bat_bits;
  }
  GVX_CUDA(cudaMemsetAsync(g.bat_counts, 0, (size_t)n_cfg * 8, g.stream));
  BatchDesc d;
  d.prefix_d = (const uint32_t*)g.bat_tr;
  d.eligible_d = (const uint8_t*)g.bat_tr + ((size_t)n_clouds + 1) * 4;
  d.transforms_d = (const float*)((const char*)g.bat_tr + hdr_bytes);
  d.n_clouds = n_clouds;
  d.dx = m->dx; d.dy = m->dy; d.dz = m->dz; d.side = m->side;
  for (uint32_t c0 = 0; c0 < n_cfg; c0 += (uint32_t)wave) {
    const uint32_t w = (uint32_t)(n_cfg - c0 < wave ? n_cfg - c0 : wave);
    GVX_CUDA(cudaMemsetAsync(g.bat_seen, 0, (size_t)w * words * 4, g.stream));
    GVX_CUDA(launch_batch_collide(lctx(), xyz_d, d, (uint32_t)n_pts, env_bits, (uint32_t*)g.bat_seen,
                                  words, c0, w, (unsigned long long*)g.bat_counts));
  }
  GVX_CUDA(cudaMemcpyAsync(g.bat_counts_h, g.bat_counts, (size_t)n_cfg * 8, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  memcpy(counts, g.bat_counts_h, (size_t)n_cfg * 8);
  return GVX_OK;
}

}  // extern "C"
