// gvx_api_batch.cu -- C ABI of the batched robot-configuration collision check (include/gvx.h, "batch" section).
#include <vector>
#include "gvx_context.h"

using namespace gvx;

extern "C" {

int gvx_batch_set_scratch_budget(uint64_t bytes) {
  std::lock_guard<ApiMutex> lk(g.mu);
  if (bytes < 4096) return fail(GVX_ERR_INVALID, "scratch budget below 4 KiB");
  g.bat_seen_budget = (size_t)bytes;
  return GVX_OK;
}

}  // extern "C"

namespace {

constexpr uint32_t kAmbCap = 4096;  // configurations whose rounding is re-done on the host, per call

struct DhArgs {  // forward kinematics instead of host transforms
  const float* dh4; const int32_t* joint_types; uint32_t n_links; const float* joint_values;
  const int32_t* link_cloud;  // [n_links] cloud index of each link or -1; NULL: link l <-> cloud l
};

// stage links | joint values | ambiguous counter, launch the chain kernel into out_d (n_cfg x out_stride x 16), wait,
// and redo on the host the configurations whose rounding could not be proven (none, in practice)
int dh_chain_run(const DhArgs& fk, uint32_t n_cfg, uint32_t out_stride, float* out_d, uint32_t* n_ambiguous) {
  if (fk.n_links == 0 || fk.n_links > 256) return fail(GVX_ERR_INVALID, "a DH chain needs 1..256 links");
  if (!fk.dh4 || !fk.joint_types || !fk.joint_values) return fail(GVX_ERR_INVALID, "null DH description");
  const size_t links_b = ((size_t)fk.n_links * sizeof(DhLinkDev) + 15) & ~(size_t)15;
  const size_t jv_b = ((size_t)n_cfg * fk.n_links * 4 + 15) & ~(size_t)15;
  const size_t amb_b = (size_t)(kAmbCap + 4) * 4;
  int rc;
  if ((rc = ensure_dev(&g.bat_fk, &g.bat_fk_cap, links_b + jv_b + amb_b))) return rc;
  if ((rc = ensure_pinned(&g.bat_fk_h, &g.bat_fk_h_cap, links_b + jv_b + amb_b))) return rc;
  DhLinkDev* links_h = (DhLinkDev*)g.bat_fk_h;
  for (uint32_t l = 0; l < fk.n_links; ++l) {
    const int jt = fk.joint_types[l];
    if (jt != 0 && jt != 1) return fail(GVX_ERR_INVALID, "link %u: illegal joint type %d", l, jt);
    const int slot = fk.link_cloud ? fk.link_cloud[l] : (int)l;
    if (slot >= (int)out_stride) return fail(GVX_ERR_INVALID, "link %u: cloud index %d out of range", l, slot);
    links_h[l] = dh_link_prepare(fk.dh4[4 * l], fk.dh4[4 * l + 1], fk.dh4[4 * l + 2], fk.dh4[4 * l + 3], jt, slot);
  }
  memcpy((char*)g.bat_fk_h + links_b, fk.joint_values, (size_t)n_cfg * fk.n_links * 4);
  uint32_t* amb_h = (uint32_t*)((char*)g.bat_fk_h + links_b + jv_b);
  amb_h[0] = 0;
  GVX_CUDA(cudaMemcpyAsync(g.bat_fk, g.bat_fk_h, links_b + jv_b + 16, cudaMemcpyHostToDevice, g.stream));
  uint32_t* amb_d = (uint32_t*)((char*)g.bat_fk + links_b + jv_b);
  GVX_CUDA(launch_dh_chain(lctx(), (const DhLinkDev*)g.bat_fk, fk.n_links, (const float*)((char*)g.bat_fk + links_b),
                           n_cfg, out_d, out_stride, g.dh_check_bits, amb_d, kAmbCap));
  GVX_CUDA(cudaMemcpyAsync(amb_h, amb_d, amb_b, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  const uint32_t n_amb = amb_h[0];
  if (n_ambiguous) *n_ambiguous = n_amb;
  if (n_amb) {
    std::vector<float> row((size_t)out_stride * 16);
    auto redo = [&](uint32_t c) -> int {
      GVX_CUDA(cudaMemcpy(row.data(), out_d + (size_t)c * out_stride * 16, row.size() * 4, cudaMemcpyDeviceToHost));
      dh_chain_host(links_h, fk.n_links, fk.joint_values + (size_t)c * fk.n_links, row.data());
      GVX_CUDA(cudaMemcpy(out_d + (size_t)c * out_stride * 16, row.data(), row.size() * 4, cudaMemcpyHostToDevice));
      return GVX_OK;
    };
    if (n_amb <= kAmbCap) { for (uint32_t i = 0; i < n_amb; ++i) if ((rc = redo(amb_h[1 + i]))) return rc; }
    else { for (uint32_t c = 0; c < n_cfg; ++c) if ((rc = redo(c))) return rc; }  // list overflowed: redo all
  }
  return GVX_OK;
}

int batch_core(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes, uint32_t n_clouds,
               const int32_t* meanings, const float* transforms, const DhArgs* fk, uint32_t n_cfg,
               const uint8_t* mask32, float thr, uint64_t* counts) {
  Map* m = find_map(env);
  if (!m) return fail(GVX_ERR_INVALID, "unknown map handle %llu", (unsigned long long)env);
  if (m->sharded()) return fail(GVX_ERR_UNSUPPORTED, "batch collide shards the BATCH across GPUs; the environment map must be whole");
  if (!cloud_sizes || !meanings || !n_clouds) return fail(GVX_ERR_INVALID, "null/empty cloud description");
  if (n_clouds > 256) return fail(GVX_ERR_UNSUPPORTED, "more than 256 sub-clouds per robot");
  if (n_cfg == 0) return GVX_OK;
  if ((!transforms && !fk) || !counts) return fail(GVX_ERR_INVALID, "null transforms or counts");
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
  const size_t hdr_bytes = (((size_t)n_clouds + 1) * 4 + n_clouds + 15) & ~(size_t)15;  // transforms 16-byte aligned
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
  if (fk) {
    // transforms are produced on the device; clouds without a link keep the identity (KinematicChain.cu:107-112
    // transforms only clouds whose name matches a link)
    GVX_CUDA(cudaMemcpyAsync(g.bat_tr, g.bat_tr_h, hdr_bytes, cudaMemcpyHostToDevice, g.stream));
    float* tr_d = (float*)((char*)g.bat_tr + hdr_bytes);
    bool all_linked = !fk->link_cloud && fk->n_links >= n_clouds;
    if (!all_linked) {
      float* tr_h = (float*)((char*)g.bat_tr_h + hdr_bytes);
      static const float ident[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1};
      for (size_t i = 0; i < (size_t)n_cfg * n_clouds; ++i) memcpy(tr_h + i * 16, ident, 64);
      GVX_CUDA(cudaMemcpyAsync(tr_d, tr_h, tr_bytes, cudaMemcpyHostToDevice, g.stream));
    }
    if (!fk->link_cloud && fk->n_links > n_clouds) return fail(GVX_ERR_INVALID, "more links than clouds needs link_cloud");
    if ((rc = dh_chain_run(*fk, n_cfg, n_clouds, tr_d, nullptr))) return rc;
  } else {
    memcpy((char*)g.bat_tr_h + hdr_bytes, transforms, tr_bytes);
    GVX_CUDA(cudaMemcpyAsync(g.bat_tr, g.bat_tr_h, tr_bytes + hdr_bytes, cudaMemcpyHostToDevice, g.stream));
  }
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

}  // namespace

extern "C" {

int gvx_batch_robot_collide(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes,
                            uint32_t n_clouds, const int32_t* meanings, const float* transforms, uint32_t n_cfg,
                            const uint8_t* mask32, float thr, uint64_t* counts) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  return batch_core(env, base_xyz, mem_kind, cloud_sizes, n_clouds, meanings, transforms, nullptr, n_cfg, mask32, thr,
                    counts);
}

int gvx_batch_robot_collide_dh(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes,
                               uint32_t n_clouds, const int32_t* meanings, const float* dh4,
                               const int32_t* joint_types, uint32_t n_links, const int32_t* link_cloud,
                               const float* joint_values, uint32_t n_cfg, const uint8_t* mask32, float thr,
                               uint64_t* counts) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  DhArgs fk{dh4, joint_types, n_links, joint_values, link_cloud};
  return batch_core(env, base_xyz, mem_kind, cloud_sizes, n_clouds, meanings, nullptr, &fk, n_cfg, mask32, thr, counts);
}

int gvx_dh_chain_transforms(const float* dh4, const int32_t* joint_types, uint32_t n_links, const float* joint_values,
                            uint32_t n_cfg, float* transforms_out, uint32_t* n_ambiguous) {
  std::lock_guard<ApiMutex> lk(g.mu);
  GVX_NEED_INIT();
  if (n_ambiguous) *n_ambiguous = 0;
  if (n_cfg == 0) return GVX_OK;
  if (!transforms_out) return fail(GVX_ERR_INVALID, "null output");
  const size_t bytes = (size_t)n_cfg * n_links * 64;
  int rc;
  if ((rc = ensure_dev(&g.bat_tr, &g.bat_tr_cap, bytes))) return rc;
  DhArgs fk{dh4, joint_types, n_links, joint_values, nullptr};
  if ((rc = dh_chain_run(fk, n_cfg, n_links, (float*)g.bat_tr, n_ambiguous))) return rc;
  GVX_CUDA(cudaMemcpyAsync(transforms_out, g.bat_tr, bytes, cudaMemcpyDeviceToHost, g.stream));
  GVX_CUDA(cudaStreamSynchronize(g.stream));
  return GVX_OK;
}

}  // extern "C"
