// gvx_internal.h -- host-side launcher prototypes shared between the kernel files and gvx_api.cu.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "gvx_device.cuh"

namespace gvx {

struct LaunchCtx {
  cudaStream_t stream;
  int sm_count;
  ReduceScratch* scratch_d;   // device
  ResultBlock* result_d;      // device copy of the result block
  ResultBlock* result_hmap_d; // device alias of the pinned, mapped host result block
  uint64_t* launches;         // kernel launch counter (host)
  PeerReduce peer;            // world <= 1: no fused all-reduce
  int tma = 0;                // GVX_OPT_COLLIDE_TMA: -1 auto, 0 never, k > 0 = TMA bulk tiles of shape k for every pair
  bool pdl = false;           // GVX_OPT_PDL: programmatic dependent launch for the per-step kernels (pdl_launch)
  bool any = false;           // boolean collisionCheck: kernels leave as soon as any CTA has found a collision
};

// Programmatic dependent launch. A kernel started through pdl_launch with pdl = true may become resident while the
// kernel before it in the stream is still in its tail (its CTAs have passed griddepcontrol.launch_dependents or left);
// such a kernel executes pdl_wait() (gvx_reduce.cuh) before its first global access, which returns once the
// predecessor grid has completed and its writes are visible -- stream order for memory, minus the launch ramp.
// Only kernels that begin with pdl_wait() may be launched with pdl = true.
template <typename... KArgs, typename... Args>
inline cudaError_t pdl_launch(void (*kernel)(KArgs...), dim3 grid, int block, cudaStream_t stream, bool pdl,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3((unsigned)block);
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl ? 1u : 0u;
  return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

constexpr int kSameThreshold = 1000;  // thr_b_or_same: side b uses side a's threshold (the one-threshold DefaultCollider)
// collideWith map x map. `thr` applies to a ProbabilisticVoxel on side a, thr_b_or_same (an int8 value) to one on side b
// (DefaultCollider(Probability, Probability), GV/voxel/DefaultCollider.hpp:45-49). a/b: device base pointers of the voxel arrays; a is indexed at i + off.
// [i_begin, i_end) is the range of i for which 0 <= i + off < n. count_out_d: optional extra device
// destination for the 64-bit count (async API), may be nullptr.
cudaError_t launch_collide_map_map(const LaunchCtx& c, void* a, int a_type, const void* b, int b_type,
                                   int64_t i_begin, int64_t i_end, int64_t off, int8_t thr, bool mark,
                                   bool types, unsigned long long* count_out_d, const uint32_t* occ_a = nullptr,
                                   const uint32_t* occ_b = nullptr, const uint32_t* coarse_a = nullptr,
                                   const uint32_t* coarse_b = nullptr, int thr_b_or_same = kSameThreshold);
// coarse[c >> 5] bit c & 31 = "some bit of summary quads [32 c, 32 c + 32) is set"; coarse must be zeroed before
cudaError_t launch_coarse_build(const LaunchCtx& c, const uint32_t* occ, long long n_quads, uint32_t* coarse);

// the voxel-streaming collide fed by TMA bulk tiles (gvx_collide_tma.cu); applicable: off == 0, whole range, at least
// one BitVectorVoxel side, 16-byte aligned bases, no meaning mask
bool collide_tma_applicable(const void* a, int a_type, const void* b, int b_type, int64_t i_begin, int64_t off,
                            bool types);
cudaError_t launch_collide_tma(const LaunchCtx& c, void* a, int a_type, const void* b, int b_type, int64_t n,
                               int thr_pack, bool mark, unsigned long long* count_out_d);

// voxelisation
struct MetaDesc {             // device-resident description of a MetaPointCloud insert
  const uint32_t* prefix_d;   // n_clouds + 1 exclusive prefix of cloud sizes
  const int32_t* meanings_d;  // n_meanings
  const float* transforms_d;  // n_clouds * 16 or nullptr
  uint32_t n_clouds;
  uint32_t n_meanings;
};
struct MapView {
  void* data;
  int type;
  uint32_t dx, dy, dz, z0, z1;
  float side;
  uint32_t* occ;  // occupancy summary to keep in sync (BitVectorVoxelMap with a valid summary), else nullptr
};
cudaError_t launch_insert_points(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n, int meaning,
                                 bool want_flag);
cudaError_t launch_insert_meta(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                               const MetaDesc& d, bool want_flag);
struct Affine12 { float a[12]; };  // row-major [R|t], passed to kernels by value (no H2D copy, no sync)
cudaError_t launch_transform(const LaunchCtx& c, const float* in_d, float* out_d, uint64_t n, const Affine12& M);
// robot_d (nullable): data of a WHOLE BitVectorVoxelMap of the same dimensions; points whose hit voxel has robot_bit
// set there are dropped (cut_real_robot)
cudaError_t launch_update_logodds(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                  const Affine12& pose, int delta, const void* robot_d = nullptr,
                                  uint32_t robot_bit = 0);
// the free-space pass of the ray caster: delta_free (< 0) on every cell from the sensor's voxel to the hit voxel
cudaError_t launch_raycast_free(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                const Affine12& pose, int delta_free, const void* robot_d, uint32_t robot_bit);
constexpr int kMaxFramesByValue = 16;
struct FramesByValue {  // a small multi-frame description passed in the kernel parameters
  uint32_t prefix[kMaxFramesByValue + 1];
  float poses[kMaxFramesByValue * 12];
  uint8_t order[kMaxFramesByValue];  // blockIdx.y -> frame: frames whose sensor sits inside this slab are scheduled first
};
// by_value != nullptr (n_frames <= kMaxFramesByValue): prefix_d / poses_d are ignored
cudaError_t launch_update_logodds_multi(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                        const uint32_t* prefix_d, const float* poses_d, uint32_t n_frames, int delta,
                                        const FramesByValue* by_value = nullptr);
// one frame per rank on a z-slab map; hits are routed to the owning rank inside the kernel (gvx_dispatch.cu)
cudaError_t launch_update_logodds_dispatch(const LaunchCtx& c, const MapView& m, const float* xyz_d, uint64_t n,
                                           const Affine12& pose, int delta, const PeerDispatch& pd,
                                           uint32_t* error_hmap_d);
// voxel &= ~mask for all voxels; occ (nullable): summary words rewritten from the result
cudaError_t launch_clear_bits(const LaunchCtx& c, void* data, uint64_t n_voxels, const uint32_t mask8[8], uint32_t* occ);


// BitVectorVoxelList build (gvx_list.cu): old list (sorted, unique) + n_new points -> out_* (sorted, unique, OR-merged);
// the number of entries lands in ResultBlock::count. All pointers are device pointers.
struct ListBuild {
  const uint32_t* old_ids; const uint32_t* old_coords; const uint8_t* old_voxels; uint64_t old_size;
  const float* xyz_d; uint64_t n_new;
  // merge (TemplateVoxelList::merge): n_new entries of ANOTHER list instead of points when add_ids != nullptr; their
  // ids / coordinates are shifted by add_id_off / add_coord_off (uint32 wrap-around like the reference's operator),
  // their voxels replaced by the single bit add_meaning when it is >= 0
  const uint32_t* add_ids = nullptr; const uint32_t* add_coords = nullptr; const uint8_t* add_voxels = nullptr;
  uint32_t add_id_off = 0; int32_t add_coord_off[3] = {0, 0, 0}; int add_meaning = -1;
  uint32_t dx, dy; float side;
  MetaDesc meta;        // prefix_d == nullptr: every point gets single_meaning
  int single_meaning;
  uint32_t *keys_in, *keys_out, *vals_in, *vals_out;  // old_size + n_new entries each
  uint8_t* new_meaning;                               // n_new bytes
  void* tmp; size_t tmp_bytes;                        // >= list_sort_temp_bytes(old_size + n_new)
  uint32_t* out_ids; uint32_t* out_coords; uint8_t* out_voxels;
};
size_t list_sort_temp_bytes(uint64_t n);
cudaError_t launch_list_build(const LaunchCtx& c, const ListBuild& b);
cudaError_t launch_collide_list_map(const LaunchCtx& c, int mode, const uint32_t* ids, void* voxels, uint64_t size,
                                    const void* map, int map_type, uint64_t n_total, uint64_t map_first,
                                    uint64_t map_n, int8_t thr_i, float thr_f, int64_t off, bool mark,
                                    const uint32_t mask_in8[8], unsigned long long* count_out_d,
                                    const uint32_t* occ = nullptr);

// i0 = lower_bound(ids, lo) -> ResultBlock::count, i1 = lower_bound(ids, hi) -> ResultBlock::per_meaning[0]
cudaError_t launch_list_bounds(const LaunchCtx& c, const uint32_t* ids, uint64_t size, uint64_t lo, uint64_t hi);

// list x list (gvx_listlist.cu): mode 0 count of common ids, 1 + OR of both voxels, 2 bit check with margin,
// 3 per-meaning histogram of a's voxels, 4 keep_d[i] = (a's id i is NOT in b). Results in the result block.
cudaError_t launch_list_list(const LaunchCtx& c, int mode, const uint32_t* a_ids, const void* a_vox, uint64_t na,
                             const uint32_t* b_ids, const void* b_vox, uint64_t nb, uint32_t margin, uint32_t* keep_d);
cudaError_t launch_exclusive_sum(const LaunchCtx& c, const uint32_t* in_d, uint32_t* out_d, uint64_t n, void* tmp,
                                 size_t tmp_bytes);
cudaError_t launch_list_compact(const LaunchCtx& c, const uint32_t* ids, const uint32_t* coords, const void* vox,
                                uint64_t n, const uint32_t* keep_d, const uint32_t* pos_d, uint32_t* o_ids,
                                uint32_t* o_coords, void* o_vox);
cudaError_t launch_list_diff(const LaunchCtx& c, const uint32_t* a_ids, const uint32_t* a_co, const void* a_vox,
                             const uint32_t* b_ids, const uint32_t* b_co, const void* b_vox, uint64_t n);

// batch of robot configurations vs one environment map (gvx_batch.cu)
struct BatchDesc {
  const uint32_t* prefix_d;    // n_clouds + 1 exclusive prefix of the link cloud sizes
  const uint8_t* eligible_d;   // n_clouds: 1 if the link's meaning passes the collision mask
  const float* transforms_d;   // n_cfg x n_clouds x 16 (row-major Matrix4f)
  uint32_t n_clouds;
  uint32_t dx, dy, dz;
  float side;
};
cudaError_t launch_env_bitmap(const LaunchCtx& c, const void* env, int env_type, uint64_t n, int8_t thr,
                              uint32_t* bits_d);
cudaError_t launch_batch_collide(const LaunchCtx& c, const float* xyz_d, const BatchDesc& d, uint32_t n_points,
                                 const uint32_t* env_bits_d, uint32_t* seen_d, uint64_t seen_words, uint32_t cfg0,
                                 uint32_t n_cfg_wave, unsigned long long* counts_d);

// swept-volume window check and ID shift (gvx_sweptvolume.cu). occ_a/occ_b: summaries usable as candidate filters
// (nullable), occ_a_mark: summary of `a` to keep in sync when a voxel is marked (nullable)
cudaError_t launch_collide_bitcheck(const LaunchCtx& c, void* a, const void* b, int64_t n, uint32_t margin,
                                    uint32_t sv_offset, bool mark, const uint32_t* occ_a, const uint32_t* occ_b,
                                    uint32_t* occ_a_mark, unsigned long long* count_out_d);
cudaError_t launch_shift_sv_ids(const LaunchCtx& c, void* data, uint64_t n_voxels, uint32_t shift, uint32_t* occ);

// batched DH forward kinematics (gvx_kinematics.cu)
struct DhLinkDev {
  float d, theta, a, ca, sa;  // ca/sa = (float)cos/sin((double)alpha), evaluated once per link on the host with libm
  int32_t joint_type;         // 0 REVOLUTE, 1 PRISMATIC (GV/robot/dh_robot/KinematicLink.h:37-41)
  int32_t out_slot;           // index of the link's cloud in the output row, < 0: none
  int32_t pad_;
};
DhLinkDev dh_link_prepare(float d, float theta, float a, float alpha, int joint_type, int out_slot);
// one configuration on the host with the reference's arithmetic; out: out_stride x 16 (rows addressed by out_slot)
void dh_chain_host(const DhLinkDev* links, uint32_t n_links, const float* joint_values, float* out);
// ambiguous_d: [0] = number of configurations whose float rounding could not be proven, [1..amb_cap] = their indices
cudaError_t launch_dh_chain(const LaunchCtx& c, const DhLinkDev* links_d, uint32_t n_links, const float* joint_values_d,
                            uint32_t n_cfg, float* out_d, uint32_t out_stride, int check_bits, uint32_t* ambiguous_d,
                            uint32_t amb_cap);

}  // namespace gvx
