/* gvx.h -- C ABI of the B200-native gpu-voxels hot path (libgvx.so).
 *
 * This is the drop-in boundary. The reference (Changliu52/gpu-voxels) has no FFI: its boundary is the
 * public C++ class API (GpuVoxels, voxelmap::BitVectorVoxelMap, voxelmap::ProbVoxelMap,
 * voxellist::BitVectorVoxelList). include/gpu_voxels/ keeps that C++ surface source-compatible for
 * the path; every method there is a few lines over the functions below. Each entry point cites the
 * reference interface it replaces ("GV/" = packages/gpu_voxels/src/gpu_voxels/ in the reference).
 *
 * Conventions
 *  - extern "C", plain pointers and sizes, opaque 64-bit handles, int status (0 = GVX_OK). Nothing
 *    throws, nothing calls exit(); gvx_last_error() returns the message of the last failure on the
 *    calling thread.
 *  - Pointer arguments are HOST pointers unless the name ends in _d or a mem_kind argument says otherwise
 *    (GVX_MEM_HOST / GVX_MEM_DEVICE).
 *  - Voxel layout is the reference's and is part of the contract (GV/voxelmap/TemplateVoxelMap.h:257-259):
 *    linear index = z*dx*dy + y*dx + x; BitVectorVoxel = 32 bytes, bit k in byte k>>3 at position k&7
 *    (GV/helpers/BitVector.h:223-228); ProbabilisticVoxel = 1 signed byte of log-odds
 *    (GV/voxel/ProbabilisticVoxel.h:34-115). Indices are 64-bit, so maps beyond the reference's 4 GiB limit
 *    (GV/voxelmap/TemplateVoxelMap.hpp:65-75) are allowed.
 *  - All work is enqueued on ONE CUDA stream per process (gvx_set_stream lets the host share it, e.g. with
 *    torch). Calls that return a host value synchronise that stream; calls whose result pointers are NULL, and
 *    the *_async variants (results into device memory), do not.
 *  - A map may be a z-slab [z0,z1) of a larger grid (multi-GPU sharding: one process per GPU, each owning a
 *    slab; not in the reference). Slab maps hold voxels z0*dx*dy .. z1*dx*dy-1 only.
 *  - One device per process. Thread-safe: a process-wide mutex serialises calls (the reference takes
 *    per-map recursive mutexes around fully synchronous calls, GV/GpuVoxelsMap.h:267-269).
 */
#ifndef GVX_H_INCLUDED
#define GVX_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef uint64_t gvx_handle_t;

enum gvx_status {
  GVX_OK = 0,
  GVX_ERR_INVALID = 1,      /* bad argument / unknown handle */
  GVX_ERR_CUDA = 2,         /* a CUDA runtime call failed (message in gvx_last_error) */
  GVX_ERR_NOMEM = 3,
  GVX_ERR_UNSUPPORTED = 4,  /* e.g. meaning != eBVM_OCCUPIED into a ProbVoxelMap, offset on a slab map */
  GVX_ERR_DIM_MISMATCH = 5, /* list x map with different reference dimensions (reference returns SSIZE_MAX) */
  GVX_ERR_NOT_INITIALIZED = 6,
  GVX_ERR_IO = 7            /* file cannot be opened, is truncated, or the write was short */
};

/* values of the reference's MapType enum (GV/helpers/common_defines.h:59-73); also the on-disk tag */
enum gvx_map_type {
  GVX_MT_BITVECTOR_VOXELMAP = 0,
  GVX_MT_BITVECTOR_VOXELLIST = 1,
  GVX_MT_PROBAB_VOXELMAP = 4
};

/* BitVoxelMeaning (GV/helpers/common_defines.h:50-58) */
enum gvx_meaning {
  GVX_BVM_FREE = 0, GVX_BVM_OCCUPIED = 1, GVX_BVM_COLLISION = 2, GVX_BVM_UNKNOWN = 3,
  GVX_BVM_SWEPT_VOLUME_START = 4, GVX_BVM_SWEPT_VOLUME_END = 254, GVX_BVM_UNDEFINED = 255
};

enum gvx_mem_kind {
  GVX_MEM_HOST = 0,   /* any host memory; the call has consumed or staged it when it returns (for page-locked memory the
                       * call waits for the upload; say GVX_MEM_HOST_PINNED to keep it asynchronous) */
  /* device memory, read by kernels on the library's stream (gvx_set_stream): the data must be complete in that
   * stream's order -- produced on the same stream, or the producing stream synchronised / joined by an event first */
  GVX_MEM_DEVICE = 1,
  /* page-locked host memory that the caller leaves untouched until the work has completed (gvx_sync, a synchronous call,
   * or an event of its own on the stream): the upload runs on a separate copy stream into one of two device staging
   * slots, so the copy of call k+1 overlaps the kernels of call k (the reference stages every host cloud with a
   * synchronous cudaMalloc + cudaMemcpy + cudaFree, GV/voxelmap/TemplateVoxelMap.hpp:563-576) */
  GVX_MEM_HOST_PINNED = 2
};

#define GVX_BITVOXEL_BYTES 32

/* ---- runtime ------------------------------------------------------------------------------------------ */
/* replaces GpuVoxels::GpuVoxels -> cuTestAndInitDevice (GV/GpuVoxels.cpp:33-42, GV/helpers/cuda_handling.cu:102-130).
 * device < 0: use the current CUDA device. Idempotent. Fails (no exit(123)) when no GPU is present. */
int gvx_init(int device);
int gvx_shutdown(void);
const char* gvx_last_error(void);
/* cudaStream_t to enqueue on (NULL = the library's own non-blocking stream). */
int gvx_set_stream(void* cuda_stream);
int gvx_sync(void);
/* Everything enqueued by later calls waits for cuda_event (a cudaEvent_t recorded on ANOTHER stream): how a
 * GVX_MEM_DEVICE buffer produced on the caller's stream is ordered before the kernels that read it. */
int gvx_wait_event(void* cuda_event);
/* getDeviceInfos / getDeviceMemoryInfo (GV/helpers/cuda_handling.cu:66-149) */
int gvx_device_info(int* sm_count, uint64_t* free_bytes, uint64_t* total_bytes, char* name64);
/* number of kernels this library has launched since gvx_init (bench.py reports it as gpu_launches) */
uint64_t gvx_launch_count(void);
/* number of collide calls that streamed the voxels of a library-owned map only because its summary was invalid (raw
 * pointer handed out and gvx_map_refresh_summary not called since); the first one per map also warns on stderr */
uint64_t gvx_summary_fallback_count(void);

/* cudaLimitMaxL2FetchGranularity (32 / 64 / 128 bytes; 0 = query only): how much DRAM the L2 fetches around a missing
 * 32-byte sector. The gather kernels (list x map, random inserts) touch one sector per access. */
int gvx_set_l2_fetch_granularity(int bytes, int* previous);

/* ---- device buffers ----------------------------------------------------------------------------------- */
/* Device-resident storage for point clouds: what PointCloud / MetaPointCloud hold in the reference
 * (cudaMalloc + cudaMemcpy in GV/helpers/PointCloud.cu:32-120, GV/helpers/MetaPointCloud.cu:24-140). The device
 * pointer can be passed wherever a mem_kind == GVX_MEM_DEVICE cloud is accepted. */
int gvx_buffer_create(gvx_handle_t* h, uint64_t bytes);
int gvx_buffer_destroy(gvx_handle_t h);
int gvx_buffer_upload(gvx_handle_t h, uint64_t offset, const void* host, uint64_t bytes);
int gvx_buffer_download(gvx_handle_t h, uint64_t offset, void* host, uint64_t bytes);
int gvx_buffer_device_ptr(gvx_handle_t h, void** data_d, uint64_t* bytes);

/* ---- dense maps --------------------------------------------------------------------------------------- */
/* GpuVoxels::addMap(MT_BITVECTOR_VOXELMAP | MT_PROBAB_VOXELMAP) -> TemplateVoxelMap ctor + clearMap
 * (GV/GpuVoxels.cpp:164-290, GV/voxelmap/TemplateVoxelMap.hpp:54-138). */
int gvx_map_create(gvx_handle_t* h, int map_type, uint32_t dx, uint32_t dy, uint32_t dz, float side);
/* same, but this process holds only z in [z0,z1) of the dx*dy*dz grid */
int gvx_map_create_slab(gvx_handle_t* h, int map_type, uint32_t dx, uint32_t dy, uint32_t dz, float side,
                        uint32_t z0, uint32_t z1);
/* non-owning ctor TemplateVoxelMap(Voxel* dev_data, ...) (GV/voxelmap/TemplateVoxelMap.hpp:139-148);
 * data_d must be 32-byte aligned; not cleared */
int gvx_map_wrap(gvx_handle_t* h, int map_type, void* data_d, uint32_t dx, uint32_t dy, uint32_t dz, float side,
                 uint32_t z0, uint32_t z1);
/* GpuVoxels::delMap (GV/GpuVoxels.cpp:292-302) / ~TemplateVoxelMap */
int gvx_map_destroy(gvx_handle_t h);
/* clearMap: memset 0x00 (Bit) / 0x80 (Prob) (GV/voxelmap/TemplateVoxelMap.hpp:205-246). Asynchronous. */
int gvx_map_clear(gvx_handle_t h);
/* clearBitVoxelMeaning: BitVoxelMap::clearBit (GV/voxelmap/BitVoxelMap.hpp:57-71,228-232); ProbVoxelMap accepts only
 * eBVM_OCCUPIED and then clears the map (GV/voxelmap/ProbVoxelMap.hpp:110-116), else GVX_ERR_UNSUPPORTED */
int gvx_map_clear_meaning(gvx_handle_t h, int meaning);
/* BitVoxelMap::clearBits (GV/voxelmap/BitVoxelMap.hpp:73-81) */
int gvx_map_clear_bits(gvx_handle_t h, const uint8_t mask32[GVX_BITVOXEL_BYTES]);
/* getDeviceDataPtr / getVoidDeviceDataPtr, getDimensions, getVoxelSideLength, getMemoryUsage */
int gvx_map_device_ptr(gvx_handle_t h, void** data_d, uint64_t* first_voxel, uint64_t* n_voxels);
/* getConstVoidDeviceDataPtr / getConstDeviceDataPtr: a read-only view; unlike gvx_map_device_ptr it leaves the map's
 * summary valid */
int gvx_map_device_ptr_const(gvx_handle_t h, const void** data_d, uint64_t* first_voxel, uint64_t* n_voxels);
int gvx_map_info(gvx_handle_t h, int* map_type, uint32_t dims3[3], float* side, uint32_t* z0, uint32_t* z1,
                 uint64_t* bytes);
/* raw voxel bytes of this process's slab, reference layout (the payload of writeToDisk/readFromDisk,
 * GV/voxelmap/TemplateVoxelMap.hpp:666-799) */
int gvx_map_download(gvx_handle_t h, void* host_buf);
int gvx_map_upload(gvx_handle_t h, const void* host_buf);

/* Occupancy summary. Every library-owned BitVectorVoxelMap carries 1 bit per voxel = "any meaning except eBVM_FREE
 * set" (BitVector::noneButEmpty negated, GV/helpers/BitVector.h:183-198), 1/256 of the map's size, kept in sync by
 * every mutating call of this API. Collision kernels read it instead of the 32-byte voxels whenever occupancy is all
 * they need; results are identical. Handing out the raw device pointer (gvx_map_device_ptr with data_d != NULL)
 * invalidates it, because the caller may then write voxels directly -- and keeps it invalid across every later call
 * (clearMap included) until gvx_map_refresh_summary (one streaming pass) declares the direct writes finished; until
 * then the kernels read the voxels. gvx_map_device_ptr_const does not invalidate. Wrapped maps have no summary. */
int gvx_map_refresh_summary(gvx_handle_t h);
/* test / debug: the summary words of this process's slab (bit i&31 of word i>>5 = voxel i) and whether it is valid */
int gvx_map_summary_download(gvx_handle_t h, uint32_t* words, uint64_t n_words, int* valid);
/* test / debug: the coarse mask (GVX_OPT_USE_COARSE; bit c&31 of word c>>5 = voxels [4096 c, 4096 c + 4096) of this
 * slab) and whether it is currently valid (built, and the map not modified since) */
int gvx_map_coarse_download(gvx_handle_t h, uint32_t* words, uint64_t n_words, int* valid);
enum gvx_option {
  GVX_OPT_USE_SUMMARY = 1,   /* 0: always stream the voxels (the reference's access pattern) */
  GVX_OPT_DH_CHECK_BITS = 2, /* forward kinematics: sin/cos brackets of |v| * 2^-bits (default 50); smaller values send
                                more configurations through the host re-computation (tests) */
  GVX_OPT_COLLIDE_TMA = 3,   /* the voxel-streaming collide kernels can be fed by TMA bulk tiles (cp.async.bulk into a
                                shared-memory ring) instead of per-thread 256-bit loads; same results. -1 (default):
                                where that measured faster (Bit x Prob on grids of >= 4 M voxels),
                                0: never, 1..5: always, the value picks the tile x stages shape (1 = 512 voxels x 4) */
  GVX_OPT_PDL = 4,           /* 1 (default): the per-step kernels (log-odds frame update, summary collide) are started
                                with programmatic dependent launch, so the launch ramp of one overlaps the tail of the
                                kernel before it; they wait for that kernel's completion before touching memory, so
                                results and ordering are those of a plain stream. 0: plain launches */
  GVX_OPT_USE_COARSE = 5     /* 1 (default): a map that is collided twice without being modified in between (a robot
                                volume, a static environment) gets a coarse mask, 1 bit per 4096 voxels, built from its
                                summary on the stream; summary collides then skip chunks that are empty in either map
                                without reading either summary. Any modifying call invalidates the mask. Same results */
};
int gvx_set_option(int option, int value);

/* ---- voxelisation ------------------------------------------------------------------------------------- */
/* insertPointCloud (GV/voxelmap/TemplateVoxelMap.hpp:563-607, kernelInsertGlobalPointCloud
 * GV/voxelmap/kernels/VoxelMapOperations.hpp:241-264). xyz = n packed Vector3f. ProbVoxelMap accepts only
 * eBVM_OCCUPIED (GV/voxelmap/ProbVoxelMap.hpp:118-140) -> GVX_ERR_UNSUPPORTED otherwise, nothing inserted.
 * any_outside (nullable): set to 1 if a point fell outside the whole grid (the reference's warning flag);
 * NULL makes the call asynchronous. */
int gvx_insert_points(gvx_handle_t h, const float* xyz, uint64_t n, int mem_kind, int meaning, int* any_outside);
/* insertMetaPointCloud, both overloads (GV/voxelmap/TemplateVoxelMap.hpp:609-664, kernels
 * GV/voxelmap/kernels/VoxelMapOperations.hpp:302-340,386-437), with the per-sub-cloud transform of
 * MetaPointCloud::transformSubCloud (GV/helpers/MetaPointCloud.cu:624-649) optionally fused in.
 * xyz: the concatenated clouds; cloud_sizes[n_clouds] (host); meanings[n_meanings] (host), n_meanings == 1 or
 * n_clouds; transforms: NULL or n_clouds row-major 4x4 fp32 matrices (host). */
int gvx_insert_meta(gvx_handle_t h, const float* xyz, int mem_kind, const uint32_t* cloud_sizes, uint32_t n_clouds,
                    const int32_t* meanings, uint32_t n_meanings, const float* transforms, int* any_outside);
/* PointCloud::transform / MetaPointCloud::transform (kernelTransformCloud,
 * GV/helpers/kernels/MetaPointCloudOperations.cu:77-91); in and out are both of mem_kind; in == out allowed */
int gvx_transform_points(const float* in_xyz, float* out_xyz, uint64_t n, int mem_kind, const float M16[16]);
/* ProbVoxelMap::insertSensorData without ray casting / robot cut: the log-odds update
 * (GV/voxelmap/ProbVoxelMap.hpp:52-102, kernelTransformSensorData GV/voxelmap/kernels/VoxelMapOperations.cu:27-71,
 * kernelInsertSensorData GV/voxelmap/kernels/VoxelMapOperations.hpp:559-613). pose12 = row-major [R|t];
 * delta_occ = cSENSOR_MODEL_OCCUPIED (72). Atomic and saturating, so no update is lost. Asynchronous. */
int gvx_update_logodds(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind, const float pose12[12],
                       int delta_occ);

/* ProbVoxelMap::insertSensorData complete (GV/voxelmap/ProbVoxelMap.hpp:52-102): enable_raycasting != 0 adds the
 * free-space update of RayCaster::rayCast (GV/voxelmap/kernels/VoxelMapOperations.h:199-323): every cell on the integer
 * line from the sensor's voxel to the hit voxel (both included) receives delta_free (cSENSOR_MODEL_FREE = -10) before
 * the hit voxel receives delta_occ. robot_map != 0 is cut_real_robot (VoxelMapOperations.hpp:578-592): points whose
 * hit voxel has bit robot_bit set in that (whole) BitVectorVoxelMap are dropped. Order of updates: all decrements of
 * the frame, then all hits -- deterministic, no lost updates; equal to the reference executed point by point wherever
 * a voxel is not both crossed and hit by different rays (the reference's own result depends on its thread schedule).
 * On a z-slab map every rank walks the rays and keeps its slab's cells. Asynchronous. */
int gvx_insert_sensor_data(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind, const float pose12[12],
                           int enable_raycasting, gvx_handle_t robot_map, int robot_bit, int delta_free,
                           int delta_occ);

/* The same update for several sensor frames, each with its own pose, in ONE launch (multi-camera rigs; z-slab ranks
 * that fold the whole rig's cloud and keep only their slab's points). xyz_sensor: the concatenated frames;
 * frame_sizes[n_frames] (host); poses12: n_frames row-major [R|t] (host). Asynchronous. */
int gvx_update_logodds_multi(gvx_handle_t h, const float* xyz_sensor, int mem_kind, const uint32_t* frame_sizes,
                             uint32_t n_frames, const float* poses12, int delta_occ);

/* ---- collision counting, map x map -------------------------------------------------------------------- */
/* collideWith (BitVoxelMap::collideWith GV/voxelmap/BitVoxelMap.hpp:191-203, ProbVoxelMap::collideWith
 * GV/voxelmap/ProbVoxelMap.hpp:144-155 -> collisionCheckWithCounterRelativeTransform
 * GV/voxelmap/TemplateVoxelMap.hpp:486-519, kernelCollideVoxelMapsDebug
 * GV/voxelmap/kernels/VoxelMapOperations.hpp:129-184). a is `this`, b is `other`; off3 (nullable) is the
 * Vector3i offset applied to a's base pointer; voxels where the shifted index leaves a's allocation are
 * skipped (the reference reads out of bounds there). mark != 0 reproduces the reference's side effect
 * a[i].insert(eBVM_COLLISION) on every colliding voxel of a. */
int gvx_collide_map_map(gvx_handle_t a, gvx_handle_t b, float thr, const int32_t off3[3], int mark, uint64_t* count);
/* same, result (one uint64) written to device memory, no synchronisation */
int gvx_collide_map_map_async(gvx_handle_t a, gvx_handle_t b, float thr, const int32_t off3[3], int mark,
                              uint64_t* count_d);
/* collisionCheck (boolean, no side effect; GV/voxelmap/TemplateVoxelMap.hpp:329-414) */
/* collisionCheckWithCounter[RelativeTransform](other, DefaultCollider(Probability threshold1, Probability threshold2))
 * (GV/voxelmap/TemplateVoxelMap.hpp:477-519, GV/voxel/DefaultCollider.hpp:45-49): the collider built from two int8
 * thresholds -- threshold1 applies to a ProbabilisticVoxel on side a, threshold2 to one on side b; BitVectorVoxels ignore
 * them. Otherwise gvx_collide_map_map. */
int gvx_collide_map_map_thresholds(gvx_handle_t a, gvx_handle_t b, int threshold1, int threshold2,
                                   const int32_t off3[3], int mark, uint64_t* count);
int gvx_collide_map_map_any(gvx_handle_t a, gvx_handle_t b, float thr, int* flag);
/* BitVoxelMap::collideWithTypes(ProbVoxelMap*) (GV/voxelmap/BitVoxelMap.hpp:212-217 -> collisionCheckBitvector :139-188,
 * kernelCollideVoxelMapsBitvector GV/voxelmap/kernels/VoxelMapOperations.hpp:187-239, SVCollider::collide
 * GV/voxel/SVCollider.hpp:96-106). mask_inout32 is OR-ed into (the reference does not reset it). */
int gvx_collide_map_map_types(gvx_handle_t a, gvx_handle_t b, float thr, int mark,
                              uint8_t mask_inout32[GVX_BITVOXEL_BYTES], uint64_t* count);
/* Swept-volume check with a bit window: BitVoxelMap::collisionCheckBitvector(BitVoxelMap* other,
 * SVCollider(thr, margin), types, sv_offset) (GV/voxelmap/BitVoxelMap.hpp:91-135; SVCollider::collide(Bit, Bit, ...)
 * GV/voxel/SVCollider.hpp:87-93; bitMarginCollisionCheck GV/helpers/BitVector.h:416-471; reference tests
 * GV/test/testing_voxelmap.cu:173-229). A voxel collides if a swept-volume bit of b lies within +-margin bits of a
 * bit of a (after moving a by sv_offset); *count = such voxels, types_inout32 |= their colliding bits of a, and with
 * mark != 0 they receive eBVM_COLLISION. BitVoxelMap::collideWithTypes(BitVectorVoxelMap*) (BitVoxelMap.hpp:205-210)
 * is margin 0, sv_offset 0. margin <= 31 and sv_offset / 8 < 32 + margin (beyond that the reference shifts a
 * 64-bit word by >= 64: undefined there, refused here). */
int gvx_collide_map_map_bitcheck(gvx_handle_t a, gvx_handle_t b, uint32_t margin, uint32_t sv_offset, int mark,
                                 uint8_t types_inout32[GVX_BITVOXEL_BYTES], uint64_t* count);
/* BitVoxelMap::shiftLeftSweptVolumeIDs (GV/voxelmap/BitVoxelMap.hpp:234-246; performLeftShift
 * GV/helpers/BitVector.h:361-402; reference test GV/test/testing_voxelmap.cu:238-304): every voxel's bit vector moves
 * `shift` bits towards bit 0 and loses its four non-swept-volume bits. shift > 56 is an error (no change), as there. */
int gvx_map_shift_sv_ids(gvx_handle_t h, uint32_t shift);

/* ---- multi-GPU: collision count all-reduced INSIDE the collide kernel over NVLink peer memory ------------- */
/* One process per GPU (not in the reference). Every rank creates a small mailbox in its HBM and exports it as a
 * 64-byte CUDA IPC handle; the host program exchanges the handles (any transport: torch.distributed, MPI, files)
 * and every rank maps all peers' mailboxes. From then on the *_allreduce collide variants finish with the last CTA
 * storing this rank's count into every peer's mailbox and summing all ranks' counts: compute + collective are ONE
 * kernel, no NCCL launch follows. All ranks must issue the same sequence of *_allreduce calls. */
int gvx_peer_mailbox_create(void* ipc_handle64_out);
int gvx_peer_mailbox_connect(int rank, int world, const void* all_handles /* world x 64 bytes, rank order */);
int gvx_peer_mailbox_disconnect(void);
/* collideWith on this rank's z-slab (offset 0), result = sum over all ranks; count (host, nullable: NULL = no sync)
 * and count_d (device, nullable) both receive the GLOBAL count */
int gvx_collide_map_map_allreduce(gvx_handle_t a, gvx_handle_t b, float thr, int mark, uint64_t* count,
                                  uint64_t* count_d);
int gvx_collide_list_map_allreduce(gvx_handle_t list, gvx_handle_t map, float thr, int mark, uint64_t* count,
                                   uint64_t* count_d);
/* per-meaning counts of the list against this rank's z-slab, the total AND the 256 counts all-reduced inside the kernel
 * over the peers' mailboxes (257 packed words per peer; no NCCL launch): counts257_d (device) receives the GLOBAL total in
 * [0] and the global per-meaning counts in [1..256] on every rank; nothing synchronises. A collective. */
int gvx_collide_list_map_per_meaning_allreduce(gvx_handle_t list, gvx_handle_t map, float thr, int mark,
                                               uint64_t* counts257_d);

/* ---- multi-GPU behind the drop-in API (SPMD, one process per GPU, no torch / MPI needed) ---------------------- */
/* Start N copies of the SAME program with GVX_SHARD=z GVX_WORLD=N GVX_RANK=r in the environment ([GVX_LOCAL_RANK],
 * [GVX_DEVICES=0,1,..], [GVX_RENDEZVOUS=name]; torchrun's RANK / WORLD_SIZE / LOCAL_RANK / MASTER_PORT and mpirun's
 * OMPI_COMM_WORLD_* are understood). gvx_init(-1) then selects the rank's device and joins the job's POSIX
 * shared-memory rendezvous; gvx_map_create (GpuVoxels::addMap, GV/GpuVoxels.cpp:164-290) allocates this rank's z-slab
 * of the grid -- so maps beyond one GPU's memory and the reference's 4 GiB limit (GV/voxelmap/TemplateVoxelMap.hpp:
 * 65-75) fit; inserts keep the slab's points; voxel lists stay replicated. Every synchronous call that returns a
 * collision result returns the GLOBAL result on every rank: gvx_collide_map_map's count is all-reduced inside the
 * collide kernel over NVLink peer memory (GVX_SHARD_FUSED=0: through the host segment), gvx_collide_map_map_any / _types /
 * _bitcheck and gvx_collide_list_map / _types / _mask / _per_meaning through the host segment (sum of counts, OR of
 * masks). Every rank must make the same calls in the same order. The *_async variants and downloads stay local to
 * the slab.
 * collideWith with an OFFSET (gvx_collide_map_map, gvx_collide_map_map_thresholds): the offset shifts the linear voxel
 * index (GV/voxelmap/TemplateVoxelMap.hpp:486-519, getVoxelIndexSigned), so the partners of a slab's voxels lie in
 * other ranks' slabs. On the first such call every rank maps the peers' slabs of the map through CUDA IPC (collective;
 * kept until the map is destroyed, which then contains a barrier); the collide then runs one launch per peer slab that
 * holds partners, reading them over NVLink, and returns the global count. Needs peer access between the ranks' GPUs
 * (or ranks sharing a GPU); the *_async / *_allreduce variants, lanes and explicit gvx_map_create_slab maps still refuse
 * offsets (GVX_ERR_UNSUPPORTED). */
int gvx_shard_info(int* rank, int* world, int* fused);
/* utility for callers with results of their own: sum of values[0..n_sum), OR of the next n_or words, over all ranks */
int gvx_shard_allreduce_u64(uint64_t* values, int n_sum, int n_or);
/* Offset collides on EXPLICIT z-slab maps (gvx_map_create_slab), for host programs with a communicator of their own:
 * every rank exports its slab of the map that will be the `other` side (gvx_map_ipc_handle: a CUDA IPC handle, 64
 * bytes), the program all-gathers the handles and every rank connects (all_handles: world x 64 bytes in rank order;
 * z_bounds: world + 1 slab boundaries, rank r owns layers [z_bounds[r], z_bounds[r + 1])). From then on
 * gvx_collide_map_map / gvx_collide_map_map_thresholds with an offset read the partner voxels that lie in other ranks'
 * slabs from the owners' GPUs and return the LOCAL count (this rank's voxels of `a`): the program sums it over the
 * ranks, and orders the ranks around the call (every rank's earlier inserts complete before any rank collides; every
 * rank's collide complete before any rank modifies the map again). Disconnect on every rank, synchronise, then destroy. */
int gvx_map_ipc_handle(gvx_handle_t map, void* handle64_out);
int gvx_map_connect_peers(gvx_handle_t map, int rank, int world, const void* all_handles, const uint32_t* z_bounds);
int gvx_map_disconnect_peers(gvx_handle_t map);
/* The launches such a collide is made of, as pure arithmetic (no GPU needed; used by the library itself and pinned by
 * brute force in the CPU tests): for the rank that owns layers [z_bounds[rank], z_bounds[rank + 1]) of `a` and every rank s,
 * seg3[3 s .. 3 s + 2] = {i_begin, i_end, off_local}: over the LOCAL indices [i_begin, i_end) of rank s's slab of `b`,
 * b_s[i'] meets a_local[i' + off_local]; i_begin == i_end: rank s holds no partner. */
int gvx_halo_plan(uint32_t dx, uint32_t dy, uint32_t dz, const uint32_t* z_bounds, int world, int rank,
                  const int32_t off3[3], int64_t* seg3);
/* every rank's `bytes` bytes (at most 2176), in rank order, into `all` (world x bytes) on every rank */
int gvx_shard_allgather(const void* mine, uint32_t bytes, void* all);
/* join / leave the job's rendezvous explicitly instead of through the environment: host side only (no GPU needed, no
 * peer mailboxes: every global result then goes through the host segment). Call before gvx_init, or in a process that
 * only needs gvx_shard_allreduce_u64 (the CPU tests do). */
int gvx_shard_connect_host(int rank, int world, const char* rendezvous);
int gvx_shard_disconnect(void);

/* ---- multi-GPU: sensor frames routed to the rank that owns the voxel, INSIDE the update kernel ---------------- */
/* insertSensorData (log-odds hits, ProbVoxelMap::insertSensorData GV/voxelmap/ProbVoxelMap.hpp:52-102 without ray
 * casting) on a map cut into z-slabs, one sensor per rank: every rank transforms and voxelises only ITS frame; hits
 * inside its slab are applied at once, a hit in another rank's slab is stored as a 32-bit slab-local voxel index into
 * an inbox in that rank's HBM (CUDA IPC mapping, NVLink), and the same kernel then applies what the other ranks sent.
 * The all-to-all is part of the kernel: no all-gather of the frames, no second launch, work per rank independent of
 * the number of ranks. The slab ends up byte-identical to that slice of the map the reference builds from all frames
 * (equal-sign saturating updates commute).
 * create: this rank's inboxes for `world` sources of up to capacity_points hits each (double-buffered), exported as
 * a 64-byte CUDA IPC handle; connect: map all peers (handles in rank order); collective calls from then on. */
int gvx_peer_dispatch_create(int world, uint64_t capacity_points, void* ipc_handle64_out);
int gvx_peer_dispatch_connect(int rank, int world, const void* all_handles /* world x 64 bytes, rank order */);
int gvx_peer_dispatch_disconnect(void);
/* COLLECTIVE: every connected rank calls it, in the same order, also with n == 0. xyz_sensor: this rank's frame
 * (n <= capacity_points) in the sensor frame, pose12 its row-major [R|t]; z_bounds (host, world + 1 entries): rank r
 * owns global z in [z_bounds[r], z_bounds[r + 1]). Not connected: the plain gvx_update_logodds.
 * Asynchronous; a rank that does not arrive within ~4 s makes the NEXT dispatch call return GVX_ERR_CUDA. */
int gvx_update_logodds_dispatch(gvx_handle_t h, const float* xyz_sensor, uint64_t n, int mem_kind,
                                const float pose12[12], int delta_occ, const uint32_t* z_bounds);

/* ---- voxel lists (BitVectorVoxelList) ----------------------------------------------------------------- */
/* GpuVoxels::addMap(MT_BITVECTOR_VOXELLIST) -> TemplateVoxelList ctor (GV/GpuVoxels.cpp:213-221,
 * GV/voxellist/TemplateVoxelList.hpp:67-86). Storage is the reference's SoA: id[u32] = linear voxel index,
 * coord[3 x u32], voxel[32 B] (GV/voxellist/TemplateVoxelList.h:250-252); always sorted by id and unique. */
int gvx_list_create(gvx_handle_t* h, uint32_t dx, uint32_t dy, uint32_t dz, float side);
int gvx_list_destroy(gvx_handle_t h);
/* TemplateVoxelList::clearMap (GV/voxellist/TemplateVoxelList.hpp:405-416) */
int gvx_list_clear(gvx_handle_t h);
/* getDimensions().x of the reference's list = number of entries (GV/voxellist/TemplateVoxelList.hpp:418-422) */
int gvx_list_size(gvx_handle_t h, uint64_t* size);
/* TemplateVoxelList::insertPointCloud (GV/voxellist/TemplateVoxelList.hpp:284-341, kernelInsertGlobalPointCloud
 * GV/voxellist/kernels/VoxelListOperations.hpp:41-61) followed by make_unique (:141-208): one entry per point with NO
 * bounds check (outside points get wrapped uint32 ids), stable sort by id, OR-merge of equal ids, unique. */
int gvx_list_insert_points(gvx_handle_t h, const float* xyz, uint64_t n, int mem_kind, int meaning);
/* TemplateVoxelList<BitVoxel>::merge (GV/voxellist/TemplateVoxelList.hpp:548-608; applyOffsetOperator :68-90): the
 * entries of `other` are appended to `list` -- ids shifted by getVoxelIndexSigned(voxel_offset), coordinates by
 * voxel_offset (both in uint32 arithmetic, no bounds check, as there), voxels replaced by the single bit new_meaning
 * unless it is -1 -- and the list is made unique again (equal ids OR-merged, the list's own entry keeps its
 * coordinates). voxel_offset nullable = 0. The reference does not compare the two lists' reference dimensions. */
int gvx_list_merge(gvx_handle_t list, gvx_handle_t other, const int32_t voxel_offset[3], int new_meaning);
/* TemplateVoxelList::insertMetaPointCloud x2 (GV/voxellist/TemplateVoxelList.hpp:343-403, kernels
 * GV/voxellist/kernels/VoxelListOperations.hpp:63-120); n_meanings == 1 or n_clouds */
int gvx_list_insert_meta(gvx_handle_t h, const float* xyz, int mem_kind, const uint32_t* cloud_sizes, uint32_t n_clouds,
                         const int32_t* meanings, uint32_t n_meanings);
/* copyCoordsToHost / the payload of writeToDisk (GV/voxellist/TemplateVoxelList.hpp:227-282, 444-471); each output
 * pointer may be NULL; buffers hold gvx_list_size entries */
int gvx_list_download(gvx_handle_t h, uint32_t* ids, uint32_t* coords, uint8_t* voxels);
/* readFromDisk's payload: replace the list by n host entries; ids must be strictly increasing; coords may be NULL */
int gvx_list_upload(gvx_handle_t h, const uint32_t* ids, const uint32_t* coords, const uint8_t* voxels, uint64_t n);
/* getDeviceIdPtr / getDeviceCoordPtr / getDeviceDataPtr (GV/voxellist/TemplateVoxelList.h:120-160) */
int gvx_list_device_ptrs(gvx_handle_t h, void** ids_d, void** coords_d, void** voxels_d, uint64_t* size);

/* ---- collision counting, list x map ------------------------------------------------------------------- */
/* BitVoxelList::collideWith(ProbVoxelMap* | BitVectorVoxelMap*) (GV/voxellist/BitVoxelList.hpp:70-84 ->
 * TemplateVoxelList::collisionCheckWithCollider GV/voxellist/TemplateVoxelList.hpp:731-774, kernelCollideWithVoxelMap
 * GV/voxellist/kernels/VoxelListOperations.hpp:126-177). The map voxel of entry e is map[id_e + linear(off3)], used
 * only if it lies inside the map. mark != 0: colliding LIST voxels get eBVM_COLLISION. The map may be a z-slab
 * (entries of other slabs are skipped; sum the counts over ranks). Different reference dimensions ->
 * GVX_ERR_DIM_MISMATCH (the reference returns SSIZE_MAX, TemplateVoxelList.hpp:737-742). */
int gvx_collide_list_map(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3], int mark,
                         uint64_t* count);
int gvx_collide_list_map_async(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3], int mark,
                               uint64_t* count_d);
/* gvx_collide_list_map_per_meaning (below) without the host round trip: counts257_d (device) receives the total in
 * [0] and the 256 per-meaning counts in [1..256]; nothing synchronises. On z-slab maps every rank gets its slab's
 * share -- sum the 257 words over the ranks (ncclAllReduce on the same stream: the path's one collective for
 * BASELINE config 5; bit j of the meaning mask = count j > 0). */
int gvx_collide_list_map_per_meaning_async(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3],
                                           int mark, uint64_t* counts257_d);
/* BitVoxelList::collideWithTypes (GV/voxellist/BitVoxelList.hpp:127-218, kernel VoxelListOperations.hpp:182-240):
 * count + OR of the colliding list voxels; mask_out32 is cleared first (BitVoxelList.hpp:163); the map voxel is
 * tested with isOccupied(thr) on the RAW float threshold (GV/voxel/ProbabilisticVoxel.hpp:94-98) */
int gvx_collide_list_map_types(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3], int mark,
                               uint8_t mask_out32[GVX_BITVOXEL_BYTES], uint64_t* count);
/* BitVoxelList::collideWithTypeMask (GV/voxellist/BitVoxelList.hpp:220-265, kernel VoxelListOperations.hpp:260-314):
 * only entries with (types_to_check & list voxel) != 0 count -- the BitVoxelMeaning collision mask */
int gvx_collide_list_map_mask(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3], int mark,
                              const uint8_t mask_in32[GVX_BITVOXEL_BYTES], uint64_t* count);
/* Per-meaning counts with the semantics of BitVoxelList::collideCountingPerMeaning (GV/voxellist/BitVoxelList.hpp:300-344,
 * list x list only in the reference, summed on the host) applied to collideWith's colliding set: every set bit j of a
 * colliding list voxel adds 1 to counts256[j] and to *total. */
int gvx_collide_list_map_per_meaning(gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3], int mark,
                                     uint64_t counts256[256], uint64_t* total);

/* ---- list x list (offsets between lists are refused by the reference, BitVoxelList.hpp:413-417; none here) ---- */
/* BitVoxelList::collideWith(BitVectorVoxelList*) (GV/voxellist/BitVoxelList.hpp:86-99 -> collideVoxellists,
 * GV/voxellist/TemplateVoxelList.hpp:227-282): the number of entries of `a` whose voxel id is also in `b` */
int gvx_collide_list_list(gvx_handle_t a, gvx_handle_t b, uint64_t* count);
/* collideWithTypes(BitVectorVoxelList*) (BitVoxelList.hpp:101-125): types_out32 = OR over the matching pairs of
 * (a's voxel | b's voxel) -- assigned, not accumulated -- and the number of matching entries */
int gvx_collide_list_list_types(gvx_handle_t a, gvx_handle_t b, uint8_t types_out32[GVX_BITVOXEL_BYTES], uint64_t* count);
/* collideWithBitcheck(BitVectorVoxelList*, margin) (BitVoxelList.hpp:267-296): matching pairs with a common bit
 * (margin 0: BitvectorCollision, ALL bits) or passing bitMarginCollisionCheck(a, b, margin, 0) (margin 1..31);
 * reference test GV/test/testing_voxellist.cu:412-460 */
int gvx_collide_list_list_bitcheck(gvx_handle_t a, gvx_handle_t b, uint32_t margin, uint64_t* count);
/* collideCountingPerMeaning(list) (BitVoxelList.hpp:299-344; tests testing_voxellist.cu:44-107,341-410): every set bit j
 * of a matching entry OF `a` adds one to counts256[j] and to *total */
int gvx_collide_list_list_per_meaning(gvx_handle_t a, gvx_handle_t b, uint64_t counts256[256], uint64_t* total);
/* TemplateVoxelList::subtract (GV/voxellist/TemplateVoxelList.hpp:609-641; test testing_voxellist.cu:541-572): removes
 * from `a` every entry whose voxel id is in `b` */
int gvx_list_subtract(gvx_handle_t a, gvx_handle_t b);
/* TemplateVoxelList::equals (TemplateVoxelList.hpp:806-825): same dimensions, side, size, ids, coordinates and voxels */
int gvx_list_equals(gvx_handle_t a, gvx_handle_t b, int* equal);
/* BitVoxelList::shiftLeftSweptVolumeIDs (GV/voxellist/BitVoxelList.hpp:560-582) */
int gvx_list_shift_sv_ids(gvx_handle_t list, uint32_t shift);

/* ---- files: the reference's binary formats, readable and writable both ways ------------------------------------ */
/* TemplateVoxelMap::writeToDisk / readFromDisk (GV/voxelmap/TemplateVoxelMap.hpp:666-799): int32 MapType, float side,
 * 3 x uint32 dims, raw voxels. Reading checks type and dimensions (a different side length is only a warning there
 * and is accepted here). Whole maps only. */
int gvx_map_write_file(gvx_handle_t h, const char* path);
int gvx_map_read_file(gvx_handle_t h, const char* path);
/* TemplateVoxelList::writeToDisk / readFromDisk (GV/voxellist/TemplateVoxelList.hpp:444-526; test
 * testing_voxellist.cu:503-539): int32 MapType, 3 x uint32 dims, float side, uint32 n, ids, coords, voxels */
int gvx_list_write_file(gvx_handle_t h, const char* path);
int gvx_list_read_file(gvx_handle_t h, const char* path);

/* ---- point-cloud files (.xyz, .pcd, .binvox) ------------------------------------------------------------------- */
/* file_handling::PointcloudFileHandler::loadPointCloud (GV/helpers/PointcloudFileHandler.cpp:64-137) with its readers
 * (GV/helpers/XyzFileReader.cpp:28-59, BinvoxFileReader.cpp:38-166, PcdFileReader.cpp:31-54): the format is chosen by the
 * first of "xyz", "pcd", "binvox" found anywhere in the path; use_model_path prepends $GPU_VOXELS_MODEL_PATH
 * (GV/helpers/common_defines.h:276-293); shift_to_zero moves the minimum corner to the origin; then every coordinate
 * becomes scaling * c + offset (offset_xyz nullable = 0). Host only: works before gvx_init and without a GPU.
 * *xyz_out receives n packed Vector3f allocated by the library (NULL for an empty cloud): release with
 * gvx_pointcloud_file_free. GVX_ERR_IO: cannot be opened / truncated; GVX_ERR_INVALID: unknown or malformed format
 * (the reference returns false for both). */
int gvx_pointcloud_file_load(const char* path, int use_model_path, int shift_to_zero, const float* offset_xyz,
                             float scaling, float** xyz_out, uint64_t* n_out);
void gvx_pointcloud_file_free(float* xyz);
/* GpuVoxelsMap::insertPointCloudFromFile (GV/GpuVoxelsMap.cpp:36-49; GpuVoxels::insertPointCloudFromFile,
 * GV/GpuVoxels.cpp:427-440): load as above, then insertPointCloud into a map / a list */
int gvx_insert_pointcloud_file(gvx_handle_t map, const char* path, int use_model_path, int meaning, int shift_to_zero,
                               const float* offset_xyz, float scaling, int* any_outside);
int gvx_list_insert_pointcloud_file(gvx_handle_t list, const char* path, int use_model_path, int meaning,
                                    int shift_to_zero, const float* offset_xyz, float scaling);

/* ---- batch of robot configurations -------------------------------------------------------------------- */
/* n_cfg times the planner's state check (gvl_ompl_planning/gvl_ompl_planner_helper.cpp:169-202): clearMap(robot);
 * transformSubCloud per link (GV/helpers/MetaPointCloud.cu:624-649); insertMetaPointCloud(robot, meanings);
 * counts[c] = robot.collideWith(env, thr) -- in ONE call, without a robot grid and without per-configuration
 * launches or syncs. base_xyz: the concatenated link clouds in link frames; meanings[n_clouds];
 * transforms: n_cfg x n_clouds row-major 4x4 fp32 (host). mask32 (nullable) is the BitVoxelMeaning collision mask:
 * a robot voxel counts only if (mask & voxel) != 0; NULL = DefaultCollider (any meaning other than eBVM_FREE).
 * No side effects on env. counts: n_cfg host uint64. */
int gvx_batch_robot_collide(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes,
                            uint32_t n_clouds, const int32_t* meanings, const float* transforms, uint32_t n_cfg,
                            const uint8_t* mask32, float thr, uint64_t* counts);
/* The same check with the forward kinematics of a DH chain done on the device: replaces, per configuration,
 * KinematicChain::setConfiguration (GV/robot/dh_robot/KinematicChain.cu:93-126: DHParameters::convertDHtoM,
 * GV/robot/dh_robot/KinematicLink.cu:30-84, and the host Matrix4f product, GV/helpers/cuda_matrices.h:772-803).
 * dh4: n_links x {d, theta, a, alpha}; joint_types[n_links]: 0 REVOLUTE, 1 PRISMATIC; joint_values: n_cfg x n_links
 * (host). link_cloud[n_links] (nullable): index of the sub-cloud that moves with link l, or -1 (the reference matches
 * them by name; a link without a cloud still enters the chain, a cloud without a link is not transformed);
 * NULL = link l moves cloud l. The cloud of link l is transformed by T_0 * ... * T_(l-1), as in the reference. */
int gvx_batch_robot_collide_dh(gvx_handle_t env, const float* base_xyz, int mem_kind, const uint32_t* cloud_sizes,
                               uint32_t n_clouds, const int32_t* meanings, const float* dh4,
                               const int32_t* joint_types, uint32_t n_links, const int32_t* link_cloud,
                               const float* joint_values, uint32_t n_cfg, const uint8_t* mask32, float thr,
                               uint64_t* counts);
/* The per-link matrices alone: transforms_out (host) n_cfg x n_links row-major 4x4 fp32, bit-identical to the
 * reference's host arithmetic (double libm sin/cos rounded to float, float products without FMA): the device result is
 * used where its rounding is provably the same, the few other configurations (*n_ambiguous, nullable; ~1e-8 per
 * value) are recomputed on the host. */
int gvx_dh_chain_transforms(const float* dh4, const int32_t* joint_types, uint32_t n_links, const float* joint_values,
                            uint32_t n_cfg, float* transforms_out, uint32_t* n_ambiguous);
/* bytes of per-wave scratch the batch call may use for its per-configuration "seen" bitmaps (default 4 GiB) */
int gvx_batch_set_scratch_budget(uint64_t bytes);
/* No counterpart in the reference. gvx_batch_robot_collide for callers
 * that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc): ALL
 * scratch -- the environment's 1-bit-per-voxel bitmap and one "seen" bitmap per configuration -- lives in one static
 * __device__ array of gvx_noalloc_scratch_bytes() = 16 MiB inside the library (every GPU that loads libgvx.so holds it
 * until the process exits). Inputs and output are the caller's: xyz_d (device, packed Vector3f of all link clouds),
 * cloud_sizes / meanings (host, n_clouds <= 16), transforms_d (device, n_cfg x n_clouds x 16 floats, row-major
 * Matrix4f), mask32 (host, nullable), counts_d (device, n_cfg x uint64, written by the call). Same result as
 * gvx_batch_robot_collide. Limits, refused with GVX_ERR_INVALID and a message: (1 + n_cfg) bitmaps of
 * ceil(voxels / 2048) * 256 bytes must fit the 16 MiB (e.g. 63 configurations of a 128^3 grid, 7 of 256^3),
 * n_clouds <= 16, n_cfg <= 65535. The call enqueues a memset, at most one bitmap kernel and one collide kernel on the
 * library's stream and returns without synchronising; it calls no cudaMalloc / cudaFree / cudaHostAlloc and no
 * allocating library routine. All calls share the array, which keeps the last call's contents (this call clears what
 * must start at zero): calls that overlap on the GPU race -- keep them ordered on one stream. */
int gvx_batch_robot_collide_noalloc(gvx_handle_t env_map, const float* xyz_d, const uint32_t* cloud_sizes,
                                    uint32_t n_clouds, const int32_t* meanings, const float* transforms_d,
                                    uint32_t n_cfg, const uint8_t* mask32, float thr, uint64_t* counts_d);
uint64_t gvx_noalloc_scratch_bytes(void);

/* ------------------------------------------------------------------ collision lanes: concurrent read-only queries
 * The reference serialises per MAP: collideWith / collisionCheck lock the two maps' m_mutex
 * (GV/voxelmap/TemplateVoxelMap.hpp:1043-1075 collisionCheckWithCounterRelativeTransform and the callers above it;
 * GV/voxellist/TemplateVoxelList.hpp:713-760), so planner threads checking different robot maps against one environment
 * run at the same time. The serial entry points above share one stream and one result block under one mutex. A lane is
 * a stream + accumulators + result blocks of its own: the library mutex is held only while the query is validated and
 * enqueued, the wait for the result runs outside it, and queries issued on different lanes (one lane per thread)
 * overlap on the GPU. One lane serialises its own callers.
 * Ordering against the serial API is by events: a lane query sees every insert / clear / collide-with-mark the process
 * enqueued before the call, and any later serial call is ordered behind the lane queries still in flight. Lane queries
 * never write voxels (no `mark`) and allocate nothing after gvx_lane_create. Under GVX_SHARD=z they are refused
 * (GVX_ERR_UNSUPPORTED): a global result needs a collective, which every rank must issue in the same order.
 * Results equal gvx_collide_map_map(..., mark = 0) / gvx_collide_map_map_any / gvx_collide_list_map(..., mark = 0). */
int gvx_lane_create(gvx_handle_t* lane);
int gvx_lane_destroy(gvx_handle_t lane); /* waits for the lane's last query */
int gvx_lane_count(uint32_t* n);
int gvx_lane_collide_map_map(gvx_handle_t lane, gvx_handle_t a, gvx_handle_t b, float thr, const int32_t off3[3],
                             uint64_t* count);
int gvx_lane_collide_map_map_any(gvx_handle_t lane, gvx_handle_t a, gvx_handle_t b, float thr, int* flag);
int gvx_lane_collide_list_map(gvx_handle_t lane, gvx_handle_t list, gvx_handle_t map, float thr, const int32_t off3[3],
                              uint64_t* count);

#ifdef __cplusplus
}
#endif
#endif /* GVX_H_INCLUDED */
