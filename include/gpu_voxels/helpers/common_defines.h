// gpu_voxels/helpers/common_defines.h -- enums and limits of the reference's public API, same names and values
// (GV/helpers/common_defines.h:33-74). Part of the header-only C++ facade over include/gvx.h.
#ifndef GPU_VOXELS_B200_HELPERS_COMMON_DEFINES_H
#define GPU_VOXELS_B200_HELPERS_COMMON_DEFINES_H

#include <stdint.h>
#include <string>

namespace gpu_voxels {

const uint32_t BIT_VECTOR_LENGTH = 256;  // GV/helpers/common_defines.h:46

// GV/helpers/common_defines.h:50-58
enum BitVoxelMeaning
{
  eBVM_FREE = 0,
  eBVM_OCCUPIED = 1,
  eBVM_COLLISION = 2,
  eBVM_UNKNOWN = 3,
  eBVM_SWEPT_VOLUME_START = 4,
  eBVM_SWEPT_VOLUME_END = 254,
  eBVM_UNDEFINED = 255
};

// GV/helpers/common_defines.h:59-73; the enum order is the on-disk tag
enum MapType
{
  MT_BITVECTOR_VOXELMAP,
  MT_BITVECTOR_VOXELLIST,
  MT_BITVECTOR_OCTREE,
  MT_BITVECTOR_MORTON_VOXELLIST,
  MT_PROBAB_VOXELMAP,
  MT_PROBAB_VOXELLIST,
  MT_PROBAB_OCTREE,
  MT_PROBAB_MORTON_VOXELLIST,
  MT_COUNTING_VOXELLIST,
  MT_DISTANCE_VOXELMAP
};

static const std::string GPU_VOXELS_MAP_TYPE_NOT_IMPLEMENTED = "THIS TYPE OF DATA STRUCTURE IS NOT YET IMPLEMENTED!";
static const std::string GPU_VOXELS_MAP_OPERATION_NOT_SUPPORTED = "THIS OPERATION IS NOT SUPPORTED BY THE DATA STRUCTURE!";
static const std::string GPU_VOXELS_MAP_OPERATION_NOT_YET_SUPPORTED = "THIS OPERATION IS NOT YET SUPPORTED BY THE DATA STRUCTURE!";
static const std::string GPU_VOXELS_MAP_ONLY_SUPPORTS_BVM_OCCUPIED = "THIS DATA STRUCTURE ONLY SUPPORTS BITVOXEL MEANING eBVM_OCCUPIED!";
static const std::string GPU_VOXELS_MAP_SWAP_FOR_COLLIDE = "TRY TO SWAP BOTH DATA STRUCTURES TO COLLIDE.";
static const std::string GPU_VOXELS_MAP_OFFSET_ON_WRONG_DATA_STRUCTURE = "OFFSET ADDITION ONLY POSSIBLE WHEN COLLIDING WITH VOXELMAP/VOXELLIST";
const uint32_t cMAX_NR_OF_DEVICES = 10;  // GV/helpers/common_defines.h:33 (unused there as well: one device per process)

typedef int8_t Probability;  // GV/voxel/ProbabilisticVoxel.h
static const Probability UNKNOWN_PROBABILITY = Probability(-128);
static const Probability MIN_PROBABILITY = Probability(-127);
static const Probability MAX_PROBABILITY = Probability(127);

// ProbabilisticVoxel sensor model (GV/voxelmap/kernels/VoxelMapOperations.h:38-39)
static const Probability cSENSOR_MODEL_FREE = -10;
static const Probability cSENSOR_MODEL_OCCUPIED = 72;

// GV/helpers/common_defines.h:294-295
const float cMBYTE2BYTE = 1024.0 * 1024.0;
const float cBYTE2MBYTE = 1.0 / cMBYTE2BYTE;

}  // namespace gpu_voxels
#endif
