// gpu_voxels/voxelmap/BitVoxelMap.h -- same include path as the reference; the classes live in VoxelMap.h.
#include <gpu_voxels/voxelmap/VoxelMap.h>
