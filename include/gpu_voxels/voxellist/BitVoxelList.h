// gpu_voxels/voxellist/BitVoxelList.h -- same include path as the reference; the class lives in VoxelList.h.
#include <gpu_voxels/voxellist/VoxelList.h>
