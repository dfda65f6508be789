"""gpu_voxels_b200 -- B200-native (sm_100a) hot path of gpu-voxels: point-cloud voxelisation and dense
voxel-grid collision counting, behind the reference's own API names.

  libgvx.so   hand-written CUDA + the C ABI of include/gvx.h     (csrc/, built by build.py)
  _lib        ctypes binding of that ABI
  api         host-side mirror of the reference's classes (GpuVoxels, BitVectorVoxelMap, ProbVoxelMap, ...)

There is no CPU fallback: every call ends in libgvx.so and fails loudly without it or without a GPU.
"""
from .api import *  # noqa: F401,F403
from .api import GpuVoxels, BitVectorVoxelMap, ProbVoxelMap, MetaPointCloud, BitVectorVoxel  # noqa: F401
from .api import batch_robot_collide, batch_robot_collide_dh, dh_chain_transforms, transform_points  # noqa: F401
from .api import PointcloudFileHandler, CollisionLane  # noqa: F401
from .voxellist import BitVectorVoxelList  # noqa: F401
from ._lib import GvxError  # noqa: F401
