"""experiment: effect of cudaLimitMaxL2FetchGranularity on the gather kernels and on the streaming collide"""
import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import gpu_voxels_b200 as gv
from gpu_voxels_b200 import _lib
from benchmarks import run_configs as rc

rc.setup(); L = rc.L
dims, side, n = (1024, 1024, 512), 0.01, 1024 * 1024 * 512
m = gv.BitVectorVoxelMap(dims, side)
rc.boxes_into_map(m, dims, side, 100, 47, 503)
pm = gv.ProbVoxelMap(dims, side)
g = torch.Generator(device="cuda"); g.manual_seed(501)
ids = torch.unique(torch.randint(0, n, (10_200_000,), device="cuda", generator=g, dtype=torch.int64))[:10_000_000]
lst = gv.BitVectorVoxelList(dims, side)
pts = rc.voxel_centres_on_device(ids, dims, side)
_lib.check(L.gvx_list_insert_points(lst.handle, rc.p_(pts), pts.shape[0], _lib.MEM_DEVICE, 7))
rnd = (torch.rand((4_000_000, 3), device="cuda", generator=g) * torch.tensor([10.24, 10.24, 5.12], device="cuda")).contiguous()
cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
pose = np.ascontiguousarray(np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32).reshape(-1))
prev = C.c_int(0)
_lib.check(L.gvx_set_l2_fetch_granularity(0, C.byref(prev)))
print("default granularity", prev.value)
for gran in (prev.value, 32, 64, 128, 32):
    _lib.check(L.gvx_set_l2_fetch_granularity(gran, None))
    t_list = rc.time_gpu(lambda: _lib.check(L.gvx_collide_list_map_async(lst.handle, m.handle, 1.0, None, 0, rc.p_(cnt))), 20)
    t_col = rc.time_gpu(lambda: _lib.check(L.gvx_collide_map_map_async(pm.handle, m.handle, 0.5, None, 1, rc.p_(cnt))), 10)
    t_ins = rc.time_gpu(lambda: rc.insert_device_points(m, rnd, 9), 10)
    t_lo = rc.time_gpu(lambda: _lib.check(L.gvx_update_logodds(pm.handle, rc.p_(rnd), rnd.shape[0], _lib.MEM_DEVICE, pose.ctypes.data_as(C.c_void_p), 72)), 10)
    print("granularity %3d: list collide 10M %.4f ms | map collide 16.5GiB %.4f ms (%.0f GB/s) | insert 4M random bit %.4f ms | logodds 4M random %.4f ms"
          % (gran, t_list[0], t_col[0], 33 * n / t_col[0] / 1e6, t_ins[0], t_lo[0]))
