#!/usr/bin/env python
"""benchmarks/profile_kernels.py -- a short driver that launches each kernel family once or twice at a size that is
large against L2 but quick to replay, for `ncu` captures (profiles/):
    list   k_collide_list_map        4 M entries vs a 512^3 BitVectorVoxelMap (4 GiB)
    batch  k_env_bitmap + k_batch_collide   128 configurations x 350k points vs a 256^3 environment
    insert k_insert / k_update_logodds      1 M points
"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib  # noqa: E402
from benchmarks import run_configs as rc  # noqa: E402


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "list,batch,insert"
    rc.setup()
    L = rc.L
    if "list" in which:
        dims, side, n = (512, 512, 512), 0.01, 512 ** 3
        m = gv.BitVectorVoxelMap(dims, side)
        rc.boxes_into_map(m, dims, side, 40, 40, 503)
        g = torch.Generator(device="cuda"); g.manual_seed(501)
        ids = torch.unique(torch.randint(0, n, (4_100_000,), device="cuda", generator=g, dtype=torch.int64))[:4_000_000]
        lst = gv.BitVectorVoxelList(dims, side)
        pts = rc.voxel_centres_on_device(ids, dims, side)
        _lib.check(L.gvx_list_insert_points(lst.handle, rc.p_(pts), pts.shape[0], _lib.MEM_DEVICE, 7))
        cnt = torch.zeros(1, dtype=torch.int64, device="cuda")
        for _ in range(3):
            _lib.check(L.gvx_collide_list_map_async(lst.handle, m.handle, 1.0, None, 0, rc.p_(cnt)))
        print("list collisions", lst.collideCountingPerMeaning(m, mark=False)[0], len(lst))
    if "batch" in which:
        dims, side = (256, 256, 256), 0.01
        env = gv.BitVectorVoxelMap(dims, side)
        rc.boxes_into_map(env, dims, side, 40, 24, 303)
        links, lengths = rc.arm_links()
        T = rc.arm_transforms(lengths, 128)
        for _ in range(2):
            c = gv.batch_robot_collide(env, gv.MetaPointCloud(links), list(range(4, 11)), T, None, 1.0)
        print("batch collisions", int(c.sum()))
    if "insert" in which:
        dims, side = (512, 512, 512), 0.01
        pm, bm = gv.ProbVoxelMap(dims, side), gv.BitVectorVoxelMap(dims, side)
        g = torch.Generator(device="cuda"); g.manual_seed(9)
        pts = (torch.rand((1_000_000, 3), device="cuda", generator=g) * 5.12).contiguous()
        pose = np.ascontiguousarray(np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32).reshape(-1))
        for _ in range(3):
            rc.insert_device_points(bm, pts, 9)
            _lib.check(L.gvx_update_logodds(pm.handle, rc.p_(pts), pts.shape[0], _lib.MEM_DEVICE, pose.ctypes.data_as(C.c_void_p), 72))
        torch.cuda.synchronize()
        print("insert done")


if __name__ == "__main__":
    main()
