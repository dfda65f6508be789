#!/usr/bin/env python
"""benchmarks/fold_probe.py -- the multi-frame log-odds fold as ONE rank of an N-rank z-slab job sees it, on a single GPU
(no peers needed): grid 512 x 512 x (512 N), this process holds slab 0, the cloud is N sensor frames of which only the
first falls into the slab. Prints the per-launch time for N = 1, 2, 4, 8 (CUDA events around 200 back-to-back launches,
8 rotating maps) -- the part of the sharded step that grows with N."""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench as B  # noqa: E402
import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib  # noqa: E402


def main():
    L = _lib.init(0)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    _lib.check(L.gvx_set_stream(C.c_void_p(stream.cuda_stream)))
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    last = len(sys.argv) > 2 and sys.argv[2] == "last"   # hold the LAST slab instead of the first
    for world in (1, 2, 4, 8):
        dims = (512, 512, 512 * world)
        z0 = 512 * (world - 1) if last else 0
        maps = [gv.ProbVoxelMap(dims, B.SIDE, slab=(z0, z0 + 512)) for _ in range(8)]
        one = torch.cat([torch.from_numpy(B.depth_frame(201 + r)).cuda() for r in range(world)]).contiguous()
        copies = [one.clone() for _ in range(8)]  # 8 x world x 3.7 MB in rotation: the cloud comes from HBM, not L2
        poses = np.ascontiguousarray(np.stack([B.sensor_pose(5.12 * r).reshape(-1) for r in range(world)]))
        sizes = np.full(world, B.N_FRAME, np.uint32)
        sp, pp = sizes.ctypes.data_as(C.c_void_p), poses.ctypes.data_as(C.c_void_p)
        fps = [C.c_void_p(t.data_ptr()) for t in copies]
        hs = [m.handle for m in maps]

        def go(k):
            _lib.check(L.gvx_update_logodds_multi(hs[k % 8], fps[k % 8], _lib.MEM_DEVICE, sp, world, pp, 72))
        for k in range(16):
            go(k)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for k in range(reps):
            go(k)
        e1.record(stream)
        torch.cuda.synchronize()
        print("frames %d: %.2f us per fold launch, known voxels in the slab %d" % (world, e0.elapsed_time(e1) / reps * 1e3,
              int((maps[0].download() != -128).sum())), flush=True)
        for m in maps:
            m.destroy()


if __name__ == "__main__":
    main()
