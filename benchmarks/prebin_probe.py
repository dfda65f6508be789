#!/usr/bin/env python
"""benchmarks/prebin_probe.py -- north_star subsystem (1) sketches "points pre-binned by voxel block so atomics stay mostly in
shared memory". Measured instead of argued, on the one config where random-sector DRAM read-modify-writes bound the
insert: C4's 8 x 1 M uniformly random points into a 2048^3 ProbVoxelMap (one GPU, z-slab of `--dz` layers).

  A  the shipped path: gvx_update_logodds_multi on the points as they arrive (warp-aggregated CAS straight to L2)
  B  the same kernel on points PRE-SORTED by voxel id -- the best access pattern any binning can produce (neighbouring
     lanes hit neighbouring bytes: one DRAM sector per 32 voxels instead of one per point), without the load / write-back
     of a shared-memory tile (a 32^3-voxel tile = 64 KiB of traffic for the ~31 points that fall into it), so B is a LOWER
     bound on the scatter time of a binned insert
  S  the binning itself: keys (voxel id per point) + radix sort of (key, index) pairs + gather of the points
     (torch.sort = CUB radix sort, a library call: only a probe, nothing of this ships)
Binning wins only if S + B < A. Also checks that A and B leave byte-identical maps (equal-sign saturating adds commute).
"""
import argparse
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench as B  # noqa: E402
import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dz", type=int, default=2048)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--gran", type=int, default=0, help="cudaLimitMaxL2FetchGranularity (32/64/128; 0 = leave)")
    ap.add_argument("--z0", type=int, default=0, help="first layer of the slab (z0 >= 2048: no point falls into it -> pure filter time)")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "prebin_probe.json"))
    a = ap.parse_args()
    torch.cuda.set_device(0)
    L = _lib.init(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    _lib.set_stream(stream.cuda_stream)
    prev = C.c_int(0)
    _lib.check(L.gvx_set_l2_fetch_granularity(a.gran, C.byref(prev)))
    dims = (B.C4_DIM,) * 3
    clouds = torch.from_numpy(np.concatenate([B.c4_cloud(k) for k in range(B.C4_CLOUDS)])).cuda()
    n = clouds.shape[0]
    sizes = np.full(B.C4_CLOUDS, B.C4_POINTS, np.uint32)
    poses = np.ascontiguousarray(np.tile(B.C4_POSE.reshape(-1), (B.C4_CLOUDS, 1)))
    if a.z0 >= B.C4_DIM:   # a slab ABOVE the cloud: every point is read, filtered and dropped
        dims = (B.C4_DIM, B.C4_DIM, a.z0 + a.dz)
    m = gv.ProbVoxelMap(dims, B.SIDE, slab=(a.z0, a.z0 + a.dz))

    def fold(pts):
        _lib.check(L.gvx_update_logodds_multi(m.handle, C.c_void_p(pts.data_ptr()), _lib.MEM_DEVICE,
                                              sizes.ctypes.data_as(C.c_void_p), B.C4_CLOUDS, poses.ctypes.data_as(C.c_void_p), 72))

    def bin_points(pts):
        w = pts + torch.tensor(B.C4_POSE[:, 3], device="cuda")
        c = torch.floor(w / B.SIDE).to(torch.int64).clamp_(0, B.C4_DIM - 1)
        key = (c[:, 2] * B.C4_DIM + c[:, 1]) * B.C4_DIM + c[:, 0]
        order = torch.sort(key).indices
        return pts[order].contiguous()

    def timed(fn, iters):
        ms = []
        for _ in range(iters):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream); fn(); e1.record(stream)
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        return {"median": ms[len(ms) // 2], "min": ms[0]}

    sorted_pts = bin_points(clouds)
    res = {"points": n, "slab_layers": a.dz, "l2_fetch_granularity": a.gran or prev.value, "l2_fetch_granularity_default": prev.value}
    for tag, pts in (("A_unsorted", clouds), ("B_presorted", sorted_pts)):
        m.clearMap()
        fold(pts); torch.cuda.synchronize()                      # first touch: every voxel still UNKNOWN
        m.clearMap(); torch.cuda.synchronize()
        res[tag + "_first_frame_ms"] = timed(lambda: fold(pts), 1)["median"]
        res[tag + "_steady_ms"] = timed(lambda: fold(pts), a.iters)   # later frames: the same voxels again
    res["S_binning_ms"] = timed(lambda: bin_points(clouds), a.iters)
    # back to back: `iters` launches between ONE event pair (no per-launch event overhead)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(a.iters):
        fold(clouds)
    e1.record(stream)
    torch.cuda.synchronize()
    res["A_unsorted_back_to_back_ms"] = e0.elapsed_time(e1) / a.iters
    # identical maps
    dims = (B.C4_DIM,) * 3
    ma, mb = gv.ProbVoxelMap(dims, B.SIDE, slab=(0, 64)), gv.ProbVoxelMap(dims, B.SIDE, slab=(0, 64))
    for mm, pts in ((ma, clouds), (mb, sorted_pts)):
        for _ in range(3):
            _lib.check(L.gvx_update_logodds_multi(mm.handle, C.c_void_p(pts.data_ptr()), _lib.MEM_DEVICE,
                                                  sizes.ctypes.data_as(C.c_void_p), B.C4_CLOUDS, poses.ctypes.data_as(C.c_void_p), 72))
    res["maps_identical"] = bool(np.array_equal(ma.download(), mb.download()))
    a_ms, b_ms, s_ms = res["A_unsorted_steady_ms"]["median"], res["B_presorted_steady_ms"]["median"], res["S_binning_ms"]["median"]
    res["verdict"] = {"A_ms": a_ms, "S_plus_B_ms": s_ms + b_ms, "binning_wins": bool(s_ms + b_ms < a_ms),
                      "A_points_per_s": n / (a_ms * 1e-3), "B_points_per_s": n / (b_ms * 1e-3)}
    print(json.dumps(res))
    os.makedirs(os.path.dirname(a.out), exist_ok=True)
    json.dump(res, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
