#!/usr/bin/env python
"""benchmarks/tma_probe.py -- A/B of the voxel-streaming collideWith kernels at 512^3 (BASELINE configs[1]):
per-thread 256-bit loads (k_collide_voxelwise, gvx_collide.cu) against TMA bulk tiles into a shared-memory ring
(k_collide_tma, gvx_collide_tma.cu; shapes 1..5 = tile x stages 512x4, 1024x3, 256x6, 512x2, 1024x2).

Both read every voxel of both grids (GVX_OPT_USE_SUMMARY = 0): 33 B/voxel for Prob x Bit and Bit x Prob = 4.43 GB per
launch, 64 B/voxel for Bit x Bit = 8.59 GB. P rotating map pairs (P x 4.4 GB >> 126 MB of L2), CUDA events on the
launching stream around `reps` back-to-back launches, counts compared between the variants. One JSON line per variant
and a summary line; written to gpurun_out/tma_probe.json when that directory exists."""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib  # noqa: E402

DIMS = (512, 512, 512)
N = DIMS[0] * DIMS[1] * DIMS[2]
SIDE = 0.01
P = 4


def cloud(seed, n, lo, hi):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)


def main():
    L = _lib.init(0)
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    _lib.check(L.gvx_set_stream(C.c_void_p(stream.cuda_stream)))
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 40
    only = sys.argv[2] if len(sys.argv) > 2 else None  # "bitprob": Bit x Prob with shapes 0 and 2 only (ncu captures)
    peak = 6448.1
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        pass
    probs, bits = [], []
    for k in range(P):
        pm, bm = gv.ProbVoxelMap(DIMS, SIDE), gv.BitVectorVoxelMap(DIMS, SIDE)
        pm.insertPointCloud(cloud(10 + k, 400000, 1.0, 4.0), gv.eBVM_OCCUPIED)
        bm.insertPointCloud(cloud(20 + k, 400000, 1.5, 3.5), 7)
        bm.insertPointCloud(cloud(30 + k, 100000, 0.0, 5.12), gv.eBVM_FREE)  # FREE alone is not occupancy
        probs.append(pm)
        bits.append(bm)
    bits2 = []
    for k in range(2):  # Bit x Bit partners (2 more 4 GiB maps)
        bm = gv.BitVectorVoxelMap(DIMS, SIDE)
        bm.insertPointCloud(cloud(40 + k, 400000, 1.5, 3.5), 9)
        bits2.append(bm)
    count_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    cp = C.c_void_p(count_d.data_ptr())
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    pairs = {
        "Prob x Bit (mark)": ([(probs[k].handle, bits[k].handle) for k in range(P)], 33, 1),
        "Bit x Prob": ([(bits[k].handle, probs[k].handle) for k in range(P)], 33, 0),
        "Bit x Bit": ([(bits[k].handle, bits2[k % 2].handle) for k in range(P)], 64, 0),
    }
    results = []
    for name, (hs, bpv, mark) in pairs.items():
        base_counts = None
        if only == "bitprob" and name != "Bit x Prob":
            continue
        for shape in ((0, 2) if only else (0, 1, 2, 3, 4, 5)):
            if shape > 1 and name == "Bit x Bit":
                continue  # one shape only for Bit x Bit (256 x 4)
            gv.set_option(gv.GVX_OPT_COLLIDE_TMA, shape)

            def go(k):
                _lib.check(L.gvx_collide_map_map_async(hs[k % P][0], hs[k % P][1], 0.5, None, mark, cp))
            counts = []
            for k in range(P):  # warm-up, and the counts of every pair
                go(k)
                torch.cuda.synchronize()
                counts.append(int(count_d.item()))
            if base_counts is None:
                base_counts = counts
            assert counts == base_counts, (name, shape, counts, base_counts)
            best = None
            for _ in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for k in range(reps):
                    go(k)
                e1.record(stream)
                torch.cuda.synchronize()
                ms = e0.elapsed_time(e1) / reps
                best = ms if best is None else min(best, ms)
            gbs = bpv * N / (best * 1e-3) / 1e9
            r = {"pair": name, "kernel": "k_collide_voxelwise (LDG.256)" if shape == 0 else "k_collide_tma shape %d" % shape,
                 "ms_per_launch": round(best, 4), "GB/s": round(gbs, 1), "frac_of_measured_peak": round(gbs / peak, 3),
                 "frac_of_8000": round(gbs / 8000, 3), "counts": counts}
            print(json.dumps(r), flush=True)
            results.append(r)
    gv.set_option(gv.GVX_OPT_COLLIDE_TMA, -1)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    out = os.path.join(os.path.dirname(__file__), "..", "gpurun_out")
    if os.path.isdir(out) and not only:
        json.dump({"grid": DIMS, "pairs_in_rotation": P, "reps": reps, "peak_gbs": peak, "results": results},
                  open(os.path.join(out, "tma_probe.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
