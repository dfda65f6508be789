#!/usr/bin/env python
"""benchmarks/cpu_host_loops.py -- the host-loop baseline of SURVEY 8(d) / BASELINE.json north_star ("a straightforward
OpenMP host loop of identical semantics on the box's own cores") for the five BASELINE configs, on BOUNDED samples of
each workload (a few seconds of CPU work per line), next to the GPU numbers of benchmarks/run_configs.py.

The loops are the CPU oracle (oracle/gvoracle.c, gcc -O2 -fopenmp; test infrastructure, used here only as the thing
timed on the host, never on the product path). Each line states the threads its loop really uses: the map x map collide
and the point transform are `omp parallel for` loops and are timed with all host threads and with one; voxelisation,
list x map and the planner sequence are sequential loops there (the defined, race-free semantics) and say threads = 1.
No GPU is touched. Reported baselines, not optimisation targets.

    python benchmarks/cpu_host_loops.py [--out gpurun_out/cpu_host_loops.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import gvoracle as orc  # noqa: E402

BIT, PROB = orc.MAP_BIT, orc.MAP_PROB


def cpu_model():
    try:
        for ln in open("/proc/cpuinfo"):
            if ln.startswith("model name"):
                return ln.split(":", 1)[1].strip()
    except OSError:
        pass
    return "unknown"


def best_of(fn, reps=3):
    ts = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        ts.append(time.perf_counter() - t0)
    return min(ts), out


def collide_both(a, ta, b, tb, n, thr, threads_all):
    """map x map collide with all threads and with one (mark = False so that repetitions see the same maps)"""
    out = {}
    for label, t in (("all_threads", threads_all), ("one_thread", 1)):
        orc.set_num_threads(t)
        dt, cnt = best_of(lambda: orc.collide_map_map(a, ta, b, tb, n, thr, 0, False))
        out[label] = {"threads": t, "ms": dt * 1e3, "voxels_per_s": n / dt, "collisions": int(cnt)}
    orc.set_num_threads(threads_all)
    return out


def cloud(seed, n, lo, hi):
    return np.random.Generator(np.random.PCG64(seed)).uniform(lo, hi, size=(n, 3)).astype(np.float32)


def c1(T):
    dims, side, n = (128, 128, 128), 0.01, 128 ** 3
    pa, pb = cloud(101, 100000, -0.02, 1.30), cloud(102, 100000, -0.02, 1.30)
    a, b = orc.new_map(BIT, dims), orc.new_map(PROB, dims)
    orc.insert_points(b, PROB, dims, side, pb, 1)
    dt, _ = best_of(lambda: orc.insert_points(a, BIT, dims, side, pa, 1))
    return {"config": "C1 128^3 BitVector x Prob, 100k-point cloud (full size)",
            "insert": {"threads": 1, "ms": dt * 1e3, "points_per_s": 1e5 / dt},
            "collide": collide_both(a, BIT, b, PROB, n, 1.0, T)}


def c2(T):
    import bench as B
    dims, z0, z1 = (512, 512, 512), 96, 128
    n = 512 * 512 * (z1 - z0)
    env, rob = orc.new_map(PROB, dims, z0, z1), orc.new_map(BIT, dims, z0, z1)
    clouds = B.robot_clouds()
    orc.insert_meta(rob, BIT, dims, B.SIDE, np.concatenate(clouds), [len(c) for c in clouds], list(range(4, 11)), None, z0, z1)
    frame, pose = B.depth_frame(201), B.sensor_pose()
    dt, _ = best_of(lambda: orc.update_logodds(env, dims, B.SIDE, frame, pose, 72, z0, z1))
    return {"config": "C2 512^3 Prob x BitVector + 640x480 log-odds frame; sample: z-slab [96,128) = %d voxels" % n,
            "update_logodds": {"threads": 1, "ms": dt * 1e3, "points_per_s": len(frame) / dt},
            "collide": collide_both(env, PROB, rob, BIT, n, 0.5, T)}


def c3(T):
    dims, side = (256, 256, 256), 0.01
    rng = np.random.Generator(np.random.PCG64(301))
    links = [rng.uniform(0.0, 0.3, size=(50000, 3)).astype(np.float32) for _ in range(7)]
    env = orc.new_map(BIT, dims)
    orc.insert_points(env, BIT, dims, side, cloud(303, 200000, 0.2, 2.3), 1)
    n_cfg = 2
    tr = np.tile(np.eye(4, dtype=np.float32), (n_cfg, 7, 1, 1))
    tr[:, :, :3, 3] = rng.uniform(0.3, 2.0, size=(n_cfg, 7, 3))
    mask = np.zeros(32, np.uint8)
    for b in range(6, 11):
        mask[b >> 3] |= 1 << (b & 7)
    xyz = np.concatenate(links)
    dt, counts = best_of(lambda: orc.batch_robot_collide(env, BIT, dims, side, xyz, [50000] * 7, list(range(4, 11)), tr, mask, 1.0), reps=2)
    return {"config": "C3 7-link arm (350k pts) vs 256^3 environment, planner sequence; sample: %d of 1024 configurations" % n_cfg,
            "per_configuration": {"threads": 1, "ms": dt * 1e3 / n_cfg, "configurations_per_s": n_cfg / dt,
                                  "points_per_s": n_cfg * 350000 / dt},
            "how": "clearMap(robot) + transform + insertMetaPointCloud + collideWith(env) per configuration, literally"}


def c4(T):
    dims, z0, z1 = (2048, 2048, 2048), 1024, 1032
    n = 2048 * 2048 * (z1 - z0)
    a, b = orc.new_map(PROB, dims, z0, z1), orc.new_map(PROB, dims, z0, z1)
    pts = cloud(401, 1000000, 0.0, 20.48)
    pts[:, 2] = np.random.Generator(np.random.PCG64(402)).uniform(10.24, 10.32, size=len(pts)).astype(np.float32)
    orc.insert_points(b, PROB, dims, 0.01, pts[::2], 1, z0, z1)
    dt, _ = best_of(lambda: orc.insert_points(a, PROB, dims, 0.01, pts, 1, z0, z1))
    return {"config": "C4 2048^3 Prob x Prob + 1M-point clouds; sample: z-slab [1024,1032) = %d voxels, one 1M-point cloud" % n,
            "insert": {"threads": 1, "ms": dt * 1e3, "points_per_s": len(pts) / dt},
            "collide": collide_both(a, PROB, b, PROB, n, 0.5, T)}


def c5(T):
    dims, side, z1 = (1024, 1024, 1024), 0.01, 32   # the first 32 z-layers: a 1 GiB BitVectorVoxelMap
    m = orc.new_map(BIT, dims, 0, z1)
    pts = cloud(501, 1200000, 0.0, 10.24)
    pts[:, 2] = np.random.Generator(np.random.PCG64(502)).uniform(0.0, z1 * side, size=len(pts)).astype(np.float32)
    orc.insert_points(m, BIT, dims, side, pts[:600000], 1, 0, z1)
    lst = orc.VoxelList(dims, side)
    t0 = time.perf_counter()
    lst.insert_points(pts[300000:], 7)
    build = time.perf_counter() - t0
    n_map = 1024 * 1024 * z1
    dt, res = best_of(lambda: lst.collide(0, m, BIT, 1.0, 0, False, None, 0, n_map))
    dt3, _ = best_of(lambda: lst.collide(3, m, BIT, 1.0, 0, False, None, 0, n_map))
    return {"config": "C5 VoxelList x 1024^3 BitVectorVoxelMap; sample: %d list entries vs the first 32 z-layers (1 GiB)" % len(lst),
            "list_build": {"threads": 1, "ms": build * 1e3, "points_per_s": 900000 / build},
            "collide": {"threads": 1, "ms": dt * 1e3, "entries_per_s": len(lst) / dt, "collisions": int(res[0]) if isinstance(res, tuple) else int(res)},
            "per_meaning": {"threads": 1, "ms": dt3 * 1e3, "entries_per_s": len(lst) / dt3}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "cpu_host_loops.json"))
    ap.add_argument("--configs", default="c1,c2,c3,c4,c5")
    args = ap.parse_args()
    orc.lib()
    T = orc.num_threads()
    res = {"cpu": cpu_model(), "host_threads": T, "os_cpus": os.cpu_count(),
           "loops": "oracle/gvoracle.c, gcc -O2 -fopenmp (bounded samples, best of 3)", "results": []}
    table = {"c1": c1, "c2": c2, "c3": c3, "c4": c4, "c5": c5}
    for name in args.configs.split(","):
        t0 = time.perf_counter()
        r = table[name](T)
        r["wall_s"] = time.perf_counter() - t0
        print(json.dumps(r), flush=True)
        res["results"].append(r)
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
