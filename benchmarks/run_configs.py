#!/usr/bin/env python
"""benchmarks/run_configs.py -- the five BASELINE.json configurations at full size on ONE B200, every number next to
the roofline its kernel is bound by and, where the reference can represent the case (< 4 GiB per map, 32-bit voxel
counts), next to the reference's own CUDA kernels rebuilt for sm_100a (oracle/_ref/libgvref_cuda.so) timed in the
same process with the reference's launch shapes and result path.

    python benchmarks/run_configs.py [--configs c1,c3,c4,c5] [--out profiles/rNN_configs.json]

bench.py stays the driver-facing headline (configs[1]); this script is the per-config evidence DESIGN.md quotes.
Timing: CUDA events on libgvx's stream, warm-up first, inputs far larger than L2 unless the line says "L2-resident".
"""
import argparse
import ctypes as C
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib  # noqa: E402

PEAK_GBS = 6449.0
try:
    PEAK_GBS = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
except Exception:
    pass

L = None
STREAM = None
REF = None


def setup():
    global L, STREAM, REF
    torch.cuda.set_device(0)
    L = _lib.init(0)
    STREAM = torch.cuda.Stream()
    torch.cuda.set_stream(STREAM)
    _lib.check(L.gvx_set_stream(C.c_void_p(STREAM.cuda_stream)))
    so = os.path.join(ROOT, "oracle", "_ref", "libgvref_cuda.so")
    if os.path.exists(so):
        REF = C.CDLL(so)
        REF.gvref_collide_map_map.restype = C.c_longlong
        REF.gvref_collide_list_map.restype = C.c_longlong


def p_(t):
    return C.c_void_p(t.data_ptr())


def time_gpu(fn, iters, warmup=3, flush=None):
    """mean / min ms of fn() over iters, CUDA events on STREAM; flush: optional tensor rewritten between iterations"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ms = []
    for _ in range(iters):
        if flush is not None:
            flush.add_(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(STREAM)
        fn()
        e1.record(STREAM)
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    return sum(ms) / len(ms), min(ms)


def time_wall(fn, iters, warmup=1):
    """for calls that synchronise internally (the reference's API shape): wall clock around fn(), device idle before"""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sum(ts) / len(ts), min(ts)


def roof(bytes_per_launch, ms):
    gbs = bytes_per_launch / (ms * 1e-3) / 1e9
    return {"achieved_GBs": gbs, "frac_of_measured_peak": gbs / PEAK_GBS, "frac_of_8000_nominal": gbs / 8000.0,
            "algorithmic_bytes": int(bytes_per_launch)}


def voxel_centres_on_device(ids, dims, side):
    """float32 (n,3) centres of linear voxel ids, computed on the device"""
    ids = ids.to(torch.int64)
    x = ids % dims[0]
    y = (ids // dims[0]) % dims[1]
    z = ids // (dims[0] * dims[1])
    c = torch.stack([x, y, z], -1).to(torch.float32)
    return ((c + 0.5) * side).contiguous()


def insert_device_points(m, pts_d, meaning):
    _lib.check(L.gvx_insert_points(m.handle, p_(pts_d), pts_d.shape[0], _lib.MEM_DEVICE, int(meaning), None))


def boxes_into_map(m, dims, side, n_boxes, box_vox, seed, meaning=1):
    """fill random axis-aligned boxes of box_vox^3 voxels (device-generated voxel centres)"""
    g = torch.Generator(device="cuda"); g.manual_seed(seed)
    r = torch.arange(box_vox, device="cuda")
    bx, by, bz = torch.meshgrid(r, r, r, indexing="ij")
    local = torch.stack([bx, by, bz], -1).reshape(-1, 3)
    for _ in range(n_boxes):
        o = torch.stack([torch.randint(0, dims[k] - box_vox, (1,), device="cuda", generator=g) for k in range(3)], -1)
        pts = ((local + o).to(torch.float32) + 0.5) * side
        insert_device_points(m, pts.contiguous(), meaning)
    torch.cuda.synchronize()


# ---- C1: 128^3 BitVectorVoxelMap @1cm, one 100k-point cloud, collideWith a 128^3 ProbVoxelMap -------------------
def config_c1():
    dims, side, n = (128, 128, 128), 0.01, 128 ** 3
    rng = np.random.Generator(np.random.PCG64(101))
    pa = rng.uniform(-0.02, 1.30, size=(100000, 3)).astype(np.float32)
    pb = np.random.Generator(np.random.PCG64(102)).uniform(-0.02, 1.30, size=(100000, 3)).astype(np.float32)
    a, b = gv.BitVectorVoxelMap(dims, side), gv.ProbVoxelMap(dims, side)
    pad, pbd = torch.from_numpy(pa).cuda(), torch.from_numpy(pb).cuda()
    insert_device_points(b, pbd, 1)
    cnt_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    flush = torch.zeros(160 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    ins = time_gpu(lambda: insert_device_points(a, pad, 1), 50)
    col = lambda: _lib.check(L.gvx_collide_map_map_async(a.handle, b.handle, 1.0, None, 1, p_(cnt_d)))  # noqa: E731
    hot = time_gpu(col, 50)
    cold = time_gpu(col, 30, flush=flush)
    host_cnt = C.c_uint64(0)
    sync_call = time_wall(lambda: _lib.check(L.gvx_collide_map_map(a.handle, b.handle, 1.0, None, 1, C.byref(host_cnt))), 50)
    host_ins = time_wall(lambda: a.insertPointCloud(pa, 1), 30)
    out = {"config": "C1 128^3 BitVector x Prob, 100k-point cloud", "collisions": int(host_cnt.value),
           "insert_device_ms": ins[0], "insert_points_per_s": 1e5 / (ins[0] * 1e-3),
           "insert_host_cloud_sync_ms": host_ins[0],
           "collide_L2_resident_ms": hot[0], "collide_L2_flushed_ms": cold[0],
           "collide_voxels_per_s_flushed": n / (cold[0] * 1e-3), "roofline_flushed": roof(0.25 * n, cold[0]),
           "collide_sync_call_latency_ms": sync_call[0]}
    if REF is not None:
        ra = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
        rb = torch.full((n,), -128, dtype=torch.int8, device="cuda")
        fl = C.c_int(0)
        REF.gvref_insert_points(p_(rb), 1, *dims, C.c_float(side), p_(pbd), 100000, 1, C.byref(fl))
        r_ins = time_wall(lambda: REF.gvref_insert_points(p_(ra), 0, *dims, C.c_float(side), p_(pad), 100000, 1, C.byref(fl)), 30)
        r_cnt = [0]

        def rc():
            r_cnt[0] = REF.gvref_collide_map_map(p_(ra), 0, p_(rb), 4, n, C.c_float(1.0))
        r_col = time_wall(rc, 30)
        out["reference_cuda"] = {"insert_call_ms": r_ins[0], "collide_call_ms": r_col[0], "collisions": int(r_cnt[0]),
                                 "speedup_collide_call": r_col[0] / sync_call[0], "speedup_insert_call": r_ins[0] / max(ins[0], 1e-6)}
        assert r_cnt[0] == host_cnt.value, (r_cnt[0], host_cnt.value)
    return out


# ---- C3: 7-link arm, 50k pts/link, per-link meanings, 1024 planner configurations vs a 256^3 environment ---------
def arm_links(seed=301, n_per_link=50000):
    rng = np.random.Generator(np.random.PCG64(seed))
    links, lengths = [], []
    for k in range(7):
        r, l = rng.uniform(0.04, 0.08), rng.uniform(0.10, 0.40)
        t, a = rng.uniform(0, l, n_per_link), rng.uniform(0, 2 * np.pi, n_per_link)
        links.append(np.stack([r * np.cos(a), r * np.sin(a), t], -1).astype(np.float32))
        lengths.append(l)
    return links, lengths


def arm_transforms(lengths, n_cfg, seed=302, base=(1.28, 1.28, 0.4)):
    """fixed DH-like chain: joint k rotates about x or z alternately, link k extends along its local z (host fp32)"""
    rng = np.random.Generator(np.random.PCG64(seed))
    q = rng.uniform(-np.pi, np.pi, size=(n_cfg, 7)).astype(np.float32)
    T = np.zeros((n_cfg, 7, 4, 4), np.float32)
    for c in range(n_cfg):
        M = np.eye(4, dtype=np.float32); M[:3, 3] = base
        for k in range(7):
            cs, sn = np.float32(np.cos(q[c, k] * 0.5)), np.float32(np.sin(q[c, k] * 0.5))
            R = np.eye(4, dtype=np.float32)
            if k % 2 == 0:
                R[:3, :3] = [[cs, -sn, 0], [sn, cs, 0], [0, 0, 1]]
            else:
                R[:3, :3] = [[1, 0, 0], [0, cs, -sn], [0, sn, cs]]
            M = (M @ R).astype(np.float32)
            T[c, k] = M
            S = np.eye(4, dtype=np.float32); S[2, 3] = lengths[k]
            M = (M @ S).astype(np.float32)
    return T


def config_c3(n_cfg=1024):
    dims, side, n = (256, 256, 256), 0.01, 256 ** 3
    env = gv.BitVectorVoxelMap(dims, side)
    boxes_into_map(env, dims, side, 40, 24, 303)
    links, lengths = arm_links()
    T = arm_transforms(lengths, n_cfg)
    mpc = gv.MetaPointCloud(links)
    meanings = list(range(4, 11))
    mask = gv.BitVectorVoxel()
    for b in (6, 7, 8, 9, 10):
        mask.setBit(b)
    counts = [None]

    def batch():
        counts[0] = gv.batch_robot_collide(env, mpc, meanings, T, mask, 1.0)
    ms = time_wall(batch, 10, warmup=2)
    pts = 7 * 50000
    out = {"config": "C3 7-link arm (350k pts) x %d planner configurations vs 256^3 BitVector env, mask {6..10}" % n_cfg,
           "batch_call_ms_host_inputs": ms[0], "batch_call_ms_min": ms[1], "configurations_per_s": n_cfg / (ms[0] * 1e-3),
           "points_per_s": n_cfg * pts / (ms[0] * 1e-3), "colliding_configurations": int((counts[0] > 0).sum()),
           "collisions_total": int(counts[0].sum())}
    # the reference-style sequence on THIS library (clear + fused transform/insert + collide per configuration)
    rob = gv.BitVectorVoxelMap(dims, side)
    k_seq = 32

    def seq():
        for c in range(k_seq):
            rob.clearMap()
            rob.insertMetaPointCloud(mpc, meanings, transforms=T[c], check_outside=False)
            rob.collideWith(env, 1.0)
    ms_seq = time_wall(seq, 3)
    out["per_configuration_sequence_this_library_ms"] = ms_seq[0] / k_seq
    if REF is not None:
        r_rob = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
        r_env = torch.from_numpy(env.download().reshape(-1)).cuda()
        base_d = torch.from_numpy(np.concatenate(links)).cuda()
        moved = torch.empty_like(base_d)
        sizes = np.array([len(l) for l in links], np.uint32)
        mg = np.array(meanings, np.int32)
        offs = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int64)
        fl = C.c_int(0)
        k_ref = 16
        r_counts = []

        def ref_seq():
            r_counts.clear()
            for c in range(k_ref):
                REF.gvref_clear_map(p_(r_rob), 0, C.c_uint64(n))
                for k in range(7):
                    M = np.ascontiguousarray(T[c, k].reshape(16))
                    REF.gvref_transform_cloud(M.ctypes.data_as(C.c_void_p), C.c_void_p(base_d.data_ptr() + int(offs[k]) * 12),
                                              C.c_void_p(moved.data_ptr() + int(offs[k]) * 12), int(sizes[k]))
                REF.gvref_insert_meta(p_(r_rob), 0, *dims, C.c_float(side), p_(moved), sizes.ctypes.data_as(C.c_void_p), 7,
                                      mg.ctypes.data_as(C.c_void_p), C.byref(fl))
                r_counts.append(REF.gvref_collide_map_map(p_(r_rob), 0, p_(r_env), 0, n, C.c_float(1.0)))
        ms_ref = time_wall(ref_seq, 2)
        full = gv.batch_robot_collide(env, mpc, meanings, T[:k_ref], None, 1.0)
        out["reference_cuda"] = {"per_configuration_ms": ms_ref[0] / k_ref, "batch_of_%d_extrapolated_ms" % n_cfg: ms_ref[0] / k_ref * n_cfg,
                                 "speedup_vs_batch_call": (ms_ref[0] / k_ref * n_cfg) / ms[0],
                                 "counts_equal_no_mask": bool(np.array_equal(np.array(r_counts, np.uint64), full)),
                                 "note": "reference kernels + launch shapes; meanings 4..10 share bytes 0/1, so the "
                                         "reference's byte RMW may drop bits but voxel occupancy (the count) is unaffected"}
    return out


# ---- C4: 2048^3 ProbVoxelMap (8.6 GB), 8 streamed 1M-point clouds per frame, collide Prob x Prob ----------------
def config_c4(dz=2048):
    """One GPU holds the z-slab [0, dz) of the 2048^3 grid (dz = 2048: the whole map, 2 x 8.6 GB; dz = 256: the share
    of one of 8 GPUs). Not representable in the reference (voxel count > uint32, map > 4 GiB)."""
    dims, side = (2048, 2048, 2048), 0.01
    n = 2048 * 2048 * dz
    env = gv.ProbVoxelMap(dims, side, slab=(0, dz))
    rob = gv.ProbVoxelMap(dims, side, slab=(0, dz))
    g = torch.Generator(device="cuda"); g.manual_seed(401)
    clouds = []
    for k in range(8):
        u = torch.rand((700000, 3), device="cuda", generator=g) * 20.48
        centres = torch.rand((50, 3), device="cuda", generator=g) * 20.48
        cl = centres[torch.randint(0, 50, (300000,), device="cuda", generator=g)] + \
            torch.randn((300000, 3), device="cuda", generator=g) * 0.05
        clouds.append(torch.cat([u, cl]).to(torch.float32).contiguous())
    robot_pts = torch.cat([clouds[0][:1000000], torch.rand((1000000, 3), device="cuda", generator=g) * 20.48]).contiguous()
    insert_device_points(rob, robot_pts, 1)
    pose = np.ascontiguousarray(np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32).reshape(-1))
    cnt_d = torch.zeros(1, dtype=torch.int64, device="cuda")

    def frame():
        for c in clouds:
            _lib.check(L.gvx_update_logodds(env.handle, p_(c), c.shape[0], _lib.MEM_DEVICE, pose.ctypes.data_as(C.c_void_p), 72))
    ins = time_gpu(frame, 10)
    col = lambda: _lib.check(L.gvx_collide_map_map_async(env.handle, rob.handle, 0.5, None, 1, p_(cnt_d)))  # noqa: E731
    ms = time_gpu(col, 10)
    clr = time_gpu(lambda: env.clearMap(), 5)
    out = {"config": "C4 2048x2048x%d slab of the 2048^3 ProbVoxelMap, 8 x 1M-point clouds per frame, Prob x Prob collide thr 0.5" % dz,
           "voxels": n, "frame_insert_ms": ins[0], "insert_points_per_s": 8e6 / (ins[0] * 1e-3),
           "collide_ms": ms[0], "collide_ms_min": ms[1], "collide_voxels_per_s": n / (ms[0] * 1e-3),
           "roofline": roof(0.25 * n, ms[0]),
           "roofline_note": "default path: the two 'known' summaries (2 x 1 bit per voxel) + the Prob bytes of candidate "
                            "groups; the reference's streaming pattern would read 2 B per voxel (%.2f GB)" % (2 * n / 1e9),
           "clearMap_ms": clr[0], "clearMap_GBs": n / (clr[0] * 1e-3) / 1e9,
           "collisions": int(cnt_d.item()), "reference_cuda": "not representable in the reference (uint32 voxel count, 4 GiB map limit)"}
    env.destroy(); rob.destroy()
    return out


# ---- C5: sparse VoxelList of 10M voxels vs 1024^3 dense BitVectorVoxelMap (32 GiB), per-meaning counts ----------
def config_c5(n_list=10_000_000):
    dims, side = (1024, 1024, 1024), 0.01
    n = 1024 ** 3
    m = gv.BitVectorVoxelMap(dims, side)
    boxes_into_map(m, dims, side, 200, 47, 503)  # 200 x 47^3 = 20.8 M voxels ~ 2 % of the grid
    g = torch.Generator(device="cuda"); g.manual_seed(501)
    ids = torch.unique(torch.randint(0, n, (int(n_list * 1.02),), device="cuda", generator=g, dtype=torch.int64))[:n_list]
    n_list = ids.numel()
    lst = gv.BitVectorVoxelList(dims, side)
    # 1-3 meanings from 4..35 per voxel: three passes over (subsets of) the same voxels; times the list build too
    centres = voxel_centres_on_device(ids, dims, side)
    t0 = time.perf_counter()
    for rnd, frac in enumerate((1.0, 0.5, 0.25)):
        sel = centres if frac == 1.0 else centres[torch.rand(n_list, device="cuda", generator=g) < frac].contiguous()
        mg = torch.randint(4, 36, (8,), device="cuda", generator=g).cpu().numpy()
        step = (sel.shape[0] + 7) // 8
        for k in range(8):
            part = sel[k * step:(k + 1) * step]
            if part.shape[0]:
                _lib.check(L.gvx_list_insert_points(lst.handle, p_(part.contiguous()), part.shape[0], _lib.MEM_DEVICE, int(mg[k])))
    torch.cuda.synchronize()
    build_s = time.perf_counter() - t0
    assert len(lst) == n_list
    # one more insert of 1M points into the 10M list, timed alone (sort + merge of 11M entries)
    extra = centres[:1000000].contiguous()
    ins1 = time_wall(lambda: _lib.check(L.gvx_list_insert_points(lst.handle, p_(extra), extra.shape[0], _lib.MEM_DEVICE, 40)), 3)
    cnt_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    col = lambda: _lib.check(L.gvx_collide_list_map_async(lst.handle, m.handle, 1.0, None, 0, p_(cnt_d)))  # noqa: E731
    ms = time_gpu(col, 20)
    colm = lambda: _lib.check(L.gvx_collide_list_map_async(lst.handle, m.handle, 1.0, None, 1, p_(cnt_d)))  # noqa: E731
    ms_mark = time_gpu(colm, 20)
    per = [None]

    def pm():
        per[0] = lst.collideCountingPerMeaning(m, None, 1.0, None, mark=False)
    ms_pm = time_wall(pm, 10)
    ty = gv.BitVectorVoxel()
    ms_ty = time_wall(lambda: lst.collideWithTypes(m, ty, 1.0, None, mark=False), 10)
    out = {"config": "C5 VoxelList of %d voxels x 1024^3 BitVectorVoxelMap (32 GiB, ~2%% occupied)" % n_list,
           "list_build_s_%d_points_in_24_inserts" % int(centres.shape[0] * 1.75): build_s,
           "list_insert_1M_into_10M_ms": ins1[0],
           "collide_ms": ms[0], "collide_entries_per_s": n_list / (ms[0] * 1e-3), "roofline_8B_per_entry": roof(8 * n_list, ms[0]),
           "roofline_note": "default path: 4 B id + one 4 B summary word gathered per entry (a 32 B DRAM sector; the ids are "
                            "sorted, so neighbouring entries share sectors: the whole 128 MB summary is read about once), "
                            "list voxels only for the ~2 % of entries whose map voxel is occupied; a gather-latency-bound "
                            "kernel, not a streaming one. The reference's pattern (4 + 32 + 32 B per entry) is 680 MB",
           "collide_mark_ms": ms_mark[0], "per_meaning_call_ms": ms_pm[0], "types_call_ms": ms_ty[0],
           "collisions": int(cnt_d.item()), "per_meaning_total": int(per[0][0]),
           "meanings_in_collision": [int(k) for k in np.flatnonzero(per[0][1])][:40],
           "reference_cuda": "map of 32 GiB is refused by the reference (4 GiB limit)"}
    m.destroy(); lst.destroy()
    return out


# ---- C2r: configs[1]'s frame with the complete sensor model (free-space ray casting) ---------------------------
def config_c2_raycast():
    import bench as B
    dims, side = (512, 512, 512), 0.01
    frames = [torch.from_numpy(B.depth_frame(201 + k)).cuda() for k in range(4)]
    pose = np.ascontiguousarray(B.sensor_pose().reshape(-1))
    m = gv.ProbVoxelMap(dims, side)
    k = [0]

    def frame(raycast):
        f = frames[k[0] % 4]; k[0] += 1
        _lib.check(L.gvx_insert_sensor_data(m.handle, p_(f), f.shape[0], _lib.MEM_DEVICE, pose.ctypes.data_as(C.c_void_p),
                                            int(raycast), 0, 1, -10, 72))
    first = time_gpu(lambda: frame(1), 1, warmup=0)   # unknown map: every cell of every ray is a first touch
    steady = time_gpu(lambda: frame(1), 30)           # the cone is saturated at -127: reads only
    hits = time_gpu(lambda: frame(0), 30)
    mm = m.download()
    out = {"config": "C2r 512^3 ProbVoxelMap, 640x480 frame, insertSensorData WITH ray casting",
           "points": int(frames[0].shape[0]), "first_frame_ms": first[0], "steady_state_ms": steady[0],
           "hits_only_ms": hits[0], "known_voxels": int((mm != -128).sum()), "free_voxels": int((mm == -127).sum()),
           "rays_per_s_steady": frames[0].shape[0] / (steady[0] * 1e-3)}
    if REF is not None:
        n = dims[0] * dims[1] * dims[2]
        rmap = torch.full((n,), -128, dtype=torch.int8, device="cuda")
        tmp = torch.empty_like(frames[0])

        def ref_frame():
            f = frames[k[0] % 4]; k[0] += 1
            REF.gvref_insert_sensor_data_full(p_(rmap), *dims, C.c_float(side), pose.ctypes.data_as(C.c_void_p), p_(f),
                                              p_(tmp), f.shape[0], 1, None, 1, 0)
        r = time_wall(ref_frame, 10, warmup=2)
        out["reference_cuda"] = {"frame_call_ms": r[0], "speedup_steady": r[0] / steady[0],
                                 "note": "the reference's kernelTransformSensorData + kernelInsertSensorData<RayCaster> "
                                         "(non-atomic, racy) rebuilt for sm_100a, same frames"}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="c1,c2r,c3,c4,c4s,c5")
    ap.add_argument("--out", default=os.path.join(ROOT, "gpurun_out", "configs.json"))
    args = ap.parse_args()
    setup()
    res = {"gpu": torch.cuda.get_device_name(0), "peak_GBs_measured": PEAK_GBS, "results": []}
    fns = {"c1": config_c1, "c3": config_c3, "c4": config_c4, "c4s": lambda: config_c4(256), "c5": config_c5,
           "c2r": config_c2_raycast}
    for name in args.configs.split(","):
        t0 = time.time()
        try:
            r = fns[name]()
        except Exception as e:  # keep the other configs' numbers
            import traceback
            r = {"config": name, "error": repr(e), "trace": traceback.format_exc()[-1500:]}
        r["wall_s"] = time.time() - t0
        res["results"].append(r)
        print(json.dumps(r), flush=True)
        torch.cuda.empty_cache()
    os.makedirs(os.path.dirname(args.out), exist_ok=True)
    json.dump(res, open(args.out, "w"), indent=1)


if __name__ == "__main__":
    main()
