#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: collideWith voxels/s, % HBM peak; insert points/s).

One step = one pass of the hot path over one batch: a 640x480 synthetic depth frame (307 200 points) is folded into
the 512^3 ProbVoxelMap with the log-odds update, then that map is collided with the robot BitVectorVoxelMap
(BASELINE.json configs[1]). value = voxels collided per second over the whole job.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

N > 1 (under torchrun, one rank per GPU, NCCL): the grid is 512 x 512 x (512*N), rank r owns z-slab r (weak scaling:
per-GPU work fixed), every rank reads the whole N-frame cloud and voxelises only its slab, the collision counts are
all-reduced (one int64) over NVLink.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DX = DY = DZ = 512
SIDE = 0.01
FRAME_W, FRAME_H = 640, 480
N_FRAME = FRAME_W * FRAME_H
BYTES_PER_VOXEL_SUMMARY = 0.25  # default path: 1 summary bit per voxel and side ("known" bit of the ProbVoxelMap,
#                                  occupancy bit of the BitVectorVoxelMap); Prob bytes only for candidate groups
BYTES_PER_VOXEL = 33  # Prob (1 B) x BitVector (32 B): algorithmic bytes of the dominant kernel
THR = 0.5
N_PAIRS = 8  # rotating (environment, robot) pairs and frames: 8 x 33.5 MB of summaries per rotation vs 126 MB of L2


# ---- synthetic inputs (seeded; SURVEY 8d "C2") ---------------------------------------------------------------
def depth_frame(seed):
    """640x480 pinhole depth image (fx=fy=525, cx=319.5, cy=239.5): back wall at 2.5 m + 3 boxes + +-1 cm noise,
    2 % NaN pixels; points in the SENSOR frame (z forward)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    u, v = np.meshgrid(np.arange(FRAME_W, dtype=np.float32), np.arange(FRAME_H, dtype=np.float32))
    z = np.full(u.shape, 2.5, np.float32)
    for (u0, u1, v0, v1, d) in ((60, 220, 200, 420, 1.4), (300, 420, 120, 300, 1.9), (470, 600, 260, 470, 1.1)):
        z[v0:v1, u0:u1] = d
    z = (z + rng.uniform(-0.01, 0.01, size=z.shape)).astype(np.float32)
    f = np.float32(525.0)
    pts = np.stack([(u - np.float32(319.5)) * z / f, (v - np.float32(239.5)) * z / f, z], -1).reshape(-1, 3)
    pts = pts.astype(np.float32)
    pts[rng.uniform(size=len(pts)) < 0.02] = np.nan
    return np.ascontiguousarray(pts)


def sensor_pose(z_shift=0.0):
    """sensor at (0.3, 2.56, 2.2 + shift) looking along +x, image-down = world -z; row-major [R|t]"""
    return np.array([[0, 0, 1, 0.3], [-1, 0, 0, 2.56], [0, -1, 0, 2.2 + z_shift]], np.float32)


def robot_clouds(seed=301, z_shift=0.0):
    """7 link clouds of 50 000 surface points each (capsules, r 4-8 cm, l 10-40 cm): an arm that starts in front of
    the sensor and reaches +x through the face of the first box of the depth scene (so there ARE collisions);
    meanings 4..10 (SURVEY 8d "C3" arm, inserted once for C2)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    clouds, base = [], np.array([1.25, 3.0, 2.0 + z_shift])
    for k in range(7):
        r, l = rng.uniform(0.04, 0.08), rng.uniform(0.10, 0.40)
        t = rng.uniform(0, l, 50000); a = rng.uniform(0, 2 * np.pi, 50000)
        d = np.array([1.0, rng.uniform(-0.25, 0.25), rng.uniform(-0.25, 0.25)]); d /= np.linalg.norm(d)
        e1 = np.cross(d, [0, 0, 1.0]); e1 /= np.linalg.norm(e1); e2 = np.cross(d, e1)
        p = base + np.outer(t, d) + r * (np.outer(np.cos(a), e1) + np.outer(np.sin(a), e2))
        clouds.append(np.ascontiguousarray(p, dtype=np.float32))
        base = base + l * d * 0.9
    return clouds


# ---- clocks sampling during the timed region (B200_PROFILING.md) ---------------------------------------------
class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index, self.samples, self._stop, self._t = index, [], threading.Event(), None

    def _run(self):
        while not self._stop.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                if len(f) >= 6:
                    self.samples.append(f)
            except Exception:
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        mhz = sorted(int(s[0]) for s in self.samples if s[0].isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": mhz[len(mhz) // 2] if mhz else None,
                "sm_max_mhz": int(self.samples[0][1]) if self.samples[0][1].isdigit() else None,
                "reasons": reasons, "samples": len(self.samples)}


def measured_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ---- CPU arms (the ONLY places bench.py touches oracle/) ------------------------------------------------------
SAMPLE_DZ = 32  # z sub-slab of the 512^3 grid used as the bounded CPU sample (8.4 M voxels, 277 MB of maps)


def _cpu_sample_maps(orc):
    """env/robot z-slab [96,128) of the same workload, built with the oracle (untimed)."""
    dims, z0, z1 = (DX, DY, DZ), 96, 96 + SAMPLE_DZ
    env = orc.new_map(orc.MAP_PROB, dims, z0, z1)
    rob = orc.new_map(orc.MAP_BIT, dims, z0, z1)
    clouds = robot_clouds()
    orc.insert_meta(rob, orc.MAP_BIT, dims, SIDE, np.concatenate(clouds), [len(c) for c in clouds],
                    list(range(4, 11)), None, z0, z1)
    return dims, z0, z1, env, rob


def cpu_arm(steps, warmup, use_ref_so=True, threads=None):
    """One step on the host cores: log-odds update of one 307 200-point frame + collideWith (Prob x BitVector) on a
    bounded z-sub-slab sample. Collide runs through oracle/_ref/libgvref_cpu.so (the reference's own voxel/collider
    headers compiled with g++, OpenMP over all host threads) when it exists, else through the oracle port."""
    from oracle import gvoracle as orc
    dims, z0, z1, env, rob = _cpu_sample_maps(orc)
    n_s = DX * DY * SAMPLE_DZ
    ref = None
    so = os.path.join(ROOT, "oracle", "_ref", "libgvref_cpu.so")
    if use_ref_so and os.path.exists(so):
        ref = C.CDLL(so)
        ref.gvrefcpu_collide_map_map.restype = C.c_uint64
        ref.gvrefcpu_collide_map_map.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, C.c_float,
                                                 C.c_int64, C.c_int]
        if threads:
            ref.gvrefcpu_set_num_threads(threads)
        cores = ref.gvrefcpu_num_threads()
    else:
        if threads:
            orc.set_num_threads(threads)
        cores = orc.num_threads()
    frames = [depth_frame(201 + k) for k in range(2)]
    pose = sensor_pose()

    def step(k):
        orc.update_logodds(env, dims, SIDE, frames[k % 2], pose, 72, z0, z1)
        if ref is not None:
            return ref.gvrefcpu_collide_map_map(env.ctypes.data, 4, rob.ctypes.data, 0, n_s, THR, 0, 1)
        return orc.collide_map_map(env, orc.MAP_PROB, rob, orc.MAP_BIT, n_s, THR, 0, True)

    for k in range(warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(steps):
        cnt = step(warmup + k)
    dt = time.perf_counter() - t0
    return {"value": n_s * steps / dt, "unit": "voxels/s", "cores": int(cores),
            "kind": "reference" if ref is not None else "port",
            "sample": "z-slab [96,128) of the 512^3 grid (%d voxels/step, Prob x BitVector collide + one 307200-point "
                      "log-odds frame), %d steps; collide = %s" % (n_s, steps, "reference voxel/collider headers (g++ -O2 "
                      "-fopenmp)" if ref is not None else "oracle C port (gcc -O2 -fopenmp)"),
            "ms_per_step": dt / steps * 1e3, "last_count": int(cnt)}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    r = cpu_arm(args.steps, args.warmup)
    line = {"impl": "reference", "metric": "collideWith voxels/s", "value": r["value"], "unit": "voxels/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "voxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0,
            "note": "the reference has no CPU path and its library cannot be built here (Boost/icl_core absent); this arm "
                    "runs the reference's own host-callable voxel/collider code in an OpenMP loop on the host cores"}
    print(json.dumps(line))


def workload_config(n_gpus):
    return {"workload": "BASELINE.json configs[1]: 512^3 ProbVoxelMap @1cm, per step one 640x480 depth frame (307200 pts) "
                        "log-odds insert then collideWith a 512^3 robot BitVectorVoxelMap (7 links, 350k pts)",
            "grid": [DX, DY, DZ * n_gpus], "voxels_per_gpu": DX * DY * DZ, "points_per_step": N_FRAME * n_gpus,
            "coll_threshold": THR, "mark_collisions": True,
            "l2": "inputs larger than L2: every timed loop rotates over %d (env, robot) map pairs and %d sensor frames, "
                  "%d x 33.5 MB of summaries (default path) or %d x 4.43 GB of grids (streaming) per rotation vs 126 MB "
                  "of L2; the single-launch figure flushes L2 (rewrites a 256 MB buffer) before every launch"
                  % (N_PAIRS, N_PAIRS, N_PAIRS, N_PAIRS),
            "collide_path": "summaries: 'known' bits of the ProbVoxelMap x occupancy bits of the BitVectorVoxelMap "
                            "(both maintained by the insert kernels), Prob bytes only for candidate groups",
            "parallelism": "z-slab x%d" % n_gpus if n_gpus > 1 else "single GPU"}


# ---- the GPU arm ------------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    import torch
    import torch.distributed as dist
    import gpu_voxels_b200 as gv
    from gpu_voxels_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        # NCCL announces its version on STDOUT when the first communicator comes up; the contract is ONE JSON line on
        # stdout, so fd 1 points at stderr until the communicator exists.
        sys.stdout.flush()
        saved_fd = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
            warm = torch.zeros(1, device="cuda")
            dist.all_reduce(warm)
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved_fd, 1)
            os.close(saved_fd)
    L = _lib.init(local_rank)
    # one explicit (non-default) stream shared by libgvx, torch events and NCCL, so CUDA events time libgvx's kernels
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    assert stream.cuda_stream != 0
    _lib.check(L.gvx_set_stream(C.c_void_p(stream.cuda_stream)))
    if args.pdl is not None:
        _lib.check(L.gvx_set_option(4, args.pdl))

    n_vox = DX * DY * DZ
    dims = (DX, DY, DZ * world)
    slab = (DZ * rank, DZ * (rank + 1))
    # N_PAIRS (environment, robot) map pairs and as many sensor frames used in rotation, so that a step never finds its
    # inputs in the 126 MB L2: one pair's summaries are 33.5 MB, so a full rotation touches 8 x 33.5 MB = 268 MB of
    # summaries (+ 8 frames, + the scattered voxels) before a pair comes round again. 8 x 4.43 GB = 35 GB of grids.
    P = N_PAIRS
    envs = [gv.ProbVoxelMap(dims, SIDE, slab=slab) for _ in range(P)]
    robs = [gv.BitVectorVoxelMap(dims, SIDE, slab=slab) for _ in range(P)]
    # robot: every slab holds its own copy of the arm (inserted once, untimed); global cloud = all copies
    rclouds = [np.concatenate([c for r in range(world) for c in [robot_clouds(z_shift=DZ * SIDE * r)[k]]]) for k in range(7)]
    for rob in robs:
        rob.insertMetaPointCloud(gv.MetaPointCloud(rclouds), list(range(4, 11)))
    # frames: N sub-frames (one per slab) concatenated into one global cloud; every rank reads all of it
    poses = [sensor_pose(DZ * SIDE * r) for r in range(world)]
    frames_h = [[depth_frame(201 + P * r + k) for r in range(world)] for k in range(P)]  # [rotation][slab] sensor frame
    pinned = [[torch.from_numpy(f).pin_memory() for f in fr] for fr in frames_h]
    frames_d = [[p.cuda(non_blocking=True) for p in fr] for fr in pinned]
    count_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    pose_arr = [np.ascontiguousarray(p.reshape(-1)) for p in poses]

    sizes_arr = np.full(world, N_FRAME, np.uint32)
    poses_arr = np.ascontiguousarray(np.stack(pose_arr))
    # [parity] all sub-frames concatenated in HBM (device-resident arm) and the all-gather target of the e2e arm
    cat_d = [torch.cat(fr).contiguous() for fr in frames_d]
    gathered = torch.empty((world * N_FRAME, 3), dtype=torch.float32, device="cuda")
    own_d = torch.empty((N_FRAME, 3), dtype=torch.float32, device="cuda")

    # ctypes argument objects are built once: creating them per call costs more host time than a 10 us kernel runs
    sizes_p, poses_p = sizes_arr.ctypes.data_as(C.c_void_p), poses_arr.ctypes.data_as(C.c_void_p)
    pose0_p = pose_arr[0].ctypes.data_as(C.c_void_p)
    env_h, rob_h = [e.handle for e in envs], [r.handle for r in robs]
    count_p = C.c_void_p(count_d.data_ptr())
    ptr_of = {}

    def dev_ptr(t):
        p = ptr_of.get(t.data_ptr())
        if p is None:
            p = ptr_of[t.data_ptr()] = C.c_void_p(t.data_ptr())
        return p

    # sharded: every rank integrates ITS sensor's frame; hits are routed to the rank owning the voxel inside the
    # update kernel over NVLink peer memory (gvx_update_logodds_dispatch). --fold multi: every rank reads all frames
    # and keeps its slab's points (gvx_update_logodds_multi; the e2e arm then all-gathers the frames with NCCL)
    dispatch = False
    if world > 1 and args.fold != "multi":
        from gpu_voxels_b200 import sharding
        dispatch = sharding.connect_peer_dispatch(N_FRAME)
        zb_arr = sharding.slab_bounds(DZ * world, world)
        zb_p = zb_arr.ctypes.data_as(C.c_void_p)
        own_pose_p = pose_arr[rank].ctypes.data_as(C.c_void_p)

    # device-resident frames on 2 ranks: reading both frames (12.5 us) is cheaper than the exchange (14.9 us, measured);
    # from 4 ranks on the exchange wins (22.9 us vs 15 us at 8). Host frames: the exchange always wins (no all-gather)
    dispatch_resident = dispatch and (args.fold == "dispatch" or world > 2)

    def fold_own(points, k, mem_kind):
        _lib.check(L.gvx_update_logodds_dispatch(env_h[k % P], dev_ptr(points), N_FRAME, mem_kind, own_pose_p, 72, zb_p))

    def fold_frames(points_d, k):
        """log-odds update of ALL sub-frames (each with its own pose) in one launch; this rank keeps its slab's points"""
        if world == 1:
            _lib.check(L.gvx_update_logodds(env_h[k % P], dev_ptr(points_d), N_FRAME, _lib.MEM_DEVICE, pose0_p, 72))
        else:
            _lib.check(L.gvx_update_logodds_multi(env_h[k % P], dev_ptr(points_d), _lib.MEM_DEVICE, sizes_p, world,
                                                  poses_p, 72))

    # sharded: the count is all-reduced INSIDE the collide kernel over NVLink peer memory (CUDA IPC mailboxes); NCCL
    # stays the fallback when the IPC mapping is refused
    fused = False
    if world > 1 and not args.nccl_count:
        from gpu_voxels_b200 import sharding
        fused = sharding.connect_peer_mailboxes()

    def collide_global(k, host_count=None):
        if fused:
            _lib.check(L.gvx_collide_map_map_allreduce(env_h[k % P], rob_h[k % P], THR, 1, host_count, count_p))
            return
        _lib.check(L.gvx_collide_map_map_async(env_h[k % P], rob_h[k % P], THR, None, 1, count_p))
        if world > 1:
            dist.all_reduce(count_d)

    def step_device(k):
        """inputs resident in HBM; result stays on the device (all-reduced when sharded)"""
        if dispatch_resident:
            fold_own(frames_d[k % P][rank], k, _lib.MEM_DEVICE)
        else:
            fold_frames(cat_d[k % P], k)
        collide_global(k)

    def step_e2e(k):
        """through the C ABI with HOST buffers: H2D of the frame inside, count read back to the host. Sharded: every
        rank uploads ITS sensor's frame from pinned host memory and the frames are all-gathered over NVLink (each
        rank needs the whole cloud to find its slab's points), instead of N x the PCIe traffic per rank."""
        if world > 1:
            if dispatch:
                fold_own(pinned[k % P][rank], k, _lib.MEM_HOST)
            else:
                own_d.copy_(pinned[k % P][rank], non_blocking=True)
                dist.all_gather_into_tensor(gathered, own_d)
                fold_frames(gathered, k)
            if fused:
                cnt = C.c_uint64(0)
                collide_global(k, C.byref(cnt))
                return cnt.value
            collide_global(k)
            return int(count_d.item())
        f = pinned[k % P][0]
        _lib.check(L.gvx_update_logodds(envs[k % P].handle, C.c_void_p(f.data_ptr()), N_FRAME, _lib.MEM_HOST,
                                        pose_arr[0].ctypes.data_as(C.c_void_p), 72))
        cnt = C.c_uint64(0)
        _lib.check(L.gvx_collide_map_map(envs[k % P].handle, robs[k % P].handle, THR, None, 1, C.byref(cnt)))
        return cnt.value

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # The same step software-pipelined by the caller, two steps in flight: frames are handed over as GVX_MEM_HOST_PINNED
    # (uploaded on libgvx's copy stream into alternating staging slots, so the upload of step k+1 overlaps the kernels of
    # step k), the collide is the asynchronous entry point, its count is copied to pinned host memory on the stream and
    # read by the host one step later. Every step still uploads its frame from host memory and the host reads every
    # step's count inside the timed region.
    count2_d = torch.zeros(2, dtype=torch.int64, device="cuda")
    count2_h = torch.zeros(2, dtype=torch.int64).pin_memory()
    count2_p = [C.c_void_p(count2_d.data_ptr() + 8 * j) for j in range(2)]
    count2_ev = [torch.cuda.Event(), torch.cuda.Event()]
    pipelined_ok = world == 1 or dispatch

    def submit_e2e(k):
        j = k & 1
        if world > 1:
            _lib.check(L.gvx_update_logodds_dispatch(env_h[k % P], dev_ptr(pinned[k % P][rank]), N_FRAME,
                                                     _lib.MEM_HOST_PINNED, own_pose_p, 72, zb_p))
            if fused:
                _lib.check(L.gvx_collide_map_map_allreduce(env_h[k % P], rob_h[k % P], THR, 1, None, count2_p[j]))
            else:
                _lib.check(L.gvx_collide_map_map_async(env_h[k % P], rob_h[k % P], THR, None, 1, count2_p[j]))
                dist.all_reduce(count2_d[j:j + 1])
        else:
            _lib.check(L.gvx_update_logodds(env_h[k % P], dev_ptr(pinned[k % P][0]), N_FRAME, _lib.MEM_HOST_PINNED,
                                            pose0_p, 72))
            _lib.check(L.gvx_collide_map_map_async(env_h[k % P], rob_h[k % P], THR, None, 1, count2_p[j]))
        count2_h[j:j + 1].copy_(count2_d[j:j + 1], non_blocking=True)
        count2_ev[j].record(stream)

    def collect_e2e(k):
        count2_ev[k & 1].synchronize()
        return int(count2_h[k & 1])

    def run_e2e_pipelined(steps):
        """-> (ms for `steps` steps, max over ranks; counts of all steps)"""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        counts = []
        barrier()
        ev0.record(stream)
        submit_e2e(0)
        for k in range(1, steps):
            submit_e2e(k)
            counts.append(collect_e2e(k - 1))
        counts.append(collect_e2e(steps - 1))
        ev1.record(stream)
        barrier()
        ms = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()), counts


    def timed(fn, steps, per_kernel=False):
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)
        for k in range(steps):
            fn(k)
        ev1.record(stream)
        barrier()
        ms = torch.tensor([ev0.elapsed_time(ev1)], device="cuda")
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for k in range(max(args.warmup, 3)):
        step_device(k)
        step_e2e(k)
    # dominant kernel alone (collide), CUDA events on the launching stream, same maps, inputs >> L2
    kev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    kev_stream = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
                  for _ in range(min(args.steps, 50))]
    launches0 = L.gvx_launch_count()
    with ClockSampler(local_rank) as clk:
        ms_total = timed(step_device, args.steps)
        launches = L.gvx_launch_count() - launches0
        ms_e2e = timed(step_e2e, args.steps)
        ms_pipe, counts_pipe = (None, None)
        if pipelined_ok:
            run_e2e_pipelined(max(args.warmup, 3))
            ms_pipe, counts_pipe = run_e2e_pipelined(args.steps)
        barrier()
        # the collide kernel alone: (1) the default path (robot map's occupancy summary + streamed ProbVoxelMap) and
        # (2) the voxel-streaming kernel (GVX_OPT_USE_SUMMARY = 0: both grids read in full, the reference's pattern)
        def collide_alone(k):
            _lib.check(L.gvx_collide_map_map_async(env_h[k % P], rob_h[k % P], THR, None, 1, count_p))
        # (1a) one event pair per launch, L2 flushed before each: includes the event pair's own ~2-3 us
        l2_flush = torch.zeros(256 << 20, dtype=torch.uint8, device="cuda")
        for k, (e0, e1) in enumerate(kev):
            l2_flush.add_(1)  # 2 x 256 MB of traffic: nothing of the 33.5 MB of summaries survives in the 126 MB L2
            e0.record(stream)
            collide_alone(k)
            e1.record(stream)
        torch.cuda.synchronize()
        del l2_flush
        # (1b) `steps` launches back to back over the rotating pairs (268 MB of summaries per rotation > L2), one event
        # pair around all of them: the average launch duration without per-launch event overhead. A ~1 ms spin kernel
        # in front lets the host queue the launches while the GPU is busy, so the figure is the device's
        # launch-to-launch time and not the Python call rate.
        evb0, evb1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(2_000_000)
        evb0.record(stream)
        for k in range(args.steps):
            collide_alone(k)
        evb1.record(stream)
        torch.cuda.synchronize()
        k_b2b = evb0.elapsed_time(evb1) / args.steps
        _lib.check(L.gvx_set_option(1, 0))
        for k in range(3):
            collide_alone(k)
        for k, (e0, e1) in enumerate(kev_stream):
            e0.record(stream)
            collide_alone(k)
            e1.record(stream)
        torch.cuda.synchronize()
        _lib.check(L.gvx_set_option(1, 1))
        # insertion alone: points/s of the log-odds frame update (atomic-bound; reported beside the headline)
        evi0, evi1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        evi0.record(stream)
        for k in range(args.steps):
            f = frames_d[k % P][rank]
            _lib.check(L.gvx_update_logodds(envs[k % P].handle, C.c_void_p(f.data_ptr()), N_FRAME, _lib.MEM_DEVICE,
                                            pose_arr[rank].ctypes.data_as(C.c_void_p), 72))
        evi1.record(stream)
        torch.cuda.synchronize()
    if args.diag:
        # where a sharded step spends its time (stderr only): the fold alone, the collide alone without / with the
        # in-kernel all-reduce, and an empty-kernel-free measure of the host's launch rate (the same calls, GPU idle)
        def t_loop(fn):
            return timed(fn, args.steps) / args.steps * 1e3
        d = {"fold": "dispatch" if dispatch_resident else "multi",
             "fold_us": t_loop((lambda k: fold_own(frames_d[k % P][rank], k, _lib.MEM_DEVICE)) if dispatch_resident else
                               (lambda k: fold_frames(cat_d[k % P], k))),
             "collide_local_us": t_loop(lambda k: _lib.check(L.gvx_collide_map_map_async(
                 envs[k % P].handle, robs[k % P].handle, THR, None, 1, C.c_void_p(count_d.data_ptr())))),
             "collide_global_us": t_loop(lambda k: collide_global(k)),
             "step_us": t_loop(step_device)}
        t0 = time.perf_counter()
        for k in range(args.steps):
            step_device(k)
        d["host_issue_us_per_step"] = (time.perf_counter() - t0) / args.steps * 1e6
        barrier()
        print("diag rank %d: %s" % (rank, json.dumps(d)), file=sys.stderr)
    k_ms = sorted(e0.elapsed_time(e1) for e0, e1 in kev)
    k_avg = sum(k_ms) / len(k_ms)
    ins_ms = evi0.elapsed_time(evi1) / args.steps
    last_count = step_e2e(0)

    if rank == 0:
        peak, peak_src = measured_peak_gbs()
        ks_ms = sorted(e0.elapsed_time(e1) for e0, e1 in kev_stream)
        ks_avg = sum(ks_ms) / len(ks_ms)
        achieved = BYTES_PER_VOXEL_SUMMARY * n_vox / (k_b2b * 1e-3) / 1e9
        achieved_stream = BYTES_PER_VOXEL * n_vox / (ks_avg * 1e-3) / 1e9

        def _traffic(name):
            try:
                return json.load(open(os.path.join(ROOT, "profiles", name))).get("dram_bytes_per_launch")
            except Exception:
                return None
        traffic, traffic_stream = _traffic("collide_summary_traffic.json"), _traffic("collide_traffic.json")
        cpu = cpu_arm(2, 1) if world == 1 and not args.no_cpu_baseline else None
        line = {
            "metric": "collideWith voxels/s", "value": n_vox * world * args.steps / (ms_total * 1e-3), "unit": "voxels/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(world),
            "e2e": {"value": n_vox * world * args.steps / ((ms_pipe or ms_e2e) * 1e-3), "unit": "voxels/s",
                    "h2d_bytes_per_step": N_FRAME * 12 * world + 48 * world, "d2h_bytes_per_step": 8 * world,
                    "mode": "caller keeps two steps in flight (GVX_MEM_HOST_PINNED frames + gvx_collide_map_map_async, "
                            "count read back one step later)" if ms_pipe else "one synchronous call after the other",
                    "note": "per step, summed over ranks: every rank uploads its own sensor frame (pinned -> HBM) through the "
                            "C ABI; " +
                            ("hits in other ranks' slabs are routed there inside the update kernel over NVLink peer memory, "
                             if dispatch else "frames are all-gathered over NVLink (NCCL), ") +
                            "the count is all-reduced and read back to the host by every rank, every step, inside the timed "
                            "region. The frame upload (3.7 MB over PCIe) is the longest stage of a step; with two steps in "
                            "flight it overlaps the previous step's kernels",
                    "ms_per_step": (ms_pipe or ms_e2e) / args.steps,
                    "synchronous_calls": {"value": n_vox * world * args.steps / (ms_e2e * 1e-3),
                                          "ms_per_step": ms_e2e / args.steps,
                                          "how": "gvx_update_logodds(host frame) then gvx_collide_map_map returning the "
                                                 "count, one after the other, nothing in flight across steps"}},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": "k_collide_summary2<Prob,Bit,mark>", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": BYTES_PER_VOXEL_SUMMARY * n_vox,
                         "algorithmic_bytes_per_voxel": BYTES_PER_VOXEL_SUMMARY, "kernel_ms_avg": k_b2b,
                         "kernel_ms_avg_how": "%d launches back to back over %d rotating map pairs (268 MB of summaries per "
                                              "rotation > L2), queued behind a 1 ms spin kernel, one CUDA event pair around all of them; "
                                              "launched with programmatic dependent launch unless --pdl 0, so the ramp of a "
                                              "launch overlaps the tail of the one before it, as in the timed step" % (args.steps, N_PAIRS),
                         "pdl": args.pdl if args.pdl is not None else 1,
                         "kernel_ms_single_launch_l2_flushed": k_avg, "kernel_ms_single_launch_min": k_ms[0],
                         "frac_of_8000_nominal": achieved / 8000.0,
                         "note": "default path: one summary bit per voxel and side (2 x 16.8 MB for 512^3) + the Prob bytes of "
                                 "candidate groups (DESIGN.md 3.1): a 33.5 MB problem, launch-latency territory; the "
                                 "33 B/voxel figure of SURVEY 8d belongs to roofline_streaming"},
            "roofline_streaming": {"bound": "hbm", "kernel": "k_collide_voxelwise<Prob,Bit,mark>", "achieved": achieved_stream,
                                   "peak": peak, "unit": "GB/s", "frac": achieved_stream / peak, "traffic": traffic_stream,
                                   "algorithmic_bytes_per_launch": BYTES_PER_VOXEL * n_vox,
                                   "algorithmic_bytes_per_voxel": BYTES_PER_VOXEL, "kernel_ms_avg": ks_avg,
                                   "kernel_ms_min": ks_ms[0], "frac_of_8000_nominal": achieved_stream / 8000.0,
                                   "voxels_per_s": n_vox / (ks_avg * 1e-3),
                                   "note": "GVX_OPT_USE_SUMMARY=0: both grids streamed in full like the reference"},
            "insert": {"metric": "insertSensorData (log-odds) points/s", "value": N_FRAME / (ins_ms * 1e-3),
                       "unit": "points/s", "ms_per_frame": ins_ms, "points": N_FRAME},
            "clocks": clk.summary(), "collisions_last_step": int(last_count),
            "count_allreduce": ("in-kernel over NVLink peer memory" if fused else "nccl") if world > 1 else "none",
            "frame_exchange": ({"value": "in-kernel dispatch over NVLink peer memory" if dispatch_resident else
                                         "every rank folds all frames",
                                "e2e": "in-kernel dispatch over NVLink peer memory" if dispatch else
                                       "NCCL all-gather of the frames, every rank folds all frames"} if world > 1 else "none"),
        }
        if cpu is not None:
            line["cpu_baseline"] = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--diag", action="store_true", help="print a per-phase timing breakdown to stderr")
    ap.add_argument("--nccl-count", action="store_true", help="all-reduce the count with NCCL instead of in-kernel")
    ap.add_argument("--pdl", type=int, default=None, choices=[0, 1],
                    help="GVX_OPT_PDL: programmatic dependent launch for the per-step kernels (library default when absent)")
    ap.add_argument("--fold", default="auto", choices=["auto", "dispatch", "multi"],
                    help="sharded runs: route each rank's frame to the owning ranks in-kernel (dispatch), fold all frames on "
                         "every rank (multi), or auto: dispatch for host frames and, from 4 ranks on, for resident frames")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
