#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (BASELINE.json: collideWith voxels/s, % HBM peak; insert points/s;
at 1/2/4/8 B200).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config all|c4|c5|c2]

Headline workload at EVERY N (so that the N = 1/2/4/8 lines are one strong-scaling series): BASELINE.json configs[3],
"C4" -- a 2048^3 ProbVoxelMap (8.6 G voxels) cut into N z-slabs, one per GPU. One step = 8 clouds of 1 M points folded
into the environment map with the log-odds update, then collideWith (Prob x Prob, threshold 0.5, collisions marked)
against a 2048^3 robot-volume map, the count all-reduced over the ranks. `value` = voxels collided per second over the
whole job with every voxel byte READ (GVX_OPT_USE_SUMMARY = 0: SURVEY 8d's 2 B per voxel, 17.18 GB per step); the
library's default path (1-bit summaries first) is timed beside it under "default_path".
Extra objects in the same line: "c5" (configs[4]: 10 M-entry BitVectorVoxelList x 1024^3 BitVectorVoxelMap, z-sharded,
count + 256 per-meaning counts all-reduced) at every N, and at N = 1 "c2" (configs[1]: 512^3 Prob x BitVector + one
640x480 depth frame per step) with the reference's own CUDA kernels rebuilt for sm_100a timed beside it ("ref_cuda").
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

SIDE = 0.01
THR = 0.5
REPS = 15  # the K-step timed region is repeated REPS times; the line reports the median (p10 / p90 beside it)

# ---- C4 (configs[3]) --------------------------------------------------------------------------------------------------
C4_DIM = 2048
C4_CLOUDS, C4_POINTS = 8, 1_000_000
C4_SETS = 2                       # cloud sets used in rotation (2 x 96 MB of points)
C4_POSE = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)  # sensor in voxel (0,0,0) of the map
C4_BYTES_PER_VOXEL = 2            # Prob x Prob, SURVEY 8d
C4_SAMPLE = (960, 1088)           # z-slab of the CPU sample (128 of 2048 layers = 537 M voxels, 2 x 0.5 GB of maps)


def c4_cloud(k, cloud_set=0):
    """cloud k: 700 000 points uniform in [0, 20.48)^3 + 300 000 in 50 Gaussian clusters (sigma 5 cm); SURVEY 8d C4"""
    rng = np.random.Generator(np.random.PCG64(401 + k + 10 * cloud_set))
    u = rng.uniform(0.0, 20.48, size=(700_000, 3))
    centres = rng.uniform(0.0, 20.48, size=(50, 3))
    cl = centres[rng.integers(0, 50, 300_000)] + rng.normal(0.0, 0.05, size=(300_000, 3))
    return np.ascontiguousarray(np.concatenate([u, cl]), dtype=np.float32)


def c4_robot_points():
    """robot volume: 2 M points, half of them shared with cloud 0 (so there ARE collisions), seed 499"""
    rng = np.random.Generator(np.random.PCG64(499))
    return np.ascontiguousarray(np.concatenate([c4_cloud(0)[:1_000_000], rng.uniform(0.0, 20.48, size=(1_000_000, 3))]),
                                dtype=np.float32)


def z_layer(points, pose=C4_POSE):
    """world z voxel layer of sensor-frame points under an axis-aligned pose (fp32, the device's arithmetic for
    z >= 0; used only to PRE-SELECT the points of a z-slab sample, the kernels / the oracle decide for themselves)"""
    wz = points[:, 2] + np.float32(pose[2, 3])
    with np.errstate(invalid="ignore"):
        return np.floor(wz / np.float32(SIDE))


# ---- C5 (configs[4]) --------------------------------------------------------------------------------------------------
C5_DIM = 1024
C5_ENTRIES = 10_000_000
C5_BYTES_PER_ENTRY = 68           # 4 B id + 32 B list voxel + 32 B map voxel, SURVEY 8d

# ---- C2 (configs[1]) --------------------------------------------------------------------------------------------------
DX = DY = DZ = 512
FRAME_W, FRAME_H = 640, 480
N_FRAME = FRAME_W * FRAME_H
C2_BYTES_PER_VOXEL = 33           # Prob (1 B) x BitVector (32 B), SURVEY 8d
C2_PAIRS = 4                      # rotating (environment, robot) pairs: 4 x 4.43 GB of grids, 4 x 33.5 MB of summaries


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


def sensor_pose(scale=1.0):
    """sensor at (0.3, 2.56, 2.2) * scale looking along +x, image-down = world -z; row-major [R|t]"""
    return np.array([[0, 0, 1, 0.3 * scale], [-1, 0, 0, 2.56 * scale], [0, -1, 0, 2.2 * scale]], np.float32)


def robot_clouds(seed=301, scale=1.0):
    """7 link clouds of 50 000 surface points each (capsules, r 4-8 cm, l 10-40 cm): an arm that starts in front of
    the sensor and reaches +x through the face of the first box of the depth scene (so there ARE collisions);
    meanings 4..10 (SURVEY 8d "C3" arm, inserted once for C2). scale 0.5 places it in a 256^3 grid."""
    rng = np.random.Generator(np.random.PCG64(seed))
    clouds, base = [], np.array([1.25, 3.0, 2.0])
    for k in range(7):
        r, l = rng.uniform(0.04, 0.08), rng.uniform(0.10, 0.40)
        t = rng.uniform(0, l, 50000); a = rng.uniform(0, 2 * np.pi, 50000)
        d = np.array([1.0, rng.uniform(-0.25, 0.25), rng.uniform(-0.25, 0.25)]); d /= np.linalg.norm(d)
        e1 = np.cross(d, [0, 0, 1.0]); e1 /= np.linalg.norm(e1); e2 = np.cross(d, e1)
        p = base + np.outer(t, d) + r * (np.outer(np.cos(a), e1) + np.outer(np.sin(a), e2))
        clouds.append(np.ascontiguousarray(p * scale, dtype=np.float32))
        base = base + l * d * 0.9
    return clouds


# ---- clocks sampling during the timed region (B200_PROFILING.md) -----------------------------------------------------
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


def traffic_of(name):
    """DRAM bytes per launch of a kernel from the committed ncu --set full digest under profiles/ (None if absent)"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", name))).get("dram_bytes_per_launch")
    except Exception:
        return None


def stats(xs):
    s = sorted(xs)
    q = lambda f: s[min(len(s) - 1, int(f * len(s)))]  # noqa: E731
    return {"median": s[len(s) // 2], "p10": q(0.1), "p90": q(0.9), "min": s[0], "n": len(s)}


# ---- the GPU side: one process per GPU ------------------------------------------------------------------------------------
class Ctx:
    """rank / stream / library handle of this process plus the timing helpers every config uses"""

    def __init__(self, args):
        import torch
        import torch.distributed as dist
        from gpu_voxels_b200 import _lib
        self.torch, self.dist, self._lib, self.args = torch, dist, _lib, args
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != args.gpus and self.world > 1:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (args.gpus, self.world))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            # NCCL announces its version on STDOUT when the first communicator comes up; the contract is ONE JSON line
            # on stdout, so fd 1 points at stderr until the communicator exists.
            sys.stdout.flush()
            saved_fd = os.dup(1)
            os.dup2(2, 1)
            try:
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
                warm = torch.zeros(1, device="cuda")
                dist.all_reduce(warm)
                torch.cuda.synchronize()
            finally:
                sys.stdout.flush()
                os.dup2(saved_fd, 1)
                os.close(saved_fd)
        self.L = _lib.init(self.local)
        # one explicit (non-default) stream shared by libgvx, torch's events and NCCL: CUDA events time libgvx's kernels
        self.stream = torch.cuda.Stream()
        torch.cuda.set_stream(self.stream)
        assert self.stream.cuda_stream != 0
        _lib.set_stream(self.stream.cuda_stream)
        if args.pdl is not None:
            _lib.check(self.L.gvx_set_option(4, args.pdl))
        self.fused = self.dispatch = False
        self._ptrs = {}

    def check(self, rc):
        self._lib.check(rc)

    def ptr(self, t):
        """ctypes pointer objects are built once per tensor: creating them per call costs more than a short kernel"""
        p = self._ptrs.get(t.data_ptr())
        if p is None:
            p = self._ptrs[t.data_ptr()] = C.c_void_p(t.data_ptr())
        return p

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, ms):
        if self.world == 1:
            return ms
        t = self.torch.tensor([ms], device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def timed(self, fn, steps):
        """EXACTLY `steps` calls of fn(k) between two CUDA events on the launching stream, barrier + synchronize on both
        sides, max over ranks -> ms"""
        torch = self.torch
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.barrier()
        ev0.record(self.stream)
        for k in range(steps):
            fn(k)
        ev1.record(self.stream)
        self.barrier()
        return self.max_over_ranks(ev0.elapsed_time(ev1))

    def timed_reps(self, fn, steps, reps=REPS):
        """the `steps`-long region repeated `reps` times -> stats of ms PER STEP"""
        return stats([self.timed(fn, steps) / steps for _ in range(reps)])

    def kernel_times(self, fn, n, flush=None):
        """one event pair per call (device time of single launches incl. the event pair's own ~2 us) -> sorted ms"""
        torch = self.torch
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        for k, (e0, e1) in enumerate(evs):
            if flush is not None:
                flush.add_(1)
            e0.record(self.stream)
            fn(k)
            e1.record(self.stream)
        torch.cuda.synchronize()
        return sorted(e0.elapsed_time(e1) for e0, e1 in evs)

    def connect_peers(self, dispatch_capacity):
        """CUDA-IPC mailboxes (count all-reduce inside the collide kernel) and dispatch inboxes (hits routed to the owning
        rank inside the update kernel) over NVLink peer memory; NCCL stays the fallback when the mapping is refused"""
        if self.world == 1:
            return
        from gpu_voxels_b200 import sharding
        if not self.args.nccl_count:
            self.fused = sharding.connect_peer_mailboxes()
        if self.args.fold != "multi":
            self.dispatch = sharding.connect_peer_dispatch(int(dispatch_capacity))


# ---- C4: 2048^3 ProbVoxelMap z-slab sharded, 8 x 1 M-point clouds per step, Prob x Prob collide ----------------------
def run_c4(ctx):
    import gpu_voxels_b200 as gv
    from gpu_voxels_b200 import sharding
    torch, dist, L, _lib, args = ctx.torch, ctx.dist, ctx.L, ctx._lib, ctx.args
    rank, world, K = ctx.rank, ctx.world, args.steps
    MD, MH, MP = _lib.MEM_DEVICE, _lib.MEM_HOST, _lib.MEM_HOST_PINNED
    dims = (C4_DIM, C4_DIM, C4_DIM)
    z0, z1 = sharding.slab_for_rank(C4_DIM, rank, world)
    n_total, n_loc = C4_DIM ** 3, C4_DIM * C4_DIM * (z1 - z0)
    env = gv.ProbVoxelMap(dims, SIDE, slab=(z0, z1))
    rob = gv.ProbVoxelMap(dims, SIDE, slab=(z0, z1))
    rob.insertPointCloud(c4_robot_points(), gv.eBVM_OCCUPIED, check_outside=False)  # every rank keeps its slab's points

    sets_h = [[c4_cloud(k, s) for k in range(C4_CLOUDS)] for s in range(C4_SETS)]
    own = list(range(rank * C4_CLOUDS // world, (rank + 1) * C4_CLOUDS // world))  # the clouds this rank's sensors deliver
    n_all, n_own = C4_CLOUDS * C4_POINTS, len(own) * C4_POINTS
    all_h = [np.concatenate(s) for s in sets_h]
    own_h = [np.concatenate([s[k] for k in own]) for s in sets_h]
    all_d = [torch.from_numpy(a).cuda() for a in all_h]          # resident arm: the whole frame in every rank's HBM
    own_d = [torch.from_numpy(a).cuda() for a in own_h]          # resident arm with dispatch: this rank's clouds only
    own_pin = [torch.from_numpy(a).pin_memory() for a in own_h]  # e2e arm: this rank's clouds in pinned host memory
    all_pin = [torch.from_numpy(a).pin_memory() for a in all_h] if world == 1 else None
    gathered = torch.empty((n_all, 3), dtype=torch.float32, device="cuda") if world > 1 else None
    own_stage = torch.empty((n_own, 3), dtype=torch.float32, device="cuda") if world > 1 else None
    pose = np.ascontiguousarray(C4_POSE.reshape(-1))
    pose_p = pose.ctypes.data_as(C.c_void_p)
    sizes = np.full(C4_CLOUDS, C4_POINTS, np.uint32)
    poses = np.ascontiguousarray(np.tile(pose, (C4_CLOUDS, 1)))
    sizes_p, poses_p = sizes.ctypes.data_as(C.c_void_p), poses.ctypes.data_as(C.c_void_p)
    zb = sharding.slab_bounds(C4_DIM, world)
    zb_p = zb.ctypes.data_as(C.c_void_p)
    count_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    count_p = ctx.ptr(count_d)
    eh, rh = env.handle, rob.handle
    ctx.connect_peers(n_own)

    def fold_multi(points_d, kind=MD):
        """every rank reads the whole 8-cloud frame and keeps its slab's points (one launch, grid.y = cloud)"""
        ctx.check(L.gvx_update_logodds_multi(eh, ctx.ptr(points_d), kind, sizes_p, C4_CLOUDS, poses_p, 72))

    def fold_dispatch(points, kind=MD):
        """every rank voxelises only ITS clouds; hits are routed to the owning rank inside the kernel over NVLink"""
        ctx.check(L.gvx_update_logodds_dispatch(eh, ctx.ptr(points), n_own, kind, pose_p, 72, zb_p))

    def collide(to_host=False, out_p=count_p):
        """collideWith of the two slabs, the count summed over the ranks; to_host: synchronous, returns the count"""
        cnt = C.c_uint64(0)
        if world == 1:
            if to_host:
                ctx.check(L.gvx_collide_map_map(eh, rh, THR, None, 1, C.byref(cnt)))
                return cnt.value
            ctx.check(L.gvx_collide_map_map_async(eh, rh, THR, None, 1, out_p))
        elif ctx.fused:  # count all-reduced inside the collide kernel over NVLink peer memory
            ctx.check(L.gvx_collide_map_map_allreduce(eh, rh, THR, 1, C.byref(cnt) if to_host else None, out_p))
            return cnt.value if to_host else None
        else:
            ctx.check(L.gvx_collide_map_map_async(eh, rh, THR, None, 1, count_p))
            dist.all_reduce(count_d)
            if to_host:
                return int(count_d.item())
            if out_p is not count_p:
                raise RuntimeError("NCCL fallback reduces count_d only")
        return None

    def collide_local(k):
        ctx.check(L.gvx_collide_map_map_async(eh, rh, THR, None, 1, count_p))

    def make_step(fold):
        if fold == "dispatch":
            return lambda k: (fold_dispatch(own_d[k % C4_SETS]), collide())
        return lambda k: (fold_multi(all_d[k % C4_SETS]), collide())

    # which fold for the resident arm: measured, not argued (both leave byte-identical maps; multi_gpu_check proves it)
    fold, fold_probe = "multi", None
    ctx.check(L.gvx_set_option(1, 0))  # GVX_OPT_USE_SUMMARY = 0: every voxel byte is read
    for k in range(max(args.warmup, 3)):
        make_step("multi")(k)
    if ctx.dispatch:
        for k in range(3):
            make_step("dispatch")(k)
        fold_probe = {"multi_ms": ctx.timed(lambda k: fold_multi(all_d[k % C4_SETS]), 5) / 5,
                      "dispatch_ms": ctx.timed(lambda k: fold_dispatch(own_d[k % C4_SETS]), 5) / 5}
        want = args.fold if args.fold != "auto" else ("dispatch" if fold_probe["dispatch_ms"] < fold_probe["multi_ms"] else "multi")
        fold = want
    step = make_step(fold)

    res = {"fold": fold, "fold_probe_ms": fold_probe}
    launches0 = L.gvx_launch_count()
    with ClockSampler(ctx.local) as clk:
        res["step"] = ctx.timed_reps(step, K)                       # headline: voxel-streaming path
        res["launches_per_region"] = int((L.gvx_launch_count() - launches0) // REPS)
        # the dominant kernel alone: K launches back to back between one event pair (2 x n_loc bytes per launch >> L2)
        res["collide_kernel"] = ctx.timed_reps(collide_local, K, reps=5)
        res["fold_alone"] = ctx.timed_reps((lambda k: fold_dispatch(own_d[k % C4_SETS])) if fold == "dispatch" else
                                           (lambda k: fold_multi(all_d[k % C4_SETS])), K, reps=5)
    res["clocks"] = clk.summary()
    collide()
    torch.cuda.synchronize()
    res["count_streaming"] = int(count_d.item())

    # ---- end to end through the C ABI with HOST buffers: every rank uploads the clouds of ITS sensors -------------------
    count2_d = torch.zeros(2, dtype=torch.int64, device="cuda")
    count2_h = torch.zeros(2, dtype=torch.int64).pin_memory()
    count2_p = [C.c_void_p(count2_d.data_ptr() + 8 * j) for j in range(2)]
    count2_ev = [torch.cuda.Event(), torch.cuda.Event()]

    def fold_host(k, kind):
        s = k % C4_SETS
        if world == 1:
            fold_multi(all_pin[s], kind)
        elif ctx.dispatch:
            fold_dispatch(own_pin[s], kind)
        else:  # no peer mapping: upload the own clouds, all-gather the frame with NCCL, fold all
            own_stage.copy_(own_pin[s], non_blocking=True)
            dist.all_gather_into_tensor(gathered, own_stage)
            fold_multi(gathered)

    def step_sync(k):
        """one synchronous call after the other: frame up, count back, nothing in flight across steps"""
        fold_host(k, MH)
        return collide(to_host=True)

    def submit(k):
        j = k & 1
        fold_host(k, MP)
        if world > 1 and not ctx.fused:
            ctx.check(L.gvx_collide_map_map_async(eh, rh, THR, None, 1, count2_p[j]))
            dist.all_reduce(count2_d[j:j + 1])
        else:
            collide(False, count2_p[j])
        count2_h[j:j + 1].copy_(count2_d[j:j + 1], non_blocking=True)
        count2_ev[j].record(ctx.stream)

    def collect(k):
        count2_ev[k & 1].synchronize()
        return int(count2_h[k & 1])

    def region_pipelined(steps):
        """the caller keeps two steps in flight: upload of step k+1 (copy stream) overlaps the kernels of step k; every
        step's count is read by the host inside the region, one step late"""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.barrier()
        ev0.record(ctx.stream)
        submit(0)
        for k in range(1, steps):
            submit(k)
            collect(k - 1)
        last = collect(steps - 1)
        ev1.record(ctx.stream)
        ctx.barrier()
        return ctx.max_over_ranks(ev0.elapsed_time(ev1)), last

    for k in range(3):
        step_sync(k)
    region_pipelined(3)
    res["e2e_sync"] = ctx.timed_reps(step_sync, K, reps=5)
    pipe = [region_pipelined(K) for _ in range(5)]
    res["e2e_pipelined"] = stats([p[0] / K for p in pipe])
    res["count_e2e"] = pipe[-1][1]
    res["h2d_bytes_per_step"] = n_all * 12 + 48 * world
    res["d2h_bytes_per_step"] = 8 * world

    # ---- the library's default path: 1-bit "known" summaries first, Prob bytes only for candidate groups -------------
    ctx.check(L.gvx_set_option(1, 1))
    for k in range(3):
        step(k)
    res["default_step"] = ctx.timed_reps(step, K)
    res["default_collide_kernel"] = ctx.timed_reps(collide_local, K, reps=5)
    collide()
    torch.cuda.synchronize()
    res["count_default"] = int(count_d.item())
    res.update({"n_total": n_total, "n_loc": n_loc, "points_per_step": n_all, "slab": [z0, z1]})
    env.destroy(); rob.destroy()
    del all_d, own_d, gathered, own_stage
    torch.cuda.empty_cache()
    return res


# ---- C5: 10 M-entry BitVectorVoxelList x 1024^3 BitVectorVoxelMap (32 GiB, z-sharded), count + per-meaning counts -----
def c5_inputs():
    """-> ids uint32[10 M] sorted unique (seed 501), voxels uint8[10 M, 32] with 1-3 meanings from 4..35 (seed 502),
    boxes int[200, 3] = origins of 47^3-voxel occupied boxes (2 % of the grid, seed 503)"""
    rng = np.random.Generator(np.random.PCG64(501))
    u = np.sort(rng.integers(0, C5_DIM ** 3, int(C5_ENTRIES * 1.01), dtype=np.int64))
    u = u[np.concatenate([[True], u[1:] != u[:-1]])]  # unique (np.unique on 10 M int64 takes 9 s in numpy 2.3)
    ids = np.sort(u[rng.permutation(len(u))[:C5_ENTRIES]]).astype(np.uint32)
    rng = np.random.Generator(np.random.PCG64(502))
    words = np.zeros((C5_ENTRIES, 8), np.uint32)
    n_mg = rng.integers(1, 4, C5_ENTRIES)
    for j in range(3):
        mg = rng.integers(4, 36, C5_ENTRIES)
        sel = n_mg > j
        rows = np.flatnonzero(sel)  # one (row, word) pair per row and pass: plain fancy indexing is exact
        words[rows, mg[sel] >> 5] |= np.uint32(1) << (mg[sel] & 31).astype(np.uint32)
    boxes = np.random.Generator(np.random.PCG64(503)).integers(0, C5_DIM - 47, size=(200, 3))
    return ids, words.view(np.uint8).reshape(C5_ENTRIES, 32), boxes


def c5_expected(ids, voxels, boxes):
    """numpy restatement on the ids (GV/voxellist/kernels/VoxelListOperations.hpp:126-177 + BitVoxelList.hpp:300-344):
    an entry collides iff its map voxel is occupied; every set bit of a colliding entry counts once"""
    occ = np.zeros((C5_DIM, C5_DIM, C5_DIM), bool)  # [z, y, x]
    for x, y, z in boxes:
        occ[z:z + 47, y:y + 47, x:x + 47] = True
    hit = occ.reshape(-1)[ids]
    bits = np.unpackbits(voxels[hit], axis=1, bitorder="little")  # (n_hit, 256)
    per = bits.sum(0).astype(np.uint64)
    return int(hit.sum()), per, hit


def run_c5(ctx):
    import gpu_voxels_b200 as gv
    from gpu_voxels_b200 import sharding
    torch, dist, L, _lib, args = ctx.torch, ctx.dist, ctx.L, ctx._lib, ctx.args
    rank, world, K = ctx.rank, ctx.world, args.steps
    dims = (C5_DIM, C5_DIM, C5_DIM)
    z0, z1 = sharding.slab_for_rank(C5_DIM, rank, world)
    ids, voxels, boxes = c5_inputs()
    m = gv.BitVectorVoxelMap(dims, SIDE, slab=(z0, z1))
    r = torch.arange(47, device="cuda")
    bz, by, bx = torch.meshgrid(r, r, r, indexing="ij")
    local = torch.stack([bx, by, bz], -1).reshape(-1, 3).to(torch.float32)
    for o in boxes:  # voxel centres of a box, generated on the device; every rank keeps its slab's part
        pts = ((local + torch.tensor(o, device="cuda", dtype=torch.float32)) + 0.5) * SIDE
        ctx.check(L.gvx_insert_points(m.handle, C.c_void_p(pts.contiguous().data_ptr()), pts.shape[0], _lib.MEM_DEVICE, 1, None))
        torch.cuda.synchronize()
    lst = gv.BitVectorVoxelList(dims, SIDE)
    coords = np.stack([ids % C5_DIM, (ids // C5_DIM) % C5_DIM, ids // (C5_DIM * C5_DIM)], -1).astype(np.uint32)
    lst.upload(ids, coords, voxels)
    out_d = torch.zeros(257, dtype=torch.int64, device="cuda")
    out_p, cnt_p = ctx.ptr(out_d), C.c_void_p(out_d.data_ptr())
    lh, mh = lst.handle, m.handle
    if world > 1 and not ctx.fused and not args.nccl_count:
        ctx.connect_peers(1)

    def step_nccl(k):
        """count + 256 per-meaning counts of this rank's slab, then ONE NCCL all-reduce of the 257 words"""
        ctx.check(L.gvx_collide_list_map_per_meaning_async(lh, mh, 1.0, None, 0, out_p))
        if world > 1:
            dist.all_reduce(out_d)

    def step_fused(k):
        """the same with the 257 words all-reduced inside the kernel over NVLink peer memory (no NCCL launch)"""
        ctx.check(L.gvx_collide_list_map_per_meaning_allreduce(lh, mh, 1.0, 0, out_p))

    step = step_fused if (world > 1 and ctx.fused) else step_nccl

    def count_only(k):
        if world > 1 and ctx.fused:
            ctx.check(L.gvx_collide_list_map_allreduce(lh, mh, 1.0, 0, None, cnt_p))
        else:
            ctx.check(L.gvx_collide_list_map_async(lh, mh, 1.0, None, 0, cnt_p))
            if world > 1:
                dist.all_reduce(out_d[:1])

    def local_kernel(k):
        ctx.check(L.gvx_collide_list_map_async(lh, mh, 1.0, None, 0, cnt_p))

    res = {}
    want_total, want_per, hit = c5_expected(ids, voxels, boxes) if rank == 0 else (None, None, None)
    for use_summary, tag in ((0, "streaming"), (1, "default")):
        ctx.check(L.gvx_set_option(1, use_summary))
        for k in range(max(args.warmup, 3)):
            step(k); count_only(k)
        r_ = {"step": ctx.timed_reps(step, K), "count_only": ctx.timed_reps(count_only, K, reps=5),
              "local_kernel": ctx.timed_reps(local_kernel, K, reps=5),
              "allreduce": "none" if world == 1 else ("in-kernel over NVLink peer memory (257 words)" if step is step_fused else "nccl")}
        if step is step_fused:  # A/B: the NCCL form of the same step
            for k in range(3):
                step_nccl(k)
            r_["step_nccl"] = ctx.timed_reps(step_nccl, K, reps=5)
        step(0)
        torch.cuda.synchronize()
        got = out_d.cpu().numpy().astype(np.uint64)
        r_["total"], r_["collisions"] = int(got[0]), None
        count_only(0)
        torch.cuda.synchronize()
        r_["collisions"] = int(out_d[0].item())
        if rank == 0:
            r_["parity"] = {"collisions_equal": r_["collisions"] == want_total,
                            "per_meaning_equal": bool(np.array_equal(got[1:], want_per)),
                            "total_equal": int(got[0]) == int(want_per.sum()),
                            "checker": "numpy restatement on the ids (c5_expected)", "expected_collisions": want_total}
        res[tag] = r_
    ctx.check(L.gvx_set_option(1, 1))
    res.update({"entries": int(len(ids)), "slab": [z0, z1], "map_bytes_per_rank": int(C5_DIM * C5_DIM * (z1 - z0) * 32)})
    m.destroy(); lst.destroy()
    torch.cuda.empty_cache()
    return res


# ---- C2 (N = 1): 512^3 Prob x BitVector + one 640x480 frame per step; reference CUDA kernels at 256^3 / 128^3 ---------
def run_c2(ctx):
    import gpu_voxels_b200 as gv
    torch, L, _lib, args = ctx.torch, ctx.L, ctx._lib, ctx.args
    K, P = args.steps, C2_PAIRS
    MD, MP = _lib.MEM_DEVICE, _lib.MEM_HOST_PINNED
    dims, n_vox = (DX, DY, DZ), DX * DY * DZ
    envs = [gv.ProbVoxelMap(dims, SIDE) for _ in range(P)]
    robs = [gv.BitVectorVoxelMap(dims, SIDE) for _ in range(P)]
    for rob in robs:
        rob.insertMetaPointCloud(gv.MetaPointCloud(robot_clouds()), list(range(4, 11)))
    frames_h = [depth_frame(201 + k) for k in range(P)]
    pinned = [torch.from_numpy(f).pin_memory() for f in frames_h]
    frames_d = [p.cuda() for p in pinned]
    pose = np.ascontiguousarray(sensor_pose().reshape(-1))
    pose_p = pose.ctypes.data_as(C.c_void_p)
    count_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    count_p = ctx.ptr(count_d)
    eh, rh = [e.handle for e in envs], [r.handle for r in robs]

    def fold(k, pts=None, kind=MD):
        ctx.check(L.gvx_update_logodds(eh[k % P], ctx.ptr(pts if pts is not None else frames_d[k % P]), N_FRAME, kind, pose_p, 72))

    def collide(k):
        ctx.check(L.gvx_collide_map_map_async(eh[k % P], rh[k % P], THR, None, 1, count_p))

    def step(k):
        fold(k); collide(k)

    res = {}
    for use_summary, tag in ((0, "streaming"), (1, "default")):
        ctx.check(L.gvx_set_option(1, use_summary))
        for k in range(max(args.warmup, 3) + P):
            step(k)
        res[tag] = {"step": ctx.timed_reps(step, K), "collide_kernel": ctx.timed_reps(collide, K, reps=5)}
        collide(0)
        torch.cuda.synchronize()
        res[tag]["count"] = int(count_d.item())
    res["fold_alone"] = ctx.timed_reps(fold, K, reps=5)
    # end to end, default path, two steps in flight (pinned frame up on the copy stream, count back one step later)
    c2_d = torch.zeros(2, dtype=torch.int64, device="cuda")
    c2_h = torch.zeros(2, dtype=torch.int64).pin_memory()
    c2_p = [C.c_void_p(c2_d.data_ptr() + 8 * j) for j in range(2)]
    ev = [torch.cuda.Event(), torch.cuda.Event()]

    def submit(k):
        j = k & 1
        fold(k, pinned[k % P], MP)
        ctx.check(L.gvx_collide_map_map_async(eh[k % P], rh[k % P], THR, None, 1, c2_p[j]))
        c2_h[j:j + 1].copy_(c2_d[j:j + 1], non_blocking=True)
        ev[j].record(ctx.stream)

    def region(steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ctx.barrier()
        e0.record(ctx.stream)
        submit(0)
        for k in range(1, steps):
            submit(k)
            ev[(k - 1) & 1].synchronize()
        ev[(steps - 1) & 1].synchronize()
        e1.record(ctx.stream)
        ctx.barrier()
        return e0.elapsed_time(e1)
    region(3)
    res["e2e_pipelined"] = stats([region(K) / K for _ in range(5)])
    res["h2d_bytes_per_step"], res["d2h_bytes_per_step"] = N_FRAME * 12 + 48, 8
    res["n_vox"] = n_vox
    for m in envs + robs:
        m.destroy()
    torch.cuda.empty_cache()
    return res


def run_ref_cuda(ctx):
    """north_star: 'the reference's own CUDA kernels rebuilt for the box where they still compile' -- oracle/_ref/
    libgvref_cuda.so = GV/voxelmap/kernels/VoxelMapOperations.{h,hpp,cu} compiled UNMODIFIED for sm_100a, launched with
    the reference's launch shape, its 128 KiB D2H and its host sum (GV/voxelmap/TemplateVoxelMap.hpp:486-519). The
    reference refuses maps >= 4 GiB (TemplateVoxelMap.hpp:65-75), so configs[1] runs at 256^3 and configs[0] at 128^3;
    this library runs the SAME inputs through its synchronous calls next to it. A reported baseline, not the target."""
    so = os.path.join(ROOT, "oracle", "_ref", "libgvref_cuda.so")
    if not os.path.exists(so):
        return {"unavailable": "oracle/_ref/libgvref_cuda.so is not built"}
    import gpu_voxels_b200 as gv
    torch, L, _lib = ctx.torch, ctx.L, ctx._lib
    REF = C.CDLL(so)
    REF.gvref_collide_map_map.restype = C.c_longlong
    K = 20
    out = {}

    def wall(fn, n=K, warm=3):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(n):
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            ts.append((time.perf_counter() - t0) * 1e3)
        return stats(ts)

    # configs[1] at 256^3: one frame folded + collideWith(Prob x BitVector), synchronous calls on both sides
    d, side, n = 256, SIDE, 256 ** 3
    frame = (depth_frame(201) * np.float32(0.5)).astype(np.float32)
    pose = np.ascontiguousarray(sensor_pose(0.5).reshape(-1))
    env, rob = gv.ProbVoxelMap((d, d, d), side), gv.BitVectorVoxelMap((d, d, d), side)
    rob.insertMetaPointCloud(gv.MetaPointCloud(robot_clouds(scale=0.5)), list(range(4, 11)))
    f_d = torch.from_numpy(frame).cuda()
    cnt = C.c_uint64(0)

    def ours():
        ctx.check(L.gvx_update_logodds(env.handle, C.c_void_p(f_d.data_ptr()), N_FRAME, _lib.MEM_DEVICE,
                                       pose.ctypes.data_as(C.c_void_p), 72))
        ctx.check(L.gvx_collide_map_map(env.handle, rob.handle, THR, None, 1, C.byref(cnt)))
    r_env = torch.full((n,), -128, dtype=torch.int8, device="cuda")
    r_rob = torch.from_numpy(rob.download().reshape(-1)).cuda()
    tmp = torch.empty_like(f_d)
    r_cnt = [0]

    def ref():
        REF.gvref_insert_sensor_data(C.c_void_p(r_env.data_ptr()), d, d, d, C.c_float(side), pose.ctypes.data_as(C.c_void_p),
                                     C.c_void_p(f_d.data_ptr()), C.c_void_p(tmp.data_ptr()), N_FRAME)
        r_cnt[0] = REF.gvref_collide_map_map(C.c_void_p(r_env.data_ptr()), 4, C.c_void_p(r_rob.data_ptr()), 0, n, C.c_float(THR))
    t_ref = wall(ref)
    res = {}
    for use_summary, tag in ((0, "streaming"), (1, "default")):
        ctx.check(L.gvx_set_option(1, use_summary))
        res[tag] = wall(ours)
    # same input on both sides: this library's environment map copied into the reference's buffer
    r_env.copy_(torch.from_numpy(env.download()).cuda())
    same = REF.gvref_collide_map_map(C.c_void_p(r_env.data_ptr()), 4, C.c_void_p(r_rob.data_ptr()), 0, n, C.c_float(THR))
    ctx.check(L.gvx_collide_map_map(env.handle, rob.handle, THR, None, 1, C.byref(cnt)))
    out["c2_at_256"] = {"workload": "256^3 ProbVoxelMap + one 307200-point frame (insertSensorData) + collideWith a 256^3 "
                                    "BitVectorVoxelMap, synchronous calls, device-resident frame",
                        "reference_cuda_ms_per_step": t_ref, "this_library_streaming_ms_per_step": res["streaming"],
                        "this_library_default_ms_per_step": res["default"],
                        "reference_voxels_per_s": n / (t_ref["median"] * 1e-3),
                        "counts_equal_on_identical_maps": bool(same == cnt.value), "count": int(cnt.value)}
    env.destroy(); rob.destroy()
    del r_env, r_rob
    # configs[0]: 128^3 BitVectorVoxelMap, insert one 100k-point cloud, collideWith a 128^3 ProbVoxelMap
    d, n = 128, 128 ** 3
    pa = np.random.Generator(np.random.PCG64(101)).uniform(-0.02, 1.30, size=(100000, 3)).astype(np.float32)
    pb = np.random.Generator(np.random.PCG64(102)).uniform(-0.02, 1.30, size=(100000, 3)).astype(np.float32)
    a, b = gv.BitVectorVoxelMap((d, d, d), side), gv.ProbVoxelMap((d, d, d), side)
    pa_d, pb_d = torch.from_numpy(pa).cuda(), torch.from_numpy(pb).cuda()
    ctx.check(L.gvx_insert_points(b.handle, C.c_void_p(pb_d.data_ptr()), 100000, _lib.MEM_DEVICE, 1, None))
    flag = C.c_int(0)

    def ours1():
        ctx.check(L.gvx_insert_points(a.handle, C.c_void_p(pa_d.data_ptr()), 100000, _lib.MEM_DEVICE, 1, C.byref(flag)))
        ctx.check(L.gvx_collide_map_map(a.handle, b.handle, 1.0, None, 1, C.byref(cnt)))
    ra = torch.zeros(n * 32, dtype=torch.uint8, device="cuda")
    rb = torch.full((n,), -128, dtype=torch.int8, device="cuda")
    fl = C.c_int(0)
    REF.gvref_insert_points(C.c_void_p(rb.data_ptr()), 1, d, d, d, C.c_float(side), C.c_void_p(pb_d.data_ptr()), 100000, 1, C.byref(fl))

    def ref1():
        REF.gvref_insert_points(C.c_void_p(ra.data_ptr()), 0, d, d, d, C.c_float(side), C.c_void_p(pa_d.data_ptr()), 100000, 1, C.byref(fl))
        r_cnt[0] = REF.gvref_collide_map_map(C.c_void_p(ra.data_ptr()), 0, C.c_void_p(rb.data_ptr()), 4, n, C.c_float(1.0))
    t_ref1, t_ours1 = wall(ref1), wall(ours1)
    out["c1"] = {"workload": "128^3 BitVectorVoxelMap: insertPointCloud(100k points) + collideWith a 128^3 ProbVoxelMap, "
                             "synchronous calls, device-resident cloud",
                 "reference_cuda_ms_per_step": t_ref1, "this_library_ms_per_step": t_ours1,
                 "counts_equal": bool(r_cnt[0] == cnt.value), "count": int(cnt.value)}
    ctx.check(L.gvx_set_option(1, 1))
    return out


# ---- CPU arms (the ONLY places bench.py touches oracle/) --------------------------------------------------------------
def _c4_sample_inputs():
    """the z-slab sample of C4: the points of every cloud set / of the robot volume whose layer lies in the sample slab
    (+- 1 layer; the oracle and the kernels apply the exact test), so that frame and grid shrink TOGETHER: voxels/s of
    the sample estimates voxels/s of the whole 2048^3 step (both parts are linear in the slab's share)"""
    z0, z1 = C4_SAMPLE

    def sel(p):
        l = z_layer(p)
        return np.ascontiguousarray(p[(l >= z0 - 1) & (l <= z1)])
    sets = [sel(np.concatenate([c4_cloud(k, s) for k in range(C4_CLOUDS)])) for s in range(C4_SETS)]
    return sets, sel(c4_robot_points())


def cpu_arm_c4(steps, warmup, threads=None, use_ref_so=True):
    """One C4 step on the host cores, on the z-slab sample: log-odds fold of the slab's points (OpenMP loop over the
    points, oracle/gvoracle.c gvo_update_logodds_omp) + collideWith Prob x Prob over the slab, through oracle/_ref/
    libgvref_cpu.so (the reference's own voxel / collider headers compiled with g++ -fopenmp) when it exists, else the
    oracle's C port. All host threads, set explicitly (torchrun exports OMP_NUM_THREADS=1)."""
    from oracle import gvoracle as orc
    threads = threads or os.cpu_count() or 1
    orc.set_num_threads(threads)
    dims, (z0, z1) = (C4_DIM, C4_DIM, C4_DIM), C4_SAMPLE
    n_s = C4_DIM * C4_DIM * (z1 - z0)
    sets, rob_pts = _c4_sample_inputs()
    env, rob = orc.new_map(orc.MAP_PROB, dims, z0, z1), orc.new_map(orc.MAP_PROB, dims, z0, z1)
    orc.insert_points(rob, orc.MAP_PROB, dims, SIDE, rob_pts, 1, z0, z1)
    ref, so = None, os.path.join(ROOT, "oracle", "_ref", "libgvref_cpu.so")
    if use_ref_so and os.path.exists(so):
        ref = C.CDLL(so)
        ref.gvrefcpu_collide_map_map.restype = C.c_uint64
        ref.gvrefcpu_collide_map_map.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint64, C.c_float,
                                                 C.c_int64, C.c_int]
        ref.gvrefcpu_set_num_threads(threads)
        cores = ref.gvrefcpu_num_threads()
    else:
        cores = orc.num_threads()
    counts = []

    def step(k):
        orc.update_logodds(env, dims, SIDE, sets[k % C4_SETS], C4_POSE, 72, z0, z1, parallel=True)
        if ref is not None:
            c = ref.gvrefcpu_collide_map_map(env.ctypes.data, 4, rob.ctypes.data, 4, n_s, THR, 0, 1)
        else:
            c = orc.collide_map_map(env, orc.MAP_PROB, rob, orc.MAP_PROB, n_s, THR, 0, True)
        counts.append(int(c))

    for k in range(warmup):
        step(k)
    t0 = time.perf_counter()
    for k in range(steps):
        step(warmup + k)
    dt = time.perf_counter() - t0
    return {"value": n_s * steps / dt, "unit": "voxels/s", "cores": int(cores),
            "kind": "reference" if ref is not None else "port",
            "sample": "z-slab [%d,%d) of the 2048^3 grid = %d voxels per step (1/16 of the grid) with the %d / %d points of "
                      "the 8 M-point frame that fall into it: log-odds fold (OpenMP over the points) + collideWith Prob x "
                      "Prob thr 0.5 with marks (%s), %d timed steps after %d warm-up" %
                      (z0, z1, n_s, len(sets[0]), len(sets[1]), "reference voxel/collider headers, g++ -O2 -fopenmp"
                       if ref is not None else "oracle C port, gcc -O2 -fopenmp", steps, warmup),
            "ms_per_step": dt / steps * 1e3, "counts": counts,
            "extrapolated_full_grid_ms_per_step": dt / steps * 1e3 * (C4_DIM / (z1 - z0))}


def c4_sample_parity(ctx, cpu_counts):
    """the CUDA path on the SAME z-slab sample, step by step: every count must equal the CPU arm's"""
    import gpu_voxels_b200 as gv
    L, _lib = ctx.L, ctx._lib
    dims = (C4_DIM, C4_DIM, C4_DIM)
    sets, rob_pts = _c4_sample_inputs()
    env, rob = gv.ProbVoxelMap(dims, SIDE, slab=C4_SAMPLE), gv.ProbVoxelMap(dims, SIDE, slab=C4_SAMPLE)
    rob.insertPointCloud(rob_pts, gv.eBVM_OCCUPIED, check_outside=False)
    got = []
    for use_summary in (1, 0):
        ctx.check(L.gvx_set_option(1, use_summary))
        env.clearMap()
        g = []
        for k in range(len(cpu_counts)):
            env.insertSensorData(sets[k % C4_SETS], C4_POSE)
            g.append(int(env.collideWith(rob, THR, mark=True)))
        got.append(g)
    ctx.check(L.gvx_set_option(1, 1))
    env.destroy(); rob.destroy()
    return {"cpu_counts": cpu_counts, "gpu_counts_default_path": got[0], "gpu_counts_streaming_path": got[1],
            "equal": got[0] == cpu_counts and got[1] == cpu_counts,
            "what": "count of every step of the z-slab sample [%d,%d): CUDA path (both collide paths) vs the CPU arm" % C4_SAMPLE}


def workload_config(n_gpus):
    return {"workload": "BASELINE.json configs[3] (C4): 2048^3 ProbVoxelMap (8.6 G voxels) z-slab sharded over %d GPU(s); "
                        "per step 8 clouds x 1 M points log-odds-folded into the environment map, then collideWith a 2048^3 "
                        "robot-volume ProbVoxelMap (2 M points), threshold 0.5, collisions marked, count all-reduced" % n_gpus,
            "grid": [C4_DIM, C4_DIM, C4_DIM], "voxels": C4_DIM ** 3, "points_per_step": C4_CLOUDS * C4_POINTS,
            "coll_threshold": THR, "mark_collisions": True,
            "collide_path": "voxel streaming (GVX_OPT_USE_SUMMARY = 0): both grids read in full, 2 B per voxel (SURVEY 8d); "
                            "the library default (1-bit summaries first) is reported under default_path",
            "l2": "inputs larger than L2: 2 x %.2f GB of voxels per GPU and step vs 126 MB of L2; two cloud sets in rotation"
                  % (C4_DIM ** 3 / n_gpus / 1e9),
            "repetitions": "every K-step region is timed %d times, median reported (p10 / p90 beside it)" % REPS,
            "parallelism": "z-slab x%d" % n_gpus if n_gpus > 1 else "single GPU"}


def run_reference_arm(args):
    if int(os.environ.get("RANK", "0")) != 0:
        return
    r = cpu_arm_c4(args.steps, args.warmup)
    line = {"impl": "reference", "metric": "collideWith voxels/s", "value": r["value"], "unit": "voxels/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "i8", "data": "synthetic",
            "config": workload_config(args.gpus),
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": "voxels/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "extrapolated_full_grid_ms_per_step": r["extrapolated_full_grid_ms_per_step"],
            "note": "the reference has no CPU path and its library cannot be built here (Boost/icl_core absent); this arm "
                    "runs the reference's own host-callable voxel/collider code in an OpenMP loop on all host threads, on "
                    "a z-slab sample whose frame shrinks with it (value = sample voxels / time = full-grid voxels/s)"}
    print(json.dumps(line))


# ---- assembling the line ----------------------------------------------------------------------------------------------
def run_gpu_arm(args):
    ctx = Ctx(args)
    rank, world, K = ctx.rank, ctx.world, args.steps
    peak, peak_src = measured_peak_gbs()
    want = set(("c4", "c5", "c2") if args.config == "all" else (args.config,))
    if world > 1:
        want.discard("c2")  # configs[1] is the single-GPU case
    c4 = run_c4(ctx) if "c4" in want else None
    c5 = run_c5(ctx) if "c5" in want else None
    c2 = run_c2(ctx) if "c2" in want else None
    ref_cuda = run_ref_cuda(ctx) if (world == 1 and rank == 0 and not args.no_ref_cuda and "c2" in want) else None
    cpu = parity = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and c4 is not None:
        cpu = cpu_arm_c4(3, 1)
        parity = c4_sample_parity(ctx, cpu["counts"])
    if world > 1:
        ctx.barrier()
    if rank == 0:
        line = {"metric": "collideWith voxels/s", "unit": "voxels/s", "n_gpus": world, "steps": K,
                "warmup": max(args.warmup, 3), "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "i8", "data": "synthetic", "config": workload_config(world)}
        if c4 is not None:
            ms = c4["step"]["median"]
            k_ms = c4["collide_kernel"]["median"]
            achieved = C4_BYTES_PER_VOXEL * c4["n_loc"] / (k_ms * 1e-3) / 1e9
            dk_ms = c4["default_collide_kernel"]["median"]
            e2e_ms = c4["e2e_pipelined"]["median"]
            line.update({
                "value": c4["n_total"] / (ms * 1e-3), "ms_per_step": ms, "ms_per_step_p10_p90": [c4["step"]["p10"], c4["step"]["p90"]],
                "e2e": {"value": c4["n_total"] / (e2e_ms * 1e-3), "unit": "voxels/s", "ms_per_step": e2e_ms,
                        "ms_per_step_p10_p90": [c4["e2e_pipelined"]["p10"], c4["e2e_pipelined"]["p90"]],
                        "h2d_bytes_per_step": c4["h2d_bytes_per_step"], "d2h_bytes_per_step": c4["d2h_bytes_per_step"],
                        "mode": "through the C ABI with HOST buffers: every rank uploads the clouds of ITS sensors from pinned "
                                "host memory (GVX_MEM_HOST_PINNED, copy stream), "
                                + ("hits are routed to the owning rank inside the update kernel over NVLink peer memory, "
                                   if ctx.dispatch else ("the frame is all-gathered with NCCL, " if world > 1 else "")) +
                                "the all-reduced count is read back by the host every step; two steps in flight",
                        "synchronous_calls": {"value": c4["n_total"] / (c4["e2e_sync"]["median"] * 1e-3),
                                              "ms_per_step": c4["e2e_sync"]["median"],
                                              "how": "GVX_MEM_HOST frame then a collide that returns the count, nothing in flight"}},
                "gpu_launches": c4["launches_per_region"],
                "roofline": {"bound": "hbm", "kernel": "k_collide_prob_prob<mark>", "achieved": achieved, "peak": peak,
                             "unit": "GB/s", "frac": achieved / peak,
                             # the ncu capture is of the whole-grid launch (N = 1): per launch like `achieved`, so null
                             # for the slab launches of N > 1 (same kernel, 1/N of the bytes; ratio to algorithmic 1.000)
                             "traffic": traffic_of("r2_collide_prob_prob_traffic.json") if world == 1 else None,
                             "traffic_captured_at": "N = 1 (profiles/r2_collide_prob_prob_traffic.json): 1.000 x algorithmic",
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": C4_BYTES_PER_VOXEL * c4["n_loc"],
                             "algorithmic_bytes_per_voxel": C4_BYTES_PER_VOXEL, "kernel_ms_avg": k_ms,
                             "kernel_ms_avg_how": "%d launches back to back between one CUDA event pair on the launching "
                                                  "stream, 5 repetitions, median; per GPU (this rank's slab)" % K,
                             "frac_of_8000_nominal": achieved / 8000.0, "share_of_step": k_ms / ms},
                "insert": {"metric": "insertSensorData (log-odds) points/s", "value": c4["points_per_step"] / (c4["fold_alone"]["median"] * 1e-3),
                           "unit": "points/s", "ms_per_frame": c4["fold_alone"]["median"], "points": c4["points_per_step"],
                           "how": c4["fold"], "probe_ms": c4["fold_probe_ms"]},
                "default_path": {"value": c4["n_total"] / (c4["default_step"]["median"] * 1e-3), "unit": "voxels/s",
                                 "ms_per_step": c4["default_step"]["median"],
                                 "ms_per_step_p10_p90": [c4["default_step"]["p10"], c4["default_step"]["p90"]],
                                 "collide_kernel_ms": dk_ms, "kernel": "k_collide_summary2<Prob,Prob,mark>",
                                 "summary_bytes_per_voxel": 0.25,
                                 "summary_GBps": 0.25 * c4["n_loc"] / (dk_ms * 1e-3) / 1e9,
                                 "note": "GVX_OPT_USE_SUMMARY = 1 (library default): 1 'known' bit per voxel and side, Prob "
                                         "bytes only for candidate groups; same counts"},
                "collisions": {"streaming": c4["count_streaming"], "default": c4["count_default"], "e2e": c4["count_e2e"]},
                "clocks": c4["clocks"],
                "count_allreduce": ("in-kernel over NVLink peer memory" if ctx.fused else "nccl") if world > 1 else "none",
            })
        if c5 is not None:
            e = c5["entries"]
            line["c5"] = {"workload": "BASELINE.json configs[4]: %d-entry BitVectorVoxelList x 1024^3 BitVectorVoxelMap (32 GiB, "
                                      "2 %% occupied, z-slab x%d), per step count + 256 per-meaning counts, all-reduced" % (e, world),
                          "scaling": "strong"}
            for tag in ("streaming", "default"):
                r = c5[tag]
                ms, kms = r["step"]["median"], r["local_kernel"]["median"]
                line["c5"][tag] = {"entries_per_s": e / (ms * 1e-3), "ms_per_step": ms, "ms_p10_p90": [r["step"]["p10"], r["step"]["p90"]],
                                   "count_only_ms": r["count_only"]["median"], "local_kernel_ms": kms,
                                   "allreduce": r["allreduce"], "ms_per_step_with_nccl_allreduce": r.get("step_nccl", {}).get("median"),
                                   "roofline_68B_per_entry_GBps": C5_BYTES_PER_ENTRY * e / world / (kms * 1e-3) / 1e9,
                                   "collisions": r["collisions"], "per_meaning_total": r["total"], "parity": r.get("parity")}
        if c2 is not None:
            n = c2["n_vox"]
            ks = c2["streaming"]["collide_kernel"]["median"]
            kd = c2["default"]["collide_kernel"]["median"]
            line["c2"] = {"workload": "BASELINE.json configs[1]: 512^3 ProbVoxelMap, one 640x480 frame (307200 pts) log-odds "
                                      "insert per step, then collideWith a 512^3 BitVectorVoxelMap (4 GiB); %d rotating map pairs" % C2_PAIRS,
                          "streaming": {"voxels_per_s": n / (c2["streaming"]["step"]["median"] * 1e-3),
                                        "ms_per_step": c2["streaming"]["step"]["median"], "collide_kernel_ms": ks,
                                        "roofline": {"kernel": "k_collide_voxelwise<Prob,Bit,mark>", "bound": "hbm",
                                                     "achieved": C2_BYTES_PER_VOXEL * n / (ks * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                                                     "frac": C2_BYTES_PER_VOXEL * n / (ks * 1e-3) / 1e9 / peak,
                                                     "frac_of_8000_nominal": C2_BYTES_PER_VOXEL * n / (ks * 1e-3) / 1e9 / 8000.0,
                                                     "traffic": traffic_of("collide_traffic.json")},
                                        "count": c2["streaming"]["count"]},
                          "default": {"voxels_per_s": n / (c2["default"]["step"]["median"] * 1e-3),
                                      "ms_per_step": c2["default"]["step"]["median"], "collide_kernel_ms": kd,
                                      "summary_GBps": 0.25 * n / (kd * 1e-3) / 1e9, "count": c2["default"]["count"]},
                          "insert_points_per_s": N_FRAME / (c2["fold_alone"]["median"] * 1e-3), "insert_ms": c2["fold_alone"]["median"],
                          "e2e": {"voxels_per_s": n / (c2["e2e_pipelined"]["median"] * 1e-3), "ms_per_step": c2["e2e_pipelined"]["median"],
                                  "h2d_bytes_per_step": c2["h2d_bytes_per_step"], "d2h_bytes_per_step": c2["d2h_bytes_per_step"]}}
        if ref_cuda is not None:
            line["ref_cuda"] = ref_cuda
        if cpu is not None:
            line["cpu_baseline"] = {k: cpu[k] for k in ("value", "unit", "cores", "kind", "sample")}
            line["cpu_baseline"]["extrapolated_full_grid_ms_per_step"] = cpu["extrapolated_full_grid_ms_per_step"]
            line["parity_check"] = parity
        print(json.dumps(line))
        if parity is not None and not parity["equal"]:
            raise SystemExit("bench.py: the CUDA path and the CPU arm disagree on the z-slab sample: %s" % json.dumps(parity))
        for tag in ("streaming", "default"):
            if c5 is not None and c5[tag].get("parity") and not all(v for k, v in c5[tag]["parity"].items() if k.endswith("_equal")):
                raise SystemExit("bench.py: C5 results differ from the numpy restatement: %s" % json.dumps(c5[tag]["parity"]))
    if world > 1:
        ctx.dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="all", choices=["all", "c4", "c5", "c2"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true")
    ap.add_argument("--nccl-count", action="store_true", help="all-reduce the count with NCCL instead of in-kernel")
    ap.add_argument("--pdl", type=int, default=None, choices=[0, 1],
                    help="GVX_OPT_PDL: programmatic dependent launch for the per-step kernels (library default when absent)")
    ap.add_argument("--fold", default="auto", choices=["auto", "dispatch", "multi"],
                    help="sharded C4: route each rank's clouds to the owning ranks in-kernel (dispatch), fold the whole frame "
                         "on every rank (multi), or auto: whichever measures faster in the warm-up")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_gpu_arm(args)


if __name__ == "__main__":
    main()
