"""Multi-GPU check of the sharded path on REAL GPUs (one rank per GPU), run as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tests/multi_gpu_check.py
Not collected by pytest (the -m gpu tier runs on one GPU; kernels that wait on other ranks must never share a GPU).
Checks, bit-exactly against the CPU oracle on the unsharded grid:
  * z-slab shards of Prob and BitVector maps filled from the whole cloud,
  * collideWith with the count all-reduced INSIDE the kernel over NVLink peer memory == NCCL all-reduce == oracle,
    repeated (epoch parity double-buffering), map x map and list x map,
  * batch of robot configurations split over ranks and all-gathered,
  * sensor frames routed to the owning rank inside the update kernel (dispatch over peer memory) == multi-frame fold.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import gpu_voxels_b200 as gv  # noqa: E402
from gpu_voxels_b200 import _lib, sharding  # noqa: E402
from oracle import gvoracle as orc  # noqa: E402

BIT, PROB = 0, 4


class pytest_raises:
    """the library must refuse (GvxError) inside the block"""

    def __enter__(self):
        return self

    def __exit__(self, et, ev, tb):
        from gpu_voxels_b200 import _lib
        assert et is not None and issubclass(et, _lib.GvxError), "expected the library to refuse the call"
        return True


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    sys.stdout.flush()
    fd = os.dup(1); os.dup2(2, 1)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dist.all_reduce(torch.zeros(1, device="cuda")); torch.cuda.synchronize()
    os.dup2(fd, 1); os.close(fd)
    L = _lib.init(local)
    stream = torch.cuda.Stream(); torch.cuda.set_stream(stream)
    _lib.check(L.gvx_set_stream(C.c_void_p(stream.cuda_stream)))
    fused = sharding.connect_peer_mailboxes()
    assert fused, "CUDA IPC peer mailboxes could not be connected"

    dims, side = (96, 80, 101), 0.01  # 101 slices: uneven slabs
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(5))
    env_pts = rng.uniform(-0.01, 1.02, size=(200000, 3)).astype(np.float32)
    rob_pts = np.concatenate([env_pts[:30000], rng.uniform(0, 1.0, size=(90000, 3)).astype(np.float32)])
    slab = sharding.slab_for_rank(dims[2], rank, world)
    env = gv.ProbVoxelMap(dims, side, slab=slab)
    rob = gv.BitVectorVoxelMap(dims, side, slab=slab)
    env.insertPointCloud(env_pts, 1)
    rob.insertMetaPointCloud(gv.MetaPointCloud([rob_pts[:50000], rob_pts[50000:]]), [9, 70])
    r_env, r_rob = orc.new_map(PROB, dims), orc.new_map(BIT, dims)
    orc.insert_points(r_env, PROB, dims, side, env_pts, 1)
    orc.insert_meta(r_rob, BIT, dims, side, rob_pts, [50000, 70000], [9, 70])
    first, n_loc = sharding.voxel_range_for_rank(dims, rank, world)
    assert np.array_equal(env.download(), r_env[first:first + n_loc]), "Prob shard differs from the slice of the full map"
    assert np.array_equal(rob.download().reshape(-1), r_rob[first * 32:(first + n_loc) * 32]), "Bit shard differs"
    want = orc.collide_map_map(r_rob.copy(), BIT, r_env, PROB, n, 1.0, 0, False)
    assert want > 1000
    cnt_d = torch.zeros(1, dtype=torch.int64, device="cuda")
    for it in range(7):  # odd and even epochs
        host = C.c_uint64(0)
        _lib.check(L.gvx_collide_map_map_allreduce(rob.handle, env.handle, 1.0, 0, C.byref(host), C.c_void_p(cnt_d.data_ptr())))
        assert host.value == want and int(cnt_d.item()) == want, (it, host.value, int(cnt_d.item()), want)
    # back-to-back without host syncs in between, then one sync
    for it in range(20):
        _lib.check(L.gvx_collide_map_map_allreduce(rob.handle, env.handle, 1.0, 0, None, C.c_void_p(cnt_d.data_ptr())))
    torch.cuda.synchronize()
    assert int(cnt_d.item()) == want
    # NCCL path gives the same
    _lib.check(L.gvx_collide_map_map_async(rob.handle, env.handle, 1.0, None, 0, C.c_void_p(cnt_d.data_ptr())))
    dist.all_reduce(cnt_d)
    assert int(cnt_d.item()) == want
    # collideWith with an OFFSET on explicit slab maps: the partners in other ranks' slabs are read from the owners' GPUs
    # (gvx_map_connect_peers on the `other` map); before the connection the call is refused
    with pytest_raises():
        rob.collideWith(env, 1.0, (0, 0, 3), mark=False)
    assert sharding.connect_map_peers(env), "CUDA IPC mapping of the peers' slabs refused"
    for off in ((0, 0, 3), (2, -1, -9), (5, 0, 0), (-1, 4, dims[2] // 2 + 1)):
        lin = off[2] * dims[0] * dims[1] + off[1] * dims[0] + off[0]
        w = orc.collide_map_map(r_rob.copy(), BIT, r_env, PROB, n, 1.0, lin, False)
        got = sharding.collide_with_offset(rob, env, 1.0, off, mark=False)
        assert got == w, ("offset collide on slab maps", off, got, w)
    off = (1, 1, -2)   # ... and with marks: the marked slab is the slice of the oracle's marked map
    lin = off[2] * dims[0] * dims[1] + off[1] * dims[0] + off[0]
    r_rob_m = r_rob.copy()
    w = orc.collide_map_map(r_rob_m, BIT, r_env, PROB, n, 1.0, lin, True)
    rob_m = gv.BitVectorVoxelMap(dims, side, slab=slab)
    rob_m.insertMetaPointCloud(gv.MetaPointCloud([rob_pts[:50000], rob_pts[50000:]]), [9, 70])
    assert sharding.collide_with_offset(rob_m, env, 1.0, off, mark=True) == w and w > 0
    assert np.array_equal(rob_m.download().reshape(-1), r_rob_m[first * 32:(first + n_loc) * 32]), "marks of an offset collide"
    _lib.check(L.gvx_map_disconnect_peers(env.handle))
    dist.barrier()
    # list x map: replicated list, sharded map
    lst = gv.BitVectorVoxelList(dims, side)
    lst.insertMetaPointCloud(gv.MetaPointCloud([rob_pts[:50000], rob_pts[50000:]]), [9, 70])
    r_lst = orc.VoxelList(dims, side)
    r_lst.insert_meta(rob_pts, [50000, 70000], [9, 70])
    l_want = r_lst.collide(0, r_env, PROB, 1.0, 0, False)[0]
    host = C.c_uint64(0)
    _lib.check(L.gvx_collide_list_map_allreduce(lst.handle, env.handle, 1.0, 0, C.byref(host), None))
    assert host.value == l_want and l_want > 0, (host.value, l_want)
    tot, per = lst.collideCountingPerMeaning(env, mark=False)
    g_tot, _, g_per = sharding.allreduce_collision_result(tot, per_meaning=per)
    w_tot, _, w_per = r_lst.collide(3, r_env, PROB, 1.0, 0, False)
    assert g_tot == w_tot and np.array_equal(g_per, w_per)
    # the same without a host round trip: 257 words stay on the device and are summed by ONE NCCL all-reduce (bench.py C5);
    # every rank reads only the entries of its slab (two binary searches on the sorted ids, cached per list version)
    out_d = torch.zeros(257, dtype=torch.int64, device="cuda")
    for it in range(3):
        _lib.check(L.gvx_collide_list_map_per_meaning_async(lst.handle, env.handle, 1.0, None, 0, C.c_void_p(out_d.data_ptr())))
        dist.all_reduce(out_d)
        got = out_d.cpu().numpy().astype(np.uint64)
        assert int(got[0]) == w_tot and np.array_equal(got[1:], w_per), (it, int(got[0]), w_tot)
    # ... and with the 257 words all-reduced INSIDE the kernel over the peers' mailboxes (no NCCL launch), repeated: both
    # epoch parities of the vector area, with scalar-only collectives in between (the vector area keeps an epoch sequence
    # of its own: 1, 2 or 0 scalar all-reduces between two vector ones shift the two sequences against each other)
    for it in range(7):
        out_d.fill_(-1)
        _lib.check(L.gvx_collide_list_map_per_meaning_allreduce(lst.handle, env.handle, 1.0, 0, C.c_void_p(out_d.data_ptr())))
        _lib.check(L.gvx_sync())
        got = out_d.cpu().numpy().astype(np.uint64)
        assert int(got[0]) == w_tot and np.array_equal(got[1:], w_per), ("in-kernel vector all-reduce", it, int(got[0]), w_tot)
        for k in range(it % 3):
            host = C.c_uint64(0)
            _lib.check(L.gvx_collide_list_map_allreduce(lst.handle, env.handle, 1.0, 0, C.byref(host), None))
            assert host.value == l_want, ("scalar all-reduce between vector ones", it, k, host.value, l_want)
    # an offset moves the slab boundaries inside the list: ids + off must fall into the slab
    off = (3, -2, 5)
    lin = off[2] * dims[0] * dims[1] + off[1] * dims[0] + off[0]
    w_off = r_lst.collide(0, r_env, PROB, 1.0, lin, False)[0]
    c_off = lst.collideWith(env, 1.0, off, mark=False)
    t_off = torch.tensor([c_off], dtype=torch.int64, device="cuda")
    dist.all_reduce(t_off)
    assert int(t_off.item()) == w_off, (int(t_off.item()), w_off)
    lst.insertPointCloud(env_pts[100000:105000], 33)   # the list changes (new voxels, all occupied in env): the cached
    r_lst.insert_points(env_pts[100000:105000], 33)    # slab range must not survive
    l_want2 = r_lst.collide(0, r_env, PROB, 1.0, 0, False)[0]
    host = C.c_uint64(0)
    _lib.check(L.gvx_collide_list_map_allreduce(lst.handle, env.handle, 1.0, 0, C.byref(host), None))
    assert host.value == l_want2 and l_want2 > l_want, (host.value, l_want2, l_want)
    # disconnect + connect in one process must not match the stale tags of the previous session
    _lib.check(L.gvx_peer_mailbox_disconnect())
    dist.barrier()
    assert sharding.connect_peer_mailboxes()
    for it in range(3):
        host = C.c_uint64(0)
        _lib.check(L.gvx_collide_map_map_allreduce(rob.handle, env.handle, 1.0, 0, C.byref(host), None))
        assert host.value == want, ("after reconnect", it, host.value, want)
    # batch split over ranks (environment replicated)
    full_env = gv.BitVectorVoxelMap(dims, side)
    full_env.insertPointCloud(env_pts, 1)
    links = [rob_pts[:4000] - 0.5, rob_pts[4000:7000] - 0.5]
    T = np.tile(np.eye(4, dtype=np.float32), (13, 2, 1, 1))
    T[:, :, :3, 3] = rng.uniform(0.3, 0.7, size=(13, 2, 3)).astype(np.float32)
    c0, c1 = sharding.batch_range_for_rank(13, rank, world)
    mine = gv.batch_robot_collide(full_env, gv.MetaPointCloud(links), [4, 5], T[c0:c1]) if c1 > c0 else np.zeros(0, np.uint64)
    allc = sharding.allgather_batch_counts(mine, 13)
    r_full = orc.new_map(BIT, dims)
    orc.insert_points(r_full, BIT, dims, side, env_pts, 1)
    w = orc.batch_robot_collide(r_full, BIT, dims, side, np.concatenate(links), [4000, 3000], [4, 5], T, None, 1.0)
    assert np.array_equal(allc, w), (allc, w)
    # sensor frames routed to the owning rank inside the update kernel (gvx_update_logodds_dispatch): one frame per
    # rank per step, most hits land in OTHER ranks' slabs; 6 steps (both inbox parities, credits, saturation at 127),
    # ragged frame sizes incl. an empty frame; result == the slab of the oracle's unsharded map == the multi-frame fold
    n_max = 60000
    assert sharding.connect_peer_dispatch(n_max), "CUDA IPC dispatch inboxes could not be connected"
    zb = sharding.slab_bounds(dims[2], world)
    d_env = gv.ProbVoxelMap(dims, side, slab=slab)
    m_env = gv.ProbVoxelMap(dims, side, slab=slab)
    o_env = orc.new_map(PROB, dims)
    for step in range(6):
        frames, poses = [], []
        for r in range(world):
            fr = np.random.Generator(np.random.PCG64(1000 + 31 * step + r))
            k = 0 if (step == 2 and r == world - 1) else int(fr.integers(n_max // 2, n_max + 1))
            pts = fr.uniform(-0.6, 0.6, size=(k, 3)).astype(np.float32)
            pts[: k // 3] = (pts[: k // 3] * 0.05).astype(np.float32)  # a dense clump: many hits per voxel -> saturation
            pts[k // 2: k // 2 + 50] = np.nan
            pose = np.eye(4, dtype=np.float32)[:3].copy()
            pose[:, 3] = (0.5, 0.4, 0.05 + 0.9 * (r + 0.5) / world)
            if step == 4 and r == 0:
                pose[:, 3] = (0.5, 0.4, 1.5)   # sensor outside the map: the whole frame is ignored
            frames.append(pts); poses.append(pose)
        d_env.insertSensorDataDispatch(frames[rank], poses[rank], zb)
        m_env.insertSensorDataMulti(frames, poses)
        for r in range(world):
            orc.update_logodds(o_env, dims, side, frames[r], poses[r])
    got_d, got_m = d_env.download(), m_env.download()
    assert np.array_equal(got_m, o_env[first:first + n_loc]), "multi-frame fold differs from the oracle"
    assert np.array_equal(got_d, o_env[first:first + n_loc]), "dispatch differs from the oracle: %d voxels" % int((got_d != o_env[first:first + n_loc]).sum())
    assert (o_env == 127).sum() > 10 and (o_env != -128).sum() > 50000
    # summaries stay in sync: the summary-path collide equals the voxel-streaming one
    c_sum = d_env.collideWith(rob, 0.3, mark=False)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 0)
    c_str = d_env.collideWith(rob, 0.3, mark=False)
    gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    assert c_sum == c_str
    # back to back without host syncs, all ranks in lock step only through the kernels themselves
    for step in range(25):
        d_env.insertSensorDataDispatch(frames[rank][: 1000 * (1 + (step + rank) % 7)], poses[rank], zb)
    _lib.check(L.gvx_sync())
    d_env.insertSensorDataDispatch(frames[rank][:0], poses[rank], zb)   # reports a time-out of the run above, if any
    _lib.check(L.gvx_sync())
    _lib.check(L.gvx_peer_dispatch_disconnect())
    dist.barrier()
    # a second dispatch session in the same process (same world and capacity: the allocation, its ready[] tags
    # and ack[] credits are reused) -- 3 more frames on top of the 6 + 26 of the first session
    assert sharding.connect_peer_dispatch(n_max)
    d2, o2 = gv.ProbVoxelMap(dims, side, slab=slab), orc.new_map(PROB, dims)
    for step in range(3):
        for r in range(world):
            fr = np.random.Generator(np.random.PCG64(5000 + 17 * step + r))
            pts = fr.uniform(-0.5, 0.5, size=(20000 + 1000 * r, 3)).astype(np.float32)
            pose = np.eye(4, dtype=np.float32)[:3].copy()
            pose[:, 3] = (0.5, 0.4, 0.05 + 0.9 * (r + 0.5) / world)
            if r == rank:
                d2.insertSensorDataDispatch(pts, pose, zb)
            orc.update_logodds(o2, dims, side, pts, pose)
    assert np.array_equal(d2.download(), o2[first:first + n_loc]), "dispatch after a reconnect differs from the oracle"
    _lib.check(L.gvx_sync())
    _lib.check(L.gvx_peer_dispatch_disconnect())
    dist.barrier()
    if rank == 0:
        print("multi_gpu_check ok: world %d, fused in-kernel all-reduce == NCCL == oracle (%d, list %d), batch split ok, frame dispatch == multi-frame fold == oracle" % (world, want, l_want))
    _lib.check(L.gvx_peer_mailbox_disconnect())
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
