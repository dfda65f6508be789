"""world_size-2 (and 3) `gloo` tests of the multi-GPU host logic (gpu_voxels_b200/sharding.py) on CPU: z-slab
partition, list id ranges, batch split, and the single all-reduce of counts / masks / per-meaning counts. The local
"GPU" work of each rank is played by the CPU oracle (tests may use it), so the test proves that slab-wise results
reduced over ranks equal the unsharded result bit for bit."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpu_voxels_b200 import sharding

BIT, PROB = 0, 4
DIMS, SIDE = (40, 36, 37), 0.01  # 37 slices: uneven split over 2 and 3 ranks


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _cloud(seed, n):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.uniform(-0.01, 0.41, size=(n, 3)).astype(np.float32)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import gvoracle as orc
        n = DIMS[0] * DIMS[1] * DIMS[2]
        z0, z1 = sharding.slab_for_rank(DIMS[2], rank, world)
        first, n_loc = sharding.voxel_range_for_rank(DIMS, rank, world)
        env_pts, rob_pts = _cloud(1, 30000), _cloud(2, 30000)
        # every rank voxelises the WHOLE cloud into its slab
        env = orc.new_map(PROB, DIMS, z0, z1)
        rob = orc.new_map(BIT, DIMS, z0, z1)
        outside = orc.insert_points(env, PROB, DIMS, SIDE, env_pts, 1, z0, z1)
        orc.insert_meta(rob, BIT, DIMS, SIDE, rob_pts, [10000, 20000], [5, 77], None, z0, z1)
        # map x map: local count, one all-reduce
        loc = orc.collide_map_map(rob, BIT, env, PROB, n_loc, 1.0, 0, False)
        cnt, _, _ = sharding.allreduce_collision_result(loc)
        # types: local mask, OR over ranks
        lc, lmask = orc.collide_bit_prob_types(rob.copy(), env, n_loc, 1.0, False)
        tcnt, tmask, _ = sharding.allreduce_collision_result(lc, mask32=lmask)
        # list x map with per-meaning counts: the (replicated) list gathers only from this rank's slab
        lst = orc.VoxelList(DIMS, SIDE)
        lst.insert_meta(rob_pts, [10000, 20000], [5, 77])
        i0, i1 = sharding.list_entries_for_rank(lst.ids, DIMS, rank, world)
        l_tot, _, l_per = lst.collide(3, env, PROB, 1.0, 0, False, None, first, n_loc)
        in_range = np.zeros(len(lst), bool); in_range[i0:i1] = True
        ids = lst.ids.astype(np.int64)
        assert np.array_equal(in_range, (ids >= first) & (ids < first + n_loc))
        ptot, _, pper = sharding.allreduce_collision_result(l_tot, per_meaning=l_per)
        any_out = sharding.allreduce_flag(outside)
        # batch split: each rank computes "its" configurations, all-gather restores the order
        n_cfg = 11
        c0, c1 = sharding.batch_range_for_rank(n_cfg, rank, world)
        counts = sharding.allgather_batch_counts(np.arange(c0, c1, dtype=np.uint64) * 3 + 1, n_cfg)
        cd = torch.tensor([loc], dtype=torch.int64)
        sharding.allreduce_count_tensor(cd)
        if rank == 0:
            out.put(dict(cnt=cnt, tcnt=tcnt, tmask=tmask, ptot=ptot, pper=pper, any_out=any_out, counts=counts,
                         cd=int(cd.item())))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_results_reduce_to_the_unsharded_result(world, oracle):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = q.get(timeout=180)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # the unsharded truth
    n = DIMS[0] * DIMS[1] * DIMS[2]
    env_pts, rob_pts = _cloud(1, 30000), _cloud(2, 30000)
    env, rob = oracle.new_map(PROB, DIMS), oracle.new_map(BIT, DIMS)
    outside = oracle.insert_points(env, PROB, DIMS, SIDE, env_pts, 1)
    oracle.insert_meta(rob, BIT, DIMS, SIDE, rob_pts, [10000, 20000], [5, 77])
    want = oracle.collide_map_map(rob, BIT, env, PROB, n, 1.0, 0, False)
    wt, wmask = oracle.collide_bit_prob_types(rob.copy(), env, n, 1.0, False)
    lst = oracle.VoxelList(DIMS, SIDE)
    lst.insert_meta(rob_pts, [10000, 20000], [5, 77])
    w_tot, _, w_per = lst.collide(3, env, PROB, 1.0, 0, False)
    assert want > 0 and got["cnt"] == want and got["cd"] == want
    assert got["tcnt"] == wt and np.array_equal(got["tmask"], wmask) and oracle.bits_of(wmask) == [5, 77]
    assert got["ptot"] == w_tot and np.array_equal(got["pper"], w_per)
    assert got["any_out"] == outside and outside
    assert np.array_equal(got["counts"], np.arange(11, dtype=np.uint64) * 3 + 1)


def test_partitions_cover_everything_exactly_once():
    for dz, world in ((512, 8), (37, 3), (5, 5), (2048, 7)):
        zs = [sharding.slab_for_rank(dz, r, world) for r in range(world)]
        assert zs[0][0] == 0 and zs[-1][1] == dz
        assert all(zs[r][1] == zs[r + 1][0] for r in range(world - 1))
        assert max(b - a for a, b in zs) - min(b - a for a, b in zs) <= 1
    for n_cfg, world in ((1024, 8), (3, 8), (0, 2), (11, 3)):
        rs = [sharding.batch_range_for_rank(n_cfg, r, world) for r in range(world)]
        assert rs[0][0] == 0 and rs[-1][1] == n_cfg and all(rs[r][1] == rs[r + 1][0] for r in range(world - 1))
    with pytest.raises(ValueError):
        sharding.slab_for_rank(4, 0, 5)
    ids = np.array([0, 5, 1439, 1440, 1441, 53279], np.uint32)  # plane = 1440
    assert sharding.list_entries_for_rank(ids, DIMS, 0, 37) == (0, 3)
    assert sharding.list_entries_for_rank(ids, DIMS, 1, 37) == (3, 5)
    assert sharding.list_entries_for_rank(ids, DIMS, 0, 37, offset=1) == (0, 2)
    assert sharding.allreduce_collision_result(7, np.arange(32, dtype=np.uint8), np.arange(256))[0] == 7  # no group: identity
