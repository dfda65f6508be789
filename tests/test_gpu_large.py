"""GPU parity at the sizes the reference cannot reach (missing #2): maps above 2^32 voxels, a 32 GiB
BitVectorVoxelMap against a 10 M-entry list, the planner batch at its real size. The oracle's dense loops do not finish at
these sizes, so expected results come from its 64-bit index arithmetic applied SPARSELY: the voxel ids of the seeded points
(oracle.map_to_voxels = GV/voxelmap/kernels/VoxelMapOperations.h:123-133 with the device's saturating convert; linear
index as in :46-58 but in 64 bits), the log-odds value each id must hold, set intersections for the counts. Every voxel of
the device map is checked: the expected ids hold the expected bytes AND the number of non-empty voxels equals the
number of expected ids (nothing else was touched). Slow (tens of seconds), collected like every other -m gpu test."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


class _Dev:
    """a raw device pointer as a torch tensor (no copy), through __cuda_array_interface__"""

    def __init__(self, ptr, n, typestr):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def _device_view(gv, m, typestr, n):
    import torch
    L = gv._lib.load()
    p = C.c_void_p()
    gv._lib.check(L.gvx_map_device_ptr_const(m.handle, C.byref(p), None, None))  # read-only: the summary stays valid
    gv._lib.check(L.gvx_sync())
    return torch.as_tensor(_Dev(p.value, n, typestr), device="cuda")


def _ids64(oracle, pts, dims, side, pose_tz=None):
    """64-bit linear voxel ids of the points that fall inside the map (the oracle's float -> voxel arithmetic)"""
    p = pts if pose_tz is None else (pts + np.array([0, 0, pose_tz], np.float32)).astype(np.float32)
    p = p[~np.isnan(p).any(1)]
    c = oracle.map_to_voxels(side, p).astype(np.int64)
    ok = (c[:, 0] < dims[0]) & (c[:, 1] < dims[1]) & (c[:, 2] < dims[2])
    c = c[ok]
    return c[:, 2] * (dims[0] * dims[1]) + c[:, 1] * dims[0] + c[:, 0]


def test_prob_map_above_2_pow_32_voxels(gv, oracle):
    """2048 x 2048 x 1056 ProbVoxelMap = 4 429 185 024 voxels (> 2^32; the reference's uint32 voxel count overflows,
    GV/voxelmap/TemplateVoxelMap.hpp:65-75). 1 M seeded points per cloud, a third of them in the layers above voxel 2^32."""
    import torch
    dims, side = (2048, 2048, 1056), 0.01
    n = dims[0] * dims[1] * dims[2]
    assert n > 2 ** 32
    rng = np.random.Generator(np.random.PCG64(7001))

    def cloud(k):
        p = rng.uniform(0.0, 20.48, size=(k, 3))
        p[:, 2] = rng.uniform(0.0, 10.56, size=k)
        p[: k // 3, 2] = rng.uniform(10.24, 10.56, size=k // 3)   # linear index >= 2^32
        p[5] = (20.475, 20.475, 10.555)                            # the last voxel of the map (its centre)
        p[6] = (0.005, 0.005, 10.245)                              # index exactly 2^32
        p[7] = (20.4801, 1.0, 1.0)                                 # outside in x
        p[8] = (1.0, 1.0, 10.5601)                                 # outside in z
        return p.astype(np.float32)
    pa, pb = cloud(1_000_000), cloud(1_000_000)
    pb[:300_000] = pa[:300_000]
    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.0]], np.float32)

    a, b = gv.ProbVoxelMap(dims, side), gv.ProbVoxelMap(dims, side)
    assert a.insertPointCloud(pa, gv.eBVM_OCCUPIED) is True      # p[7], p[8] are outside
    ia = _ids64(oracle, pa, dims, side)
    ua = np.unique(ia)
    assert ua.max() == n - 1 and (ua == 2 ** 32).any() and (ua > 2 ** 32).sum() > 300_000
    va = _device_view(gv, a, "|i1", n)
    ids_d = torch.from_numpy(ua).cuda()
    assert bool((va[ids_d] == 127).all()), "a voxel of the cloud does not hold MAX_PROBABILITY"
    assert int((va != -128).sum()) == len(ua), "voxels outside the cloud were written"
    # log-odds: the same frame three times -> clamp(-128 + 72 c) per voxel with c hits per frame
    for _ in range(3):
        b.insertSensorData(pb, pose)
    ib = _ids64(oracle, pb, dims, side, 0.0)
    ub, cb = np.unique(ib, return_counts=True)
    want_b = np.minimum(-128 + 72 * 3 * cb, 127).astype(np.int8)
    vb = _device_view(gv, b, "|i1", n)
    got_b = vb[torch.from_numpy(ub).cuda()].cpu().numpy()
    assert np.array_equal(got_b, want_b)
    assert int((vb != -128).sum()) == len(ub)
    # collideWith, threshold 0.5 -> int8 0: a is 127 everywhere it is known; b collides where its value >= 0
    want = len(np.intersect1d(ua, ub[want_b >= 0]))
    assert want > 250_000
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            assert a.collideWith(b, 0.5, mark=False) == want
            assert b.collideWith(a, 0.5, mark=False) == want
            assert a.collisionCheck(b, 0.5) is True
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    # marks: b's colliding voxels become 127, nothing else changes
    assert b.collideWith(a, 0.5, mark=True) == want
    want_b2 = want_b.copy()
    want_b2[np.isin(ub, ua) & (want_b >= 0)] = 127
    assert np.array_equal(vb[torch.from_numpy(ub).cuda()].cpu().numpy(), want_b2)
    assert int((vb != -128).sum()) == len(ub)
    # z-slabs above 2^32: the top slab alone equals that slice of the whole map
    z0 = 1030
    top = gv.ProbVoxelMap(dims, side, slab=(z0, dims[2]))
    top.insertPointCloud(pa, gv.eBVM_OCCUPIED)
    first = z0 * dims[0] * dims[1]
    vt = _device_view(gv, top, "|i1", n - first)
    assert bool(torch.equal(vt, va[first:]))
    del va, vb, vt
    a.destroy(); b.destroy(); top.destroy()
    torch.cuda.empty_cache()


def test_bitvector_map_above_4_gib_insert_and_collide(gv, oracle):
    """1024 x 1024 x 160 BitVectorVoxelMap = 5 GiB (byte offsets above 2^32; refused by the reference): per-cloud
    meanings, every touched voxel compared word by word, collideWith a ProbVoxelMap with marks"""
    import torch
    dims, side = (1024, 1024, 160), 0.01
    n = dims[0] * dims[1] * dims[2]
    rng = np.random.Generator(np.random.PCG64(7002))
    p1 = (rng.uniform(0, 1, size=(400_000, 3)) * (10.24, 10.24, 1.6)).astype(np.float32)
    p2 = np.concatenate([p1[:100_000], (rng.uniform(0, 1, size=(300_000, 3)) * (10.24, 10.24, 1.6)).astype(np.float32)])
    p2[-1] = (10.235, 10.235, 1.595)   # centre of the last voxel
    m = gv.BitVectorVoxelMap(dims, side)
    m.insertMetaPointCloud(gv.MetaPointCloud([p1, p2]), [9, 70])
    i1, i2 = np.unique(_ids64(oracle, p1, dims, side)), np.unique(_ids64(oracle, p2, dims, side))
    assert i2.max() == n - 1 and i2.max() * 32 > 2 ** 32
    words = _device_view(gv, m, "<i4", n * 8).reshape(n, 8)
    allv = np.union1d(i1, i2)
    want = np.zeros((len(allv), 8), np.uint32)
    want[np.isin(allv, i1), 9 >> 5] |= np.uint32(1 << 9)
    want[np.isin(allv, i2), 70 >> 5] |= np.uint32(1 << (70 & 31))
    got = words[torch.from_numpy(allv).cuda()].cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want)
    assert int((words != 0).any(1).sum()) == len(allv)
    env = gv.ProbVoxelMap(dims, side)
    env.insertPointCloud(p2, gv.eBVM_OCCUPIED)
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            assert m.collideWith(env, 1.0, mark=False) == len(i2)
            assert env.collideWith(m, 1.0, mark=False) == len(i2)
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    types = gv.BitVectorVoxel()
    assert m.collideWithTypes(env, types, 1.0) == len(i2)      # marks eBVM_COLLISION on the colliding voxels
    assert types.bits() == [9, 70]
    want[np.isin(allv, i2), 0] |= np.uint32(4)
    got = words[torch.from_numpy(allv).cuda()].cpu().numpy().view(np.uint32)
    assert np.array_equal(got, want)
    del words
    m.destroy(); env.destroy()
    torch.cuda.empty_cache()


def test_c5_full_size_list_vs_32_gib_map(gv, oracle):
    """BASELINE.json configs[4] at its real size: 10 M-entry BitVectorVoxelList x 1024^3 BitVectorVoxelMap (32 GiB),
    count, 256-bit meaning mask, 256 per-meaning counts, type-mask count and the eBVM_COLLISION marks on the list, against
    the numpy restatement on the ids (bench.c5_expected; semantics GV/voxellist/kernels/VoxelListOperations.hpp:126-314,
    GV/voxellist/BitVoxelList.hpp:300-344). Summary path and voxel-streaming path; whole map and 3 uneven z-slabs."""
    import torch
    import bench as B
    L = gv._lib.load()
    dims = (B.C5_DIM,) * 3
    ids, voxels, boxes = B.c5_inputs()
    want_cnt, want_per, hit = B.c5_expected(ids, voxels, boxes)
    want_mask = np.packbits((want_per > 0).astype(np.uint8), bitorder="little")
    assert want_cnt > 150_000

    def fill(m):
        r = torch.arange(47, device="cuda")
        bz, by, bx = torch.meshgrid(r, r, r, indexing="ij")
        local = torch.stack([bx, by, bz], -1).reshape(-1, 3).to(torch.float32)
        for o in boxes:
            pts = (((local + torch.tensor(o, device="cuda", dtype=torch.float32)) + 0.5) * B.SIDE).contiguous()
            gv._lib.check(L.gvx_insert_points(m.handle, C.c_void_p(pts.data_ptr()), pts.shape[0], gv._lib.MEM_DEVICE, 1, None))
            gv._lib.check(L.gvx_sync())

    lst = gv.BitVectorVoxelList(dims, B.SIDE)
    coords = np.stack([ids % 1024, (ids // 1024) % 1024, ids // (1024 * 1024)], -1).astype(np.uint32)
    lst.upload(ids, coords, voxels)
    m = gv.BitVectorVoxelMap(dims, B.SIDE)
    fill(m)
    mk = gv.BitVectorVoxel(); mk.setBit(7); mk.setBit(20)
    sel = (np.unpackbits(voxels[hit], axis=1, bitorder="little")[:, [7, 20]].any(1)).sum()
    for use_summary in (1, 0):
        gv.set_option(gv.GVX_OPT_USE_SUMMARY, use_summary)
        try:
            assert lst.collideWith(m, 1.0, None, mark=False) == want_cnt
            tot, per = lst.collideCountingPerMeaning(m, None, 1.0, None, mark=False)
            assert tot == int(want_per.sum()) and np.array_equal(per, want_per)
            types = gv.BitVectorVoxel()
            assert lst.collideWithTypes(m, types, 1.0, None, mark=False) == want_cnt
            assert np.array_equal(types.bytes, want_mask)
            assert lst.collideWithTypeMask(m, mk, 1.0, None, mark=False) == int(sel)
        finally:
            gv.set_option(gv.GVX_OPT_USE_SUMMARY, 1)
    m.destroy()
    torch.cuda.empty_cache()
    # z-slabs (what the ranks of a sharded run hold): partial results add up; marks land on exactly the colliding entries
    cuts = [0, 300, 700, 1024]
    tot_cnt, tot_per = 0, np.zeros(256, np.uint64)
    for z0, z1 in zip(cuts[:-1], cuts[1:]):
        s = gv.BitVectorVoxelMap(dims, B.SIDE, slab=(z0, z1))
        fill(s)
        lo, hi = z0 * 1024 * 1024, z1 * 1024 * 1024
        in_slab = (ids >= lo) & (ids < hi)
        c = lst.collideWith(s, 1.0, None, mark=True)
        assert c == int((hit & in_slab).sum())
        t, per = lst.collideCountingPerMeaning(s, None, 1.0, None, mark=False)
        tot_cnt += c
        tot_per += per
        s.destroy()
        torch.cuda.empty_cache()
    assert tot_cnt == want_cnt
    want_per_marked = want_per.copy()
    # the marks of earlier slabs are bits of the list voxels: eBVM_COLLISION (bit 2) now counts for marked entries of LATER calls
    assert np.array_equal(np.delete(tot_per, 2), np.delete(want_per_marked, 2))
    _, _, vx = lst.download()
    marked = (vx[:, 0] & 4) != 0
    assert np.array_equal(marked, hit)
    vx[:, 0] &= 0xFB
    assert np.array_equal(vx, voxels)
    lst.destroy()


def test_c3_real_size_batch_on_a_16_configuration_sample(gv, oracle):
    """BASELINE.json configs[2] at its real size: 256^3 environment (40 boxes), 7 links x 50 000 points with meanings
    4..10, mask {6..10}, 1024 planner configurations in one call; the oracle's literal planner sequence
    (clear, transform, insert, collide per configuration) on 16 of them."""
    from benchmarks.run_configs import arm_links, arm_transforms
    dims, side = (256, 256, 256), 0.01
    rng = np.random.Generator(np.random.PCG64(303))
    env, r_env = gv.BitVectorVoxelMap(dims, side), oracle.new_map(BIT, dims)
    g = np.arange(24)
    local = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    pts = np.concatenate([((local + rng.integers(0, 256 - 24, 3)) + 0.5) * side for _ in range(40)]).astype(np.float32)
    env.insertPointCloud(pts, 1)
    oracle.insert_points(r_env, BIT, dims, side, pts, 1)
    links, lengths = arm_links()
    T = arm_transforms(lengths, 1024)
    meanings = list(range(4, 11))
    mask = gv.BitVectorVoxel()
    for b in (6, 7, 8, 9, 10):
        mask.setBit(b)
    got = gv.batch_robot_collide(env, gv.MetaPointCloud(links), meanings, T, mask, 1.0)
    got_nomask = gv.batch_robot_collide(env, gv.MetaPointCloud(links), meanings, T, None, 1.0)
    assert len(got) == 1024 and (got > 0).sum() > 50 and (got_nomask >= got).all()
    sample = np.r_[0:4, 509:515, 1018:1024]
    base = np.concatenate(links)
    want = oracle.batch_robot_collide(r_env, BIT, dims, side, base, [len(l) for l in links], meanings, T[sample], mask.bytes, 1.0)
    assert np.array_equal(got[sample], want)
    want = oracle.batch_robot_collide(r_env, BIT, dims, side, base, [len(l) for l in links], meanings, T[sample], None, 1.0)
    assert np.array_equal(got_nomask[sample], want)
    assert want.max() > 0
