"""The occupancy summary of BitVectorVoxelMaps (1 bit per voxel, maintained by every mutating call) and the collide
kernels that read it instead of the 32-byte voxels: the summary must equal the occupancy recomputed from the voxel
bytes after any sequence of operations, and every collision result must be identical with the summary path on and off
(GVX_OPT_USE_SUMMARY) and identical to the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
BIT, PROB = 0, 4


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    yield g
    g.api.set_option(g.api.GVX_OPT_USE_SUMMARY, 1)


def _occ_from_bytes(vox):
    v = vox.reshape(-1, 32).copy()
    v[:, 0] &= 0xFE
    return v.any(axis=1)


def _check(m):
    bits, valid = m.summary()
    assert valid
    assert np.array_equal(bits, _occ_from_bytes(m.download())), "summary differs from the voxels' occupancy"


def _cloud(seed, n, hi):
    rng = np.random.Generator(np.random.PCG64(seed))
    return rng.uniform(-0.02, hi, size=(n, 3)).astype(np.float32)


@pytest.mark.parametrize("slab", [None, (7, 30)])
def test_summary_tracks_every_mutation(gv, oracle, slab):
    dims, side = (50, 37, 41), 0.01  # 75850 voxels: not a multiple of 32
    m = gv.BitVectorVoxelMap(dims, side, slab=slab)
    _check(m)
    m.insertPointCloud(_cloud(1, 20000, 0.5), 0)       # eBVM_FREE only: bits set, voxels NOT occupied
    _check(m)
    assert not m.summary()[0].any()
    m.insertPointCloud(_cloud(2, 20000, 0.5), 17)
    _check(m)
    T = np.tile(np.eye(4, dtype=np.float32), (3, 1, 1)); T[1, :3, 3] = [0.05, -0.02, 0.1]
    m.insertMetaPointCloud(gv.MetaPointCloud([_cloud(3, 5000, 0.4), _cloud(4, 7000, 0.4), _cloud(5, 1, 0.4)]), [0, 200, 3],
                           transforms=T)
    _check(m)
    m.clearBit(17)
    _check(m)
    mask = gv.BitVectorVoxel(); mask.setBit(200); mask.setBit(3); mask.setBit(99)
    m.clearBits(mask)
    _check(m)
    assert not m.summary()[0].any()  # only eBVM_FREE bits are left
    m.insertPointCloud(_cloud(6, 30000, 0.5), 1)
    z0, z1 = slab if slab else (0, dims[2])
    p = gv.ProbVoxelMap(dims, side, slab=slab)
    p.insertPointCloud(_cloud(6, 30000, 0.5)[:9000], 1)
    assert m.collideWith(p, 1.0) > 0  # marks eBVM_COLLISION on m: occupancy unchanged
    _check(m)
    m.clearBit(1)                     # collided voxels stay occupied through eBVM_COLLISION
    _check(m)
    assert m.summary()[0].any()
    up = np.zeros((m.numVoxels(), 32), np.uint8)
    up[::7, 5] = 1; up[::11, 0] = 1   # bit 40 / bit 0
    m.upload(up)
    _check(m)
    m.clearMap()
    _check(m)
    m.getVoidDeviceDataPtr()          # raw pointer handed out: summary no longer trusted ...
    assert not m.summary()[1]
    m.insertPointCloud(_cloud(7, 1000, 0.5), 9)
    assert m.collideWith(p, 1.0) == m.collideWith(p, 1.0)  # ... collide falls back to the voxels
    m.refreshSummary()
    _check(m)


def test_prob_known_summary_tracks_every_mutation(gv, oracle):
    """ProbVoxelMap: summary bit = "known" (occupancy != UNKNOWN); a clear bit must always mean UNKNOWN"""
    dims, side = (50, 37, 41), 0.01
    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)
    for slab in (None, (5, 33)):
        m = gv.ProbVoxelMap(dims, side, slab=slab)

        def check():
            bits, valid = m.summary()
            assert valid and np.array_equal(bits, m.download() != -128)
        check()
        m.insertPointCloud(_cloud(1, 20000, 0.5), 1)
        check()
        m.insertSensorData(_cloud(2, 30000, 0.5), pose)
        m.insertSensorDataMulti([_cloud(3, 5000, 0.5), _cloud(4, 5000, 0.5)], [pose, pose], delta_occupied=-10)
        check()
        b = gv.BitVectorVoxelMap(dims, side, slab=slab)
        b.insertPointCloud(_cloud(2, 30000, 0.5)[:8000], 5)
        assert m.collideWith(b, 0.3) > 0   # marks 127
        check()
        up = np.full(m.numVoxels(), -128, np.int8); up[::5] = -127; up[::9] = 40
        m.upload(up)
        check()
        m.clearMap()
        check()
        assert not m.summary()[0].any()
        m.getVoidDeviceDataPtr()
        assert not m.summary()[1]
        m.refreshSummary()
        check()


@pytest.mark.parametrize("ta,tb", [(BIT, PROB), (PROB, BIT), (BIT, BIT), (PROB, PROB)])
def test_summary_and_streaming_paths_agree_with_the_oracle(gv, oracle, ta, tb):
    dims, side = (70, 45, 53), 0.01  # 166950 voxels, tail group of 6
    n = dims[0] * dims[1] * dims[2]
    pa, pb = _cloud(11, 60000, 0.72), np.concatenate([_cloud(11, 60000, 0.72)[:15000], _cloud(12, 45000, 0.72)])
    pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0.001]], np.float32)
    results = {}
    for use in (1, 0):
        gv.api.set_option(gv.api.GVX_OPT_USE_SUMMARY, use)
        a = (gv.ProbVoxelMap if ta == PROB else gv.BitVectorVoxelMap)(dims, side)
        b = (gv.ProbVoxelMap if tb == PROB else gv.BitVectorVoxelMap)(dims, side)
        ra, rb = oracle.new_map(ta, dims), oracle.new_map(tb, dims)
        for mp_, r, t, p in ((a, ra, ta, pa), (b, rb, tb, pb)):
            if t == PROB:
                for rep in range(2):
                    mp_.insertSensorData(p[: len(p) // (rep + 1)], pose)
                    oracle.update_logodds(r, dims, side, p[: len(p) // (rep + 1)], pose)
            else:
                mp_.insertMetaPointCloud(gv.MetaPointCloud([p[:20000], p[20000:]]), [0, 33])
                oracle.insert_meta(r, t, dims, side, p, [20000, len(p) - 20000], [0, 33])
        out = []
        for thr, mark in ((0.5, False), (0.2, True), (1.0, True), (0.5, True)):
            got = a.collideWith(b, thr, mark=mark)
            want = oracle.collide_map_map(ra, ta, rb, tb, n, thr, 0, mark)
            assert got == want
            assert np.array_equal(a.download().reshape(-1).view(np.uint8), ra.view(np.uint8))
            out.append(got)
        assert a.collisionCheck(b, 0.5) == (out[0] > 0)
        if ta == BIT and tb == PROB:
            ty = gv.BitVectorVoxel(); ty.setBit(250)
            c = a.collideWithTypes(b, ty, 0.5)
            wc, wm = oracle.collide_bit_prob_types(ra, rb.view(np.int8), n, 0.5, True, _m250())
            assert c == wc and np.array_equal(ty.bytes, wm)
            out.append(c)
        results[use] = out
        assert max(out) > 0
    assert results[0] == results[1]


def _m250():
    m = np.zeros(32, np.uint8)
    m[250 >> 3] |= 1 << (250 & 7)
    return m


def test_list_and_batch_use_the_map_summary(gv, oracle):
    dims, side = (64, 64, 40), 0.01
    env_pts = _cloud(21, 80000, 0.62)
    lst_pts = np.concatenate([env_pts[:10000], _cloud(22, 50000, 0.62)])
    res = {}
    for use in (1, 0):
        gv.api.set_option(gv.api.GVX_OPT_USE_SUMMARY, use)
        m = gv.BitVectorVoxelMap(dims, side)
        m.insertMetaPointCloud(gv.MetaPointCloud([env_pts[:40000], env_pts[40000:]]), [0, 5])  # half of it eBVM_FREE only
        lst = gv.BitVectorVoxelList(dims, side)
        lst.insertPointCloud(lst_pts, 8)
        ty = gv.BitVectorVoxel()
        links = [lst_pts[:3000] - 0.3, lst_pts[3000:5000] - 0.3]
        T = np.tile(np.eye(4, dtype=np.float32), (5, 2, 1, 1)); T[:, :, :3, 3] = np.linspace(0.2, 0.4, 30).reshape(5, 2, 3)
        res[use] = (lst.collideWith(m, 1.0, (1, 0, -1), mark=False), lst.collideWithTypes(m, ty, 1.0, mark=False),
                    lst.collideCountingPerMeaning(m, mark=False)[0],
                    gv.batch_robot_collide(m, gv.MetaPointCloud(links), [4, 5], T).tolist())
    assert res[0] == res[1] and res[1][0] > 0 and sum(res[1][3]) > 0
    r_m = oracle.new_map(BIT, dims)
    oracle.insert_meta(r_m, BIT, dims, side, env_pts, [40000, 40000], [0, 5])
    r_l = oracle.VoxelList(dims, side)
    r_l.insert_points(lst_pts, 8)
    assert res[1][0] == r_l.collide(0, r_m, BIT, 1.0, oracle.voxel_index_signed(dims, (1, 0, -1)), False)[0]
