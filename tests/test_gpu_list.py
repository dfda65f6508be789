"""GPU parity of the BitVectorVoxelList path (build + list x map collisions) against the CPU oracle, through the C ABI.
Bit-exact: ids, coords, 256-bit voxels, counts, masks and per-meaning counts compared with ==."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4
DIMS = (89, 123, 74)  # GV/test/testing_fixtures.hpp:131-166
N = DIMS[0] * DIMS[1] * DIMS[2]


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _two_boxes(oracle):
    b1 = oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1)
    b2 = oracle.create_box_of_points((2.1, 2.1, 2.1), (4.9, 4.9, 4.9), 0.1)
    return b1, b2


def _same_list(lst, ref):
    ids, co, vx = lst.download()
    assert len(ids) == len(ref)
    assert np.array_equal(ids, ref.ids)
    assert np.array_equal(co, ref.coords)
    assert np.array_equal(vx, ref.voxels)


def _mask(bits):
    m = np.zeros(32, np.uint8)
    for b in bits:
        m[b >> 3] |= 1 << (b & 7)
    return m


# ---- the reference's own known-answer tests on the GPU path ----------------------------------------------------
def test_ref_insert_metapointcloud(gv, oracle):
    """GV/test/testing_voxellist.cu:282-338"""
    b1, b2 = _two_boxes(oracle)
    lst = gv.BitVectorVoxelList(DIMS, 1.0)
    lst.insertMetaPointCloud(gv.MetaPointCloud([b1, b2]), [11, 12])
    assert lst.getDimensions()[0] == 46
    ids, co, vx = lst.download()
    assert np.all(np.diff(ids.astype(np.int64)) > 0)
    both = [i for i in range(46) if oracle.bits_of(vx[i]) == [11, 12]]
    assert len(both) == 8 and all(2 <= c <= 3 for i in both for c in co[i])


@pytest.mark.parametrize("t", [PROB, BIT])
def test_ref_collide_with_map_and_types(gv, oracle, t):
    """GV/test/testing_voxellist.cu:109-236: 46 collisions, types == {11,12}, for Prob and BitVector maps"""
    b1, b2 = _two_boxes(oracle)
    mpc = gv.MetaPointCloud([b1, b2])
    lst = gv.BitVectorVoxelList(DIMS, 1.0)
    lst.insertMetaPointCloud(mpc, [11, 12])
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(DIMS, 1.0)
    m.insertMetaPointCloud(mpc, gv.eBVM_OCCUPIED)
    assert lst.collideWith(m, 1.0) == 46
    types = gv.BitVectorVoxel()
    types.setBit(77)  # must be cleared by the call (GV/voxellist/BitVoxelList.hpp:163)
    assert lst.collideWithTypes(m, types, 1.0) == 46
    assert types.bits() == [2, 11, 12]  # eBVM_COLLISION was set on every entry by the marking collideWith above


def test_ref_collide_shifting(gv, oracle):
    """GV/test/testing_voxellist.cu:239-279"""
    b1 = oracle.create_box_of_points((1.1, 1.1, 1.1), (3.9, 3.9, 3.9), 0.1)
    lst = gv.BitVectorVoxelList(DIMS, 1.0)
    lst.insertPointCloud(b1, gv.eBVM_OCCUPIED)
    m = gv.ProbVoxelMap(DIMS, 1.0)
    m.insertPointCloud(b1, gv.eBVM_OCCUPIED)
    for sx, expect in ((0, 27), (1, 18), (2, 9), (3, 0)):
        assert lst.collideWith(m, 1.0, (sx, 0, 0)) == expect


def test_ref_bitmasked_collision(gv, oracle):
    """GV/test/testing_voxellist.cu:739-864, side 0.025 (fp32 rounding cases at voxel boundaries)"""
    side = 0.025
    lst = gv.BitVectorVoxelList(DIMS, side)
    m = gv.ProbVoxelMap(DIMS, side)
    m.insertPointCloud(oracle.create_box_of_points((0.111, 0.111, 0.111), (0.799, 0.399, 0.399), 0.005), gv.eBVM_OCCUPIED)
    for mn, mx, mg in [((0.111, 0.311, 0.111), (0.399, 0.599, 0.399), 34),
                       ((0.311, 0.311, 0.111), (0.599, 0.599, 0.399), 63),
                       ((0.511, 0.311, 0.111), (0.799, 0.599, 0.399), 102)]:
        lst.insertPointCloud(oracle.create_box_of_points(mn, mx, 0.005), mg)
    for off3, expect in [((0, 0, 0), (0, 0, 9, 15, 21)), ((0, -4, 0), (0, 0, 18, 30, 42)),
                         ((0, -4, -4), (0, 0, 12, 20, 28))]:
        mask = gv.BitVectorVoxel()
        got = [lst.collideWithTypeMask(m, mask, 1.0, off3)]
        for bit in (33, 34, 63, 102):
            mask.setBit(bit)
            got.append(lst.collideWithTypeMask(m, mask, 1.0, off3))
        assert got == [e * 64 for e in expect]


def test_dimension_mismatch_returns_ssize_max(gv):
    lst = gv.BitVectorVoxelList(DIMS, 1.0)
    lst.insertPointCloud(np.array([[1.5, 1.5, 1.5]], np.float32), 5)
    m = gv.ProbVoxelMap((64, 64, 64), 1.0)
    assert lst.collideWith(m) == gv.api.SSIZE_MAX


# ---- randomised parity --------------------------------------------------------------------------------------------
def _cloud(seed, n, lo, hi):
    rng = np.random.Generator(np.random.PCG64(seed))
    p = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    if n >= 16:
        p[3] = [np.nan, 0.5, 0.5]
        p[4] = [-0.3, 0.4, 0.4]
        p[5] = [1e20, 0.1, 0.1]
        p[6] = [0.29999998, 0.3, 0.30000004]
    return p


def test_build_incremental_with_outside_points(gv, oracle):
    """three inserts (meta, single, empty) incl. NaN / negative / huge / outside points: wrapped uint32 ids, stable
    coords of the first entry per id, OR-merged bits"""
    dims, side = (64, 48, 40), 0.01
    lst, ref = gv.BitVectorVoxelList(dims, side), oracle.VoxelList(dims, side)
    clouds = [_cloud(11, 30000, -0.05, 0.70), _cloud(12, 1, 0.1, 0.2), _cloud(13, 20000, 0.0, 0.45)]
    lst.insertMetaPointCloud(gv.MetaPointCloud(clouds), [7, 200, 33])
    ref.insert_meta(np.concatenate(clouds), [len(c) for c in clouds], [7, 200, 33])
    _same_list(lst, ref)
    more = _cloud(14, 50000, 0.0, 0.5)
    lst.insertPointCloud(more, 63)
    ref.insert_points(more, 63)
    _same_list(lst, ref)
    lst.insertPointCloud(np.zeros((0, 3), np.float32), 9)
    _same_list(lst, ref)
    lst.insertMetaPointCloud(gv.MetaPointCloud([more[:100], more[100:300]]), 254)  # one meaning for all sub-clouds
    ref.insert_meta(more[:300], [100, 200], [254])
    _same_list(lst, ref)
    lst.clearMap()
    assert len(lst) == 0
    lst.insertPointCloud(more[:10], 1)
    ref2 = oracle.VoxelList(dims, side)
    ref2.insert_points(more[:10], 1)
    _same_list(lst, ref2)


@pytest.mark.parametrize("t", [PROB, BIT])
def test_collide_modes_random(gv, oracle, t):
    dims, side = (96, 80, 72), 0.01
    n = dims[0] * dims[1] * dims[2]
    lst, ref = gv.BitVectorVoxelList(dims, side), oracle.VoxelList(dims, side)
    clouds = [_cloud(21 + k, 40000, -0.02, 0.98) for k in range(4)]
    meanings = [0, 5, 40, 41]  # meaning 0 (eBVM_FREE) entries never collide under DefaultCollider
    lst.insertMetaPointCloud(gv.MetaPointCloud(clouds), meanings)
    ref.insert_meta(np.concatenate(clouds), [len(c) for c in clouds], meanings)
    m = (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side)
    r_m = oracle.new_map(t, dims)
    env = _cloud(31, 150000, 0.0, 0.96)
    if t == PROB:
        pose = np.array([[1, 0, 0, 0.1], [0, 1, 0, 0.1], [0, 0, 1, 0.1]], np.float32)
        for k in range(2):  # log-odds 16 / 88 / saturated values so thresholds matter
            m.insertSensorData(env[: 75000 * (k + 1)], pose)
            oracle.update_logodds(r_m, dims, side, env[: 75000 * (k + 1)], pose)
    else:
        m.insertPointCloud(env, 9)
        oracle.insert_points(r_m, t, dims, side, env, 9)
    for thr in (1.0, 0.55, 0.25):
        for off3 in ((0, 0, 0), (3, -2, 1), (-5, 0, 0), (0, 0, 80)):
            off = oracle.voxel_index_signed(dims, off3)
            assert lst.collideWith(m, thr, off3, mark=False) == ref.collide(0, r_m, t, thr, off, mark=False)[0]
            ty = gv.BitVectorVoxel()
            c1 = lst.collideWithTypes(m, ty, thr, off3, mark=False)
            r1 = ref.collide(1, r_m, t, thr, off, mark=False)
            assert c1 == r1[0] and np.array_equal(ty.bytes, r1[1])
            mk = _mask([5, 41])
            assert lst.collideWithTypeMask(m, mk, thr, off3, mark=False) == ref.collide(2, r_m, t, thr, off, False, mk)[0]
            tot, per = lst.collideCountingPerMeaning(m, None, thr, off3, mark=False)
            r3 = ref.collide(3, r_m, t, thr, off, mark=False)
            assert tot == r3[0] and np.array_equal(per, r3[2])
    # marking collide: list voxels must carry eBVM_COLLISION exactly where the oracle's do
    assert lst.collideWith(m, 0.55, (1, 1, 0)) == ref.collide(0, r_m, t, 0.55, oracle.voxel_index_signed(dims, (1, 1, 0)))[0]
    _same_list(lst, ref)
    assert np.array_equal(m.download().reshape(-1).view(np.uint8), r_m.view(np.uint8))  # the map is never written


def test_collide_against_slab_shards_sums_to_full(gv, oracle):
    """z-slab sharding: counts / masks / per-meaning counts over the shards add up to the unsharded result"""
    dims, side = (64, 64, 64), 0.01
    lst = gv.BitVectorVoxelList(dims, side)
    lst.insertMetaPointCloud(gv.MetaPointCloud([_cloud(41, 60000, 0, 0.64), _cloud(42, 60000, 0, 0.64)]), [8, 90])
    env = _cloud(43, 120000, 0, 0.64)
    full = gv.BitVectorVoxelMap(dims, side)
    full.insertPointCloud(env, 1)
    want = lst.collideWith(full, 1.0, mark=False)
    want_tot, want_per = lst.collideCountingPerMeaning(full, mark=False)
    got, got_per, got_mask = 0, np.zeros(256, np.uint64), np.zeros(32, np.uint8)
    for z0, z1 in ((0, 16), (16, 17), (17, 64)):
        shard = gv.BitVectorVoxelMap(dims, side, slab=(z0, z1))
        shard.insertPointCloud(env, 1)
        got += lst.collideWith(shard, 1.0, mark=False)
        got_per += lst.collideCountingPerMeaning(shard, mark=False)[1]
        ty = gv.BitVectorVoxel()
        lst.collideWithTypes(shard, ty, mark=False)
        got_mask |= ty.bytes
    assert got == want and want > 0
    assert np.array_equal(got_per, want_per) and int(got_per.sum()) == want_tot
    ty = gv.BitVectorVoxel()
    lst.collideWithTypes(full, ty, mark=False)
    assert np.array_equal(got_mask, ty.bytes)


def test_upload_download_roundtrip_and_validation(gv):
    lst = gv.BitVectorVoxelList((32, 32, 32), 1.0)
    ids = np.array([3, 9, 1000], np.uint32)
    vx = np.zeros((3, 32), np.uint8); vx[:, 1] = [1, 2, 4]
    co = np.array([[3, 0, 0], [9, 0, 0], [8, 31, 0]], np.uint32)
    lst.upload(ids, co, vx)
    i2, c2, v2 = lst.download()
    assert np.array_equal(i2, ids) and np.array_equal(c2, co) and np.array_equal(v2, vx)
    with pytest.raises(gv.GvxError):
        lst.upload(ids[::-1].copy(), co, vx)
