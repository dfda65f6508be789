"""GPU parity: the sm_100a path (through the C ABI, via gpu_voxels_b200.api) against the CPU oracle on the same
seeded inputs. Bit-exact: voxel bytes, int8 log-odds and counts are compared with ==, no tolerance."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BIT, PROB = 0, 4
DIMS = (89, 123, 74)  # the reference's default test fixture (GV/test/testing_fixtures.hpp:131-166)
N = DIMS[0] * DIMS[1] * DIMS[2]


@pytest.fixture(scope="module")
def gv():
    import gpu_voxels_b200 as g
    g._lib.init()
    return g


def _mk(gv, t, dims=DIMS, side=1.0, slab=None):
    return (gv.ProbVoxelMap if t == PROB else gv.BitVectorVoxelMap)(dims, side, slab=slab)


def _flat(m):
    return m.download().reshape(-1).view(np.uint8)


def _cloud(seed, n, lo, hi, specials=True):
    rng = np.random.Generator(np.random.PCG64(seed))
    p = rng.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    if specials and n >= 64:
        p[5] = [np.nan, 0.5, 0.5]
        p[6] = [0.5, np.inf, 0.5]
        p[7] = [-np.inf, 0.5, 0.5]
        p[8] = [-0.3, 0.4, 0.4]        # negative -> saturates to column 0 on the device, NOT rejected
        p[9] = [1e20, 0.1, 0.1]        # >= 2^32 voxels -> 0xFFFFFFFF -> rejected
        p[10] = [-0.0, 0.0, 0.0]
        p[11] = [0.29999998, 0.3, 0.30000004]  # neighbours of a voxel boundary at side 0.01
    return p


# ---- the reference's own known-answer tests, run on the GPU path ------------------------------------------
def test_ref_collision(gv, oracle):
    """GV/test/testing_voxelmap.cu:95-114"""
    pts = oracle.create_equidistant_points(10000, DIMS, 1.0)
    m1, m2 = _mk(gv, PROB), _mk(gv, PROB)
    m1.insertPointCloud(pts, gv.eBVM_OCCUPIED)
    m2.insertPointCloud(pts, gv.eBVM_OCCUPIED)
    assert m1.collideWith(m2, 0.1) == 10000


def test_ref_collision_with_offset(gv, oracle):
    """GV/test/testing_voxelmap.cu:118-141"""
    m1, m2 = _mk(gv, PROB), _mk(gv, PROB)
    m1.insertPointCloud(oracle.create_box_of_points((2.1, 2.1, 2.1), (4.1, 4.1, 4.1), 0.5), gv.eBVM_OCCUPIED)
    m2.insertPointCloud(oracle.create_box_of_points((3.1, 3.1, 3.1), (5.1, 5.1, 5.1), 0.5), gv.eBVM_OCCUPIED)
    assert m1.collideWith(m2, 0.1) == 8
    assert m1.collideWith(m2, 0.1, (-1, 0, -1)) == 18
    assert m2.collideWith(m1, 0.1, (1, 0, 1)) == 18


def test_ref_no_collision(gv, oracle):
    """GV/test/testing_voxelmap.cu:143-167"""
    black, white = oracle.create_equidistant_points(10000, DIMS, 1.0, checkerboard=True)
    m1, m2 = _mk(gv, PROB), _mk(gv, PROB)
    m1.insertPointCloud(black, gv.eBVM_OCCUPIED)
    m2.insertPointCloud(white, gv.eBVM_OCCUPIED)
    assert m1.collideWith(m2, 0.1) == 0
    assert not m1.collisionCheck(m2, 0.1)


def test_ref_collision_with_types_probab(gv):
    """GV/test/testing_voxelmap.cu:312-360"""
    pts = np.array([[x, y, z] for x in (2, 3) for y in (2, 3) for z in (2, 3)], np.float32)
    m1, m2 = _mk(gv, BIT), _mk(gv, PROB)
    for mg in (14, 46, 103):
        m1.insertPointCloud(pts, mg)
    m2.insertPointCloud(pts, gv.eBVM_OCCUPIED)
    types = gv.BitVectorVoxel()
    assert m1.collideWithTypes(m2, types) == 8
    assert types.bits() == [14, 46, 103]
    types2 = gv.BitVectorVoxel()
    assert m1.collideWithTypes(m2, types2) == 8 and types2.bits() == [2, 14, 46, 103]  # write-back of eBVM_COLLISION


# ---- voxelisation ------------------------------------------------------------------------------------------
@pytest.mark.parametrize("t", [BIT, PROB])
@pytest.mark.parametrize("n", [0, 1, 31, 33, 100000])
def test_insert_points_c1(gv, oracle, t, n):
    """config 1 shape: 128^3 @ 1 cm, cloud in [-0.02, 1.30)^3 (part of it outside), plus NaN/inf/negatives"""
    dims, side = (128, 128, 128), 0.01
    pts = _cloud(101, n, -0.02, 1.30)
    m = _mk(gv, t, dims, side)
    ref = oracle.new_map(t, dims)
    out_gpu = m.insertPointCloud(pts, gv.eBVM_OCCUPIED)
    out_ref = oracle.insert_points(ref, t, dims, side, pts, 1)
    assert out_gpu == out_ref
    assert np.array_equal(_flat(m), ref.view(np.uint8))
    if t == BIT and n:
        m.insertPointCloud(pts[: n // 2], 77)
        oracle.insert_points(ref, t, dims, side, pts[: n // 2], 77)
        assert np.array_equal(_flat(m), ref.view(np.uint8))


def test_insert_prob_rejects_other_meanings(gv):
    """GV/voxelmap/ProbVoxelMap.hpp:118-140: logged no-op"""
    m = _mk(gv, PROB, (16, 16, 16), 1.0)
    m.insertPointCloud(np.ones((10, 3), np.float32), 5)
    assert np.all(m.download() == -128)


def test_insert_meta_per_cloud_meanings_and_transforms(gv, oracle):
    """config 3 shape (scaled): 7 link clouds, meanings 4..10 (bits of bytes 0 and 1 -> the reference's racy case),
    per-link Matrix4f fused into the insert"""
    dims, side = (96, 96, 96), 0.01
    rng = np.random.Generator(np.random.PCG64(301))
    sizes = [5000, 1, 0, 7000, 333, 4096, 2500]
    clouds = [rng.normal(0, 0.05, size=(s, 3)).astype(np.float32) for s in sizes]
    meanings = list(range(4, 11))
    T = []
    for k in range(7):
        a = rng.uniform(-np.pi, np.pi)
        c, s = np.cos(a), np.sin(a)
        T.append(np.array([[c, -s, 0, 0.3 + 0.05 * k], [s, c, 0, 0.45], [0, 0, 1, 0.2 + 0.08 * k], [0, 0, 0, 1]], np.float32))
    T = np.stack(T)
    mpc = gv.MetaPointCloud(clouds)
    for tr in (None, T):
        m = _mk(gv, BIT, dims, side)
        ref = oracle.new_map(BIT, dims)
        o1 = m.insertMetaPointCloud(mpc, meanings, transforms=tr)
        o2 = oracle.insert_meta(ref, BIT, dims, side, mpc.accumulated(), sizes, meanings, tr)
        assert o1 == o2
        assert np.array_equal(_flat(m), ref)
    # single meaning overload + Prob map
    m = _mk(gv, PROB, dims, side)
    ref = oracle.new_map(PROB, dims)
    m.insertMetaPointCloud(mpc, gv.eBVM_OCCUPIED, transforms=T)
    oracle.insert_meta(ref, PROB, dims, side, mpc.accumulated(), sizes, [1], T)
    assert np.array_equal(m.download(), ref)


def test_transform_points(gv, oracle):
    rng = np.random.Generator(np.random.PCG64(7))
    p = rng.normal(0, 3, size=(10007, 3)).astype(np.float32)
    M = rng.normal(0, 1, size=(4, 4)).astype(np.float32)
    got = gv.transform_points(p, M)
    want = oracle.transform_points(M, p)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


def _depth_frame(seed, w=160, h=120):
    rng = np.random.Generator(np.random.PCG64(seed))
    u, v = np.meshgrid(np.arange(w, dtype=np.float32), np.arange(h, dtype=np.float32))
    z = (1.5 + 0.3 * np.sin(u / 17) + rng.uniform(-0.01, 0.01, size=u.shape)).astype(np.float32)
    f = np.float32(131.25)
    pts = np.stack([(u - w / 2) * z / f, (v - h / 2) * z / f, z], -1).reshape(-1, 3).astype(np.float32)
    pts[rng.uniform(size=len(pts)) < 0.02] = np.nan
    return pts


def test_update_logodds_frames(gv, oracle):
    """config 2 shape (scaled): depth frames with +72 log-odds per frame, saturating at 127; sensor looks along +x,
    part of the frustum falls below z = 0 (re-projection branch) and outside the map"""
    dims, side = (128, 128, 64), 0.02
    pose = np.array([[0, 0, 1, 0.1], [-1, 0, 0, 1.28], [0, -1, 0, 0.5]], np.float32)
    m = _mk(gv, PROB, dims, side)
    ref = oracle.new_map(PROB, dims)
    for f in range(4):
        pts = _depth_frame(201 + (f % 2))
        m.insertSensorData(pts, pose)
        oracle.update_logodds(ref, dims, side, pts, pose)
        assert np.array_equal(m.download(), ref), "frame %d" % f
        if f == 0:
            assert (ref == -56).any() and (ref == 16).any()  # one hit: -128+72; two pixels in one voxel: +144
    assert ref.max() == 127


def test_pinned_host_frames_pipelined_upload(gv, oracle):
    """GVX_MEM_HOST_PINNED: frames of different sizes from page-locked memory go up on the copy stream into alternating
    staging slots while earlier kernels run; no sync between the calls. Maps and counts equal the oracle's sequence."""
    import ctypes as C
    import torch
    from gpu_voxels_b200 import _lib
    L = _lib.load()
    dims, side = (128, 128, 64), 0.02
    pose = np.ascontiguousarray(np.array([[0, 0, 1, 0.1], [-1, 0, 0, 1.28], [0, -1, 0, 0.5]], np.float32))
    m, bm = _mk(gv, PROB, dims, side), _mk(gv, BIT, dims, side)
    ref, ref_b = oracle.new_map(PROB, dims), oracle.new_map(BIT, dims)
    frames = [_depth_frame(300 + f)[: 19200 - 3000 * (f % 3)] for f in range(7)]
    pinned = [torch.from_numpy(np.ascontiguousarray(f)).pin_memory() for f in frames]
    # the BitVector map receives the same surface in world coordinates (NaN pixels -> voxel 0), so the maps do collide
    M = np.vstack([pose, np.array([[0, 0, 0, 1]], np.float32)])
    world = [np.ascontiguousarray(oracle.transform_points(M, f)) for f in frames]
    pinned_w = [torch.from_numpy(w).pin_memory() for w in world]
    count_d = torch.zeros(7, dtype=torch.int64, device="cuda")
    for f, t in enumerate(pinned):
        _lib.check(L.gvx_update_logodds(m.handle, C.c_void_p(t.data_ptr()), len(t), _lib.MEM_HOST_PINNED,
                                        pose.ctypes.data_as(C.c_void_p), 72))
        # a second consumer of the same kind of upload, other map type, asynchronous (any_outside = NULL)
        _lib.check(L.gvx_insert_points(bm.handle, C.c_void_p(pinned_w[f].data_ptr()), len(t), _lib.MEM_HOST_PINNED, 9 + f, None))
        _lib.check(L.gvx_collide_map_map_async(m.handle, bm.handle, 0.1, None, 0, C.c_void_p(count_d.data_ptr() + 8 * f)))
    _lib.check(L.gvx_sync())
    want = []
    for f, pts in enumerate(frames):
        oracle.update_logodds(ref, dims, side, pts, pose)
        oracle.insert_points(ref_b, BIT, dims, side, world[f], 9 + f)
        want.append(oracle.collide_map_map(ref, PROB, ref_b, BIT, dims[0] * dims[1] * dims[2], 0.1, 0, False))
    assert np.array_equal(m.download(), ref)
    assert np.array_equal(_flat(bm), ref_b.reshape(-1).view(np.uint8))
    assert count_d.cpu().tolist() == want and max(want) > 0
    # a mem_kind nobody defined is refused
    assert L.gvx_insert_points(bm.handle, C.c_void_p(pinned[0].data_ptr()), 10, 7, 9, None) == _lib.ERR_INVALID


def test_step_loop_programmatic_dependent_launch(gv, oracle):
    """GVX_OPT_PDL: frame update -> collide -> frame update ... queued back to back from device-resident frames, the
    launch of each kernel overlapping the tail of the one before it. Every step's count and the final map equal the
    oracle's sequence, with the option on and off (the collide of step k must see all of update k, update k+1 must
    not overtake collide k, which marks nothing here but reads the bytes update k+1 raises)."""
    import ctypes as C
    import torch
    from gpu_voxels_b200 import _lib
    L = _lib.load()
    dims, side, steps = (128, 128, 64), 0.02, 36
    pose = np.ascontiguousarray(np.array([[0, 0, 1, 0.1], [-1, 0, 0, 1.28], [0, -1, 0, 0.5]], np.float32))
    M = np.vstack([pose, np.array([[0, 0, 0, 1]], np.float32)])
    frames = [_depth_frame(500 + f) for f in range(3)]
    frames_d = [torch.from_numpy(np.ascontiguousarray(f)).cuda() for f in frames]
    ref_b = oracle.new_map(BIT, dims)
    for f in frames:
        oracle.insert_points(ref_b, BIT, dims, side, np.ascontiguousarray(oracle.transform_points(M, f)), 11)
    ref, want = oracle.new_map(PROB, dims), []
    for k in range(steps):
        oracle.update_logodds(ref, dims, side, frames[k % 3], pose, 23)
        want.append(oracle.collide_map_map(ref, PROB, ref_b, BIT, dims[0] * dims[1] * dims[2], 0.7, 0, False))
    assert len(set(want)) > 3  # voxels cross the threshold step after step
    try:
        for pdl in (1, 0, 1):
            _lib.check(L.gvx_set_option(gv.GVX_OPT_PDL, pdl))
            m, bm = _mk(gv, PROB, dims, side), _mk(gv, BIT, dims, side)
            for f in frames:
                bm.insertPointCloud(np.ascontiguousarray(oracle.transform_points(M, f)), 11)
            count_d = torch.zeros(steps, dtype=torch.int64, device="cuda")
            _lib.check(L.gvx_sync())
            for k in range(steps):
                t = frames_d[k % 3]
                _lib.check(L.gvx_update_logodds(m.handle, C.c_void_p(t.data_ptr()), len(t), _lib.MEM_DEVICE,
                                                pose.ctypes.data_as(C.c_void_p), 23))
                _lib.check(L.gvx_collide_map_map_async(m.handle, bm.handle, 0.7, None, 0, C.c_void_p(count_d.data_ptr() + 8 * k)))
            _lib.check(L.gvx_sync())
            assert count_d.cpu().tolist() == want, pdl
            assert np.array_equal(m.download(), ref), pdl
    finally:
        _lib.check(L.gvx_set_option(gv.GVX_OPT_PDL, 1))


# ---- collisions --------------------------------------------------------------------------------------------
@pytest.mark.parametrize("n_real", [5, 15, 16, 70])  # <= 16 frames travel in the kernel parameters, more in a descriptor
def test_update_logodds_multi_frame_equals_sequence(gv, oracle, n_real):
    """gvx_update_logodds_multi: several frames, each with its own pose, in one launch == the frames one by one"""
    dims, side = (128, 96, 80), 0.02
    frames = [_depth_frame(40 + k, 96, 64) for k in range(n_real)] + [np.zeros((0, 3), np.float32)]
    frames[1], frames[2] = frames[1][:-3], frames[2][:-1001]  # frames 2.. start off the 4-point grid: unstaged path
    poses = [np.array([[0, 0, 1, 0.05 * (k % 9)], [-1, 0, 0, 0.5], [0, -1, 0, 0.4 + 0.02 * (k % 7)]], np.float32) for k in range(n_real)]
    poses.append(np.array([[1, 0, 0, 9.0], [0, 1, 0, 0], [0, 0, 1, 0]], np.float32))  # empty frame, sensor outside
    poses[3][0, 3] = -0.5  # sensor origin outside the map: that frame must contribute nothing
    for slab in (None, (17, 53)):
        m = gv.ProbVoxelMap(dims, side, slab=slab)
        z0, z1 = slab if slab else (0, dims[2])
        r = oracle.new_map(PROB, dims, z0, z1)
        for rep in range(2):
            m.insertSensorDataMulti(frames, poses)
            for f, q in zip(frames, poses):
                oracle.update_logodds(r, dims, side, f, q, 72, z0, z1)
        assert np.array_equal(m.download(), r)
        assert (r != -128).sum() > 1000


def _filled_pair(gv, oracle, ta, tb, dims, side, seed):
    hi = dims[0] * side
    pa = _cloud(seed, 60000, -0.05 * hi, 1.05 * hi)
    pb = np.concatenate([pa[:20000] + np.float32(0.2 * side), _cloud(seed + 1, 40000, 0, hi)])
    a, b = _mk(gv, ta, dims, side), _mk(gv, tb, dims, side)
    ra, rb = oracle.new_map(ta, dims), oracle.new_map(tb, dims)
    for (m, r, t, p) in ((a, ra, ta, pa), (b, rb, tb, pb)):
        if t == PROB:
            pose = np.array([[1, 0, 0, 0], [0, 1, 0, 0], [0, 0, 1, 0]], np.float32)
            m.insertSensorData(p, pose); oracle.update_logodds(r, dims, side, p, pose)
            m.insertSensorData(p[:30000], pose); oracle.update_logodds(r, dims, side, p[:30000], pose)
        else:
            m.insertPointCloud(p, 1); oracle.insert_points(r, t, dims, side, p, 1)
            m.insertPointCloud(p[:1000], 200); oracle.insert_points(r, t, dims, side, p[:1000], 200)
            m.insertPointCloud(p[1000:1500], 0); oracle.insert_points(r, t, dims, side, p[1000:1500], 0)  # eBVM_FREE only
    return a, b, ra, rb


@pytest.mark.parametrize("ta,tb", [(BIT, PROB), (PROB, BIT), (PROB, PROB), (BIT, BIT)])
def test_collide_all_pairs(gv, oracle, ta, tb):
    dims, side = (67, 53, 41), 0.02  # odd sizes: unaligned tails
    n = dims[0] * dims[1] * dims[2]
    a, b, ra, rb = _filled_pair(gv, oracle, ta, tb, dims, side, 11)
    for thr in (1.0, 0.5, 0.4, 0.0):
        assert a.collideWith(b, thr, mark=False) == oracle.collide_map_map(ra, ta, rb, tb, n, thr, 0, False), thr
    for off in [(1, 0, 0), (-1, 0, -1), (3, -2, 1), (0, 0, 16), (-16, 0, 0), (16, 0, 0), (0, 0, -40), (0, 0, 41), (0, 0, -500)]:
        lin = oracle.voxel_index_signed(dims, off)
        assert a.collideWith(b, 0.4, off, mark=False) == oracle.collide_map_map(ra, ta, rb, tb, n, 0.4, lin, False), off
    # the side effect: colliding voxels of `this` carry eBVM_COLLISION / occupancy 127 afterwards
    assert a.collideWith(b, 0.4, (2, 1, 0)) == oracle.collide_map_map(ra, ta, rb, tb, n, 0.4, oracle.voxel_index_signed(dims, (2, 1, 0)), True)
    assert np.array_equal(_flat(a), ra.view(np.uint8))
    assert a.collideWith(b, 0.4) == oracle.collide_map_map(ra, ta, rb, tb, n, 0.4, 0, True)
    assert np.array_equal(_flat(a), ra.view(np.uint8))
    assert np.array_equal(_flat(b), rb.view(np.uint8))  # `other` is never written


def test_collide_types_random(gv, oracle):
    dims, side = (64, 48, 40), 0.02
    n = dims[0] * dims[1] * dims[2]
    a, b, ra, rb = _filled_pair(gv, oracle, BIT, PROB, dims, side, 23)
    types = gv.BitVectorVoxel(); types.setBit(250)
    cnt = a.collideWithTypes(b, types, 0.4)
    mask = np.zeros(32, np.uint8); mask[250 >> 3] |= 1 << (250 & 7)
    want, mask = oracle.collide_bit_prob_types(ra, rb, n, 0.4, True, mask)
    assert cnt == want and np.array_equal(types.bytes, mask)
    assert np.array_equal(_flat(a), ra)


def test_clear_meaning_and_bits(gv, oracle):
    dims, side = (40, 40, 40), 0.05
    a, b, ra, rb = _filled_pair(gv, oracle, BIT, BIT, dims, side, 5)
    n = 40 ** 3
    a.clearBitVoxelMeaning(200)
    mask = np.zeros(32, np.uint8); mask[200 >> 3] = 1 << (200 & 7)
    oracle.clear_bits(ra, n, mask)
    assert np.array_equal(_flat(a), ra)
    bits = gv.BitVectorVoxel(); bits.setBit(0); bits.setBit(1)
    a.clearBits(bits)
    oracle.clear_bits(ra, n, bits.bytes)
    assert np.array_equal(_flat(a), ra) and not ra.any()
    a.insertPointCloud(np.full((3, 3), 0.5, np.float32), 9)
    a.clearMap()
    assert not _flat(a).any()
    p = _mk(gv, PROB, dims, side)
    p.insertPointCloud(np.full((3, 3), 0.5, np.float32), 1)
    p.clearBitVoxelMeaning(5)           # logged no-op
    assert (p.download() == 127).sum() == 1
    p.clearBitVoxelMeaning(gv.eBVM_OCCUPIED)
    assert np.all(p.download() == -128)


# ---- z-slab shards are byte-identical slices of the single map ---------------------------------------------
@pytest.mark.parametrize("t", [BIT, PROB])
def test_slab_shards_concatenate_to_full_map(gv, oracle, t):
    dims, side = (48, 40, 37), 0.025
    pts = _cloud(401, 50000, -0.05, 1.3)
    full = _mk(gv, t, dims, side)
    out_full = full.insertPointCloud(pts, 1)
    other = _mk(gv, PROB, dims, side)
    other.insertPointCloud(_cloud(402, 50000, 0, 1.2, False), 1)
    total = full.collideWith(other, 1.0, mark=False)
    cuts = [0, 9, 10, 25, 37]
    parts, cnt = [], 0
    for z0, z1 in zip(cuts[:-1], cuts[1:]):
        s = _mk(gv, t, dims, side, slab=(z0, z1))
        assert s.insertPointCloud(pts, 1) == out_full  # "outside" means outside the WHOLE grid
        so = _mk(gv, PROB, dims, side, slab=(z0, z1))
        so.insertPointCloud(_cloud(402, 50000, 0, 1.2, False), 1)
        cnt += s.collideWith(so, 1.0, mark=False)
        parts.append(_flat(s))
        ref = oracle.new_map(t, dims, z0, z1)
        oracle.insert_points(ref, t, dims, side, pts, 1, z0, z1)
        assert np.array_equal(parts[-1], ref.view(np.uint8))
    assert np.array_equal(np.concatenate(parts), _flat(full))
    assert cnt == total


# ---- full-size properties (BASELINE.json config 2: 512^3) ---------------------------------------------------
def test_c2_full_size_properties(gv, oracle):
    """512^3 Prob x BitVector (4 GiB, refused by the reference). The oracle is too slow at this size, so check
    size-independent properties: count == |voxels(cloud A) & voxels(cloud B)| computed from oracle voxel indices,
    symmetry, idempotence of insertion, and count after clear == 0."""
    dims, side = (512, 512, 512), 0.01
    pa = _cloud(201, 307200, 0.0, 5.12, False)
    pb = np.concatenate([pa[:100000], _cloud(202, 200000, 0.0, 5.12, False)])
    env, rob = _mk(gv, PROB, dims, side), _mk(gv, BIT, dims, side)
    env.insertPointCloud(pa, 1)
    rob.insertPointCloud(pb, 1)
    rob.insertPointCloud(pb, 1)  # idempotent
    def vox(p):
        c = (np.floor(p / np.float32(side))).astype(np.int64)  # all coordinates positive and finite here
        return np.unique(c[:, 2] * 512 * 512 + c[:, 1] * 512 + c[:, 0])
    spot = oracle.map_to_voxels(side, pa[:2000]).astype(np.int64)
    assert np.array_equal(spot, np.floor(pa[:2000] / np.float32(side)).astype(np.int64))
    want = len(np.intersect1d(vox(pa), vox(pb)))
    assert want >= 100000 * 0.99
    assert env.collideWith(rob, 1.0, mark=False) == want
    assert rob.collideWith(env, 1.0, mark=False) == want
    assert rob.collideWith(env, 1.0) == want            # marks bit 2
    assert rob.collideWith(env, 1.0) == want            # still the same count
    rob.clearBitVoxelMeaning(gv.eBVM_OCCUPIED)
    assert rob.collideWith(env, 1.0, mark=False) == want  # only eBVM_COLLISION is left on the colliding voxels
    rob.clearMap()
    assert rob.collideWith(env, 1.0) == 0


def test_dispatch_entry_point_on_one_rank(gv, oracle):
    """gvx_update_logodds_dispatch without peers: unconnected it is the plain update; connected with world 1 the
    dispatch kernel runs its local phase only. (The exchange itself needs one GPU per rank: tests/multi_gpu_check.py.)"""
    import ctypes as C
    from gpu_voxels_b200 import _lib
    L = _lib.init()
    dims, side = (64, 48, 40), 0.02
    rng = np.random.Generator(np.random.PCG64(77))
    pts = rng.uniform(-0.7, 0.7, size=(50000, 3)).astype(np.float32)
    pts[:20000] *= np.float32(0.03)
    pts[100:130] = np.nan
    pose = np.eye(4, dtype=np.float32)[:3].copy()
    pose[:, 3] = (0.6, 0.5, 0.4)
    want = oracle.new_map(PROB, dims)
    for _ in range(3):
        oracle.update_logodds(want, dims, side, pts, pose)
    zb = np.asarray([0, dims[2]], np.uint32)
    a = gv.ProbVoxelMap(dims, side)
    for _ in range(3):
        a.insertSensorDataDispatch(pts, pose, zb)
    assert np.array_equal(a.download(), want)
    handle = (C.c_ubyte * 64)()
    _lib.check(L.gvx_peer_dispatch_create(1, 50000, handle))
    _lib.check(L.gvx_peer_dispatch_connect(0, 1, None))
    try:
        b = gv.ProbVoxelMap(dims, side)
        for _ in range(3):
            b.insertSensorDataDispatch(pts, pose, zb)
        assert np.array_equal(b.download(), want)
        assert b.collideWith(a, 0.9, mark=False) == oracle.collide_map_map(want.copy(), PROB, want, PROB, want.size, 0.9, 0, False)
        with pytest.raises(gv.GvxError):   # more points than the inboxes hold
            b.insertSensorDataDispatch(np.zeros((50001, 3), np.float32), pose, zb)
        with pytest.raises(gv.GvxError):   # bounds that do not describe the map
            b.insertSensorDataDispatch(pts, pose, np.asarray([0, dims[2] - 1], np.uint32))
    finally:
        _lib.check(L.gvx_peer_dispatch_disconnect())
